"""Oracle pinning for SAO: the oracle's restatement of the four SAO strategies
(strategies-sao.h:37-60, sao-generic.c:50-182) and of kvz_sao_reconstruct (sao.c:277-337) against the compiled
reference (generic and selected AVX2 strategies); and the identity the batch face relies on: both ddistortion
functions are exact functions of the per-category / per-band statistics."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
import sao_oracle as so
from sao_oracle import _pp, ints

needs_ref = pytest.mark.skipif(not kvzlibs.have_ref(), reason="compiled reference not present")


def _blocks(rs, bw, bh, noise=6):
    base = rs.integers(0, 256, (bh, bw)).astype(np.int32)
    base = (base + np.roll(base, 1, 0) + np.roll(base, 1, 1) + np.roll(base, 2, 1)) // 4      # some structure
    orig = np.clip(base, 0, 255).astype(np.uint8)
    rec = np.clip(base + rs.integers(-noise, noise + 1, base.shape), 0, 255).astype(np.uint8)
    return orig, rec


@needs_ref
@pytest.mark.parametrize("cpuid", [0, 1])
@pytest.mark.parametrize("size", [(64, 64), (32, 32), (64, 56), (24, 32), (8, 8), (16, 4)])
def test_edge_strategies_match_reference(size, cpuid):
    O = kvzlibs.oracle()
    bw, bh = size
    rs = np.random.default_rng(bw * 100 + bh)
    orig, rec = _blocks(rs, bw, bh)
    dd = kvzlibs.ref_strategy("kvz_sao_edge_ddistortion", C.c_int,
                              [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)], cpuid=cpuid)
    ed = kvzlibs.ref_strategy("kvz_calc_sao_edge_dir", None,
                              [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)], cpuid=cpuid)
    for e in range(4):
        offs = [0] + [int(v) for v in rs.integers(-7, 8, 4)]
        assert O.kvzo_sao_edge_ddistortion_8(_pp(orig), _pp(rec), bw, bh, e, ints(offs)) == dd(_pp(orig), _pp(rec), bw, bh, e, ints(offs))
        a, b = (C.c_int * 10)(), (C.c_int * 10)()
        O.kvzo_calc_sao_edge_dir_8(_pp(orig), _pp(rec), e, bw, bh, a)
        ed(_pp(orig), _pp(rec), e, bw, bh, b)
        assert a[:] == b[:]
        # the batch face's identity: ddistortion = sum over categories of n * h^2 - 2 * h * E  (sao.c:381-388)
        want = sum(a[5 + c] * offs[c] ** 2 - 2 * offs[c] * a[c] for c in range(5))
        assert want == dd(_pp(orig), _pp(rec), bw, bh, e, ints(offs))


@needs_ref
@pytest.mark.parametrize("cpuid", [0, 1])
@pytest.mark.parametrize("size", [(64, 64), (32, 32), (40, 64), (8, 8)])
def test_band_ddistortion_matches_reference(size, cpuid):
    O = kvzlibs.oracle()
    R = kvzlibs.ref(cpuid)
    bw, bh = size
    rs = np.random.default_rng(bw + bh)
    orig, rec = _blocks(rs, bw, bh, noise=12)
    acc = (C.c_int * 64)()
    O.kvzo_calc_sao_bands_8(_pp(orig), _pp(rec), bw, bh, acc)
    s = np.array(acc[:]).reshape(2, 32)
    assert s[1].sum() == bw * bh
    assert np.array_equal(s[0], np.bincount(rec.reshape(-1) >> 3, (orig.astype(np.int64) - rec).reshape(-1), 32).astype(np.int64))
    for pos in (0, 5, int(rec.mean()) >> 3, 28):
        bands = [int(v) for v in rs.integers(-7, 8, 4)]
        got = O.kvzo_sao_band_ddistortion_8(_pp(orig), _pp(rec), bw, bh, pos, ints(bands))
        assert got == R.ref_shim_sao_band_ddistortion(_pp(orig), _pp(rec), bw, bh, pos, ints(bands))
        assert got == sum(int(s[1][pos + b]) * bands[b] ** 2 - 2 * bands[b] * int(s[0][pos + b]) for b in range(4))


@needs_ref
@pytest.mark.parametrize("cpuid", [0, 1])
def test_reconstruct_color_matches_reference(cpuid):
    O = kvzlibs.oracle()
    R = kvzlibs.ref(cpuid)
    rs = np.random.default_rng(4)
    for bw, bh in ((64, 64), (32, 32), (54, 10), (1, 7), (9, 1)):
        _, plane = _blocks(rs, bw + 2, bh + 2, noise=20)
        for p in so.random_params(rs, 12, p_none=0.0):
            for color in (0, 1, 2):
                s = so.info_of(p)
                a, b = np.zeros((bh, bw), np.uint8), np.zeros((bh, bw), np.uint8)
                src = plane[1:, 1:]
                O.kvzo_sao_reconstruct_color_8(_pp(src), _pp(a), C.byref(s), bw + 2, bw, bw, bh, color)
                R.ref_shim_sao_reconstruct_color(_pp(src), _pp(b), C.byref(s), bw + 2, bw, bw, bh, color)
                assert np.array_equal(a, b), (bw, bh, p, color)


@needs_ref
def test_reconstruct_picture_borders_match_reference():
    """kvz_sao_reconstruct on every CTU of a small 4:2:0 picture (partial CTUs at the right / bottom)"""
    O = kvzlibs.oracle()
    R = kvzlibs.ref(1)
    rs = np.random.default_rng(9)
    W, H = 152, 104
    for color, (w, h, ctu) in enumerate(((W, H, 64), (W // 2, H // 2, 32), (W // 2, H // 2, 32))):
        _, pre = _blocks(rs, w, h, noise=20)
        n = ((h + ctu - 1) // ctu) * ((w + ctu - 1) // ctu)
        params = so.random_params(rs, n)
        a, b = pre.copy(), pre.copy()
        for i, y0, x0, bh, bw in so.ctu_blocks(pre, ctu):
            s = so.info_of(params[i])
            O.kvzo_sao_reconstruct_8(_pp(pre), _pp(a), w, h, x0, y0, bw, bh, C.byref(s), color)
            R.ref_shim_sao_reconstruct(_pp(pre), _pp(b), W, H, x0, y0, bw, bh, C.byref(s), color)
        assert np.array_equal(a, b)
        assert not np.array_equal(a, pre)
