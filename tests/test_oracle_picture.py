"""Pins the oracle's SAD / SATD / SSD restatement (SURVEY.md 8a rows A1, A2, A6, A7, A8,
A16) on the reference's own known-answer vectors (tests/kat.py) and on the compiled,
unmodified reference (oracle/_ref), generic and AVX2 strategies both."""
import ctypes as C

import numpy as np
import pytest

import kat
import kvzlibs
from kvzlibs import ptr

PIXP = C.POINTER(C.c_uint8)


def _p8(a):
    return a.ctypes.data_as(PIXP)


# ---------------- known-answer vectors (no reference needed) ----------------

@pytest.mark.parametrize("case", kat.SAD_EDGE_CASES, ids=lambda c: f"{c[0][0]}_{c[0][1]}")
def test_image_sad_edge_kat(oracle_lib, case):
    (rx, ry), want = case
    pic, ref = kat.sad_edge_pictures()
    assert oracle_lib.kvzo_image_sad_8(ptr(pic), ptr(ref), 8, 8, 0, 0, rx, ry, 8, 8) == want


@pytest.mark.parametrize("wh", kat.REG_SAD_SHAPES, ids=lambda s: f"{s[0]}x{s[1]}")
def test_reg_sad_shapes_kat(oracle_lib, wh):
    w, h = wh
    pic, ref = kat.reg_sad_pictures()
    naive = int(np.abs(pic[:h, :w].astype(np.int64) - ref[:h, :w].astype(np.int64)).sum())
    assert oracle_lib.kvzo_reg_sad_8(ptr(pic), ptr(ref), w, h, 64, 64) == naive


def test_reg_sad_overflow_kat(oracle_lib):
    # sad_tests.c:294-316: 0 vs PIXEL_MAX over 64x64
    a = np.zeros((64, 64), np.uint8)
    b = np.full((64, 64), 255, np.uint8)
    assert oracle_lib.kvzo_reg_sad_8(ptr(a), ptr(b), 64, 64, 64, 64) == 64 * 64 * 255


@pytest.mark.parametrize("test", [0, 1, 2])
@pytest.mark.parametrize("n", [4, 8, 16, 32, 64])
def test_satd_kat(oracle_lib, test, n):
    a, b = kat.satd_buffers(test, n)
    r1 = oracle_lib.kvzo_satd_nxn_8(n, ptr(a), ptr(b))
    r2 = oracle_lib.kvzo_satd_nxn_8(n, ptr(b), ptr(a))
    assert r1 == r2 == kat.SATD_GOLDEN[test][n]


def test_coeff_abs_sum_kat(oracle_lib):
    c = kat.coeff_series()
    assert oracle_lib.kvzo_coeff_abs_sum(ptr(c), c.size) == int(np.abs(c.astype(np.int64)).sum())


# ---------------- against the compiled reference ----------------

@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_reg_sad_vs_reference(oracle_lib, ref_lib, cpuid):
    f = kvzlibs.ref_strategy("kvz_reg_sad", C.c_uint, [PIXP, PIXP, C.c_int, C.c_int, C.c_uint, C.c_uint], cpuid)
    rs = np.random.default_rng(1)
    a = rs.integers(0, 256, (80, 96), dtype=np.uint8)
    b = rs.integers(0, 256, (72, 112), dtype=np.uint8)
    for (w, h) in kat.REG_SAD_SHAPES + [(3, 5), (17, 9), (64, 1)]:
        for _ in range(3):
            ox, oy = int(rs.integers(0, 8)), int(rs.integers(0, 8))
            want = f(_p8(a[oy:, ox:]), _p8(b[oy:, ox:]), w, h, 96, 112)
            got = oracle_lib.kvzo_reg_sad_8(ptr(a[oy:, ox:]), ptr(b[oy:, ox:]), w, h, 96, 112)
            assert got == want, (w, h)


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_image_sad_satd_vs_reference(oracle_lib, ref_lib, cpuid):
    L = kvzlibs.ref(cpuid)
    W, H = 96, 80
    f = kvzlibs.synth_luma(5, W, H, 2)
    pic, ref = f[1], f[0]
    rs = np.random.default_rng(2)
    for (bw, bh) in [(8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (8, 16), (32, 16), (24, 32), (12, 16), (16, 4)]:
        for _ in range(12):
            px = int(rs.integers(0, (W - bw) // 4 + 1)) * 4
            py = int(rs.integers(0, (H - bh) // 4 + 1)) * 4
            rx = px + int(rs.integers(-90, 91))
            ry = py + int(rs.integers(-90, 91))
            want = L.ref_shim_image_calc_sad(ptr(pic), ptr(ref), W, H, px, py, rx, ry, bw, bh)
            got = oracle_lib.kvzo_image_sad_8(ptr(pic), ptr(ref), W, H, px, py, rx, ry, bw, bh)
            assert got == want, ("sad", bw, bh, px, py, rx, ry)
            want = L.ref_shim_image_calc_satd(ptr(pic), ptr(ref), W, H, px, py, rx, ry, bw, bh)
            got = oracle_lib.kvzo_image_satd_8(ptr(pic), ptr(ref), W, H, px, py, rx, ry, bw, bh)
            assert got == want, ("satd", bw, bh, px, py, rx, ry)


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_nxn_costs_vs_reference(oracle_lib, ref_lib, cpuid):
    rs = np.random.default_rng(3)
    for n in (4, 8, 16, 32, 64):
        sad = kvzlibs.ref_strategy(f"kvz_sad_{n}x{n}", C.c_uint, [PIXP, PIXP], cpuid)
        satd = kvzlibs.ref_strategy(f"kvz_satd_{n}x{n}", C.c_uint, [PIXP, PIXP], cpuid)
        for _ in range(8):
            a = kvzlibs.aligned_copy(rs.integers(0, 256, n * n, dtype=np.uint8))
            b = kvzlibs.aligned_copy(rs.integers(0, 256, n * n, dtype=np.uint8))
            assert oracle_lib.kvzo_sad_nxn_8(n, ptr(a), ptr(b)) == sad(_p8(a), _p8(b))
            assert oracle_lib.kvzo_satd_nxn_8(n, ptr(a), ptr(b)) == satd(_p8(a), _p8(b))


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_satd_any_size_and_quad_vs_reference(oracle_lib, ref_lib, cpuid):
    anys = kvzlibs.ref_strategy("kvz_satd_any_size", C.c_uint,
                                [C.c_int, C.c_int, PIXP, C.c_int, PIXP, C.c_int], cpuid)
    quad = kvzlibs.ref_strategy("kvz_satd_any_size_quad", None,
                                [C.c_int, C.c_int, C.POINTER(PIXP), C.c_int, PIXP, C.c_int, C.c_uint,
                                 C.POINTER(C.c_uint), C.POINTER(C.c_int8)], cpuid)
    rs = np.random.default_rng(4)
    for (w, h) in [(8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (8, 16), (64, 32), (12, 16), (16, 12),
                   (4, 8), (8, 4), (24, 32), (4, 4), (48, 64), (64, 16)]:
        a = rs.integers(0, 256, (64, 80), dtype=np.uint8)
        preds = [kvzlibs.aligned_copy(rs.integers(0, 256, (64, 64), dtype=np.uint8)) for _ in range(4)]
        for k in range(4):
            want = anys(w, h, _p8(a), 80, _p8(preds[k]), 64)
            assert oracle_lib.kvzo_satd_any_size_8(w, h, ptr(a), 80, ptr(preds[k]), 64) == want, (w, h)
        parr = (PIXP * 4)(*[_p8(p) for p in preds])
        want = (C.c_uint * 4)()
        valid = (C.c_int8 * 4)(1, 1, 1, 1)
        quad(w, h, parr, 64, _p8(a), 80, 4, want, valid)
        got = (C.c_uint * 4)()
        vp = (C.c_void_p * 4)(*[p.ctypes.data for p in preds])
        oracle_lib.kvzo_satd_any_size_quad_8(w, h, vp, 64, ptr(a), 80, got)
        assert list(got) == list(want), (w, h)


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_ssd_vs_reference(oracle_lib, ref_lib, cpuid):
    ssd = kvzlibs.ref_strategy("kvz_pixels_calc_ssd", C.c_uint, [PIXP, PIXP, C.c_int, C.c_int, C.c_int], cpuid)
    rs = np.random.default_rng(5)
    a = rs.integers(0, 256, (64, 72), dtype=np.uint8)
    b = rs.integers(0, 256, (64, 64), dtype=np.uint8)
    for n in (4, 8, 16, 32, 64):
        assert oracle_lib.kvzo_ssd_8(ptr(a), ptr(b), 72, 64, n) == ssd(_p8(a), _p8(b), 72, 64, n)
