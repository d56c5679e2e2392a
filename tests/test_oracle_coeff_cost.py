"""Oracle pinning for the CABAC-exact coefficient bit cost (SURVEY.md 8f row 2, A15's kvz_get_coeff_cost):
kvzo_coeff_cabac_bits against the compiled reference's kvz_get_coeff_cost (rdo.c:158-222) with
kvz_encode_coeff_nxn (encode_coding_tree.c:103-345) behind it, random context states and coder ranges."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
import coeff_cases as cc

needs_ref = pytest.mark.skipif(not kvzlibs.have_ref(), reason="compiled reference not present")


def oracle_bits(ctx, rng, blk, width, typ, scan, signhide=1, trskip=0):
    O = kvzlibs.oracle()
    O.kvzo_coeff_cabac_bits.restype = C.c_uint32
    return int(O.kvzo_coeff_cabac_bits(C.byref(ctx), C.c_uint32(rng), blk.ctypes.data_as(C.c_void_p), width, typ, scan,
                                       signhide, trskip))


@needs_ref
@pytest.mark.parametrize("shape", cc.SHAPES, ids=[f"{w}x{w}-t{t}-s{s}" for w, t, s in cc.SHAPES])
def test_bits_match_reference(shape):
    width, typ, scan = shape
    R = kvzlibs.ref(1)
    R.ref_shim_coeff_cost.restype = C.c_uint32
    rs = np.random.default_rng(width * 100 + typ * 10 + scan)
    n = 0
    for i, blk in enumerate(cc.blocks(rs, width, 36)):
        ctx = kvzlibs.random_rdoq_ctx(rs)
        rng = int(rs.integers(256, 511))
        signhide, trskip = i % 2, (i // 2) % 2
        before = bytes(ctx)
        got = oracle_bits(ctx, rng, blk, width, typ, scan, signhide, trskip)
        want = int(R.ref_shim_coeff_cost(C.byref(ctx), C.c_uint32(rng), blk.ctypes.data_as(C.c_void_p), width, typ, scan,
                                         signhide, trskip))
        assert got == want, (i, got, want)
        assert bytes(ctx) == before
        n += got > 0
    assert n >= 30


def test_zero_block_costs_nothing():
    rs = np.random.default_rng(1)
    ctx = kvzlibs.random_rdoq_ctx(rs)
    assert oracle_bits(ctx, 400, np.zeros((8, 8), np.int16), 8, 0, 0) == 0


@needs_ref
def test_slice_init_contexts():
    """the contexts a slice starts with (kvz_init_contexts) instead of random ones"""
    R = kvzlibs.ref(1)
    R.ref_shim_coeff_cost.restype = C.c_uint32
    rs = np.random.default_rng(2)
    for qp, intra in ((22, 1), (32, 0), (45, 0)):
        ctx = kvzlibs.RdoqCtx()
        R.ref_shim_init_rdoq_ctx(qp, intra, C.byref(ctx))
        for width, typ, scan in cc.SHAPES:
            for blk in cc.blocks(rs, width, 6):
                got = oracle_bits(ctx, 510, blk, width, typ, scan)
                assert got == int(R.ref_shim_coeff_cost(C.byref(ctx), C.c_uint32(510), blk.ctypes.data_as(C.c_void_p), width, typ, scan, 1, 0))
