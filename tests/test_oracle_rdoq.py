"""A15 (rdo.c:548-885): the oracle's RDOQ against the unmodified reference compiled in oracle/_ref."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
import rdoq_cases
from kvzlibs import ptr

needs_ref = pytest.mark.skipif(not kvzlibs.have_ref(), reason="oracle/_ref not built")


def run_oracle(case, ctx):
    O = kvzlibs.oracle()
    O.kvzo_rdoq.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    O.kvzo_rdoq.restype = None
    oc = kvzlibs.OracleQCfg(case["qp"], 8, case["slice_is_intra"], case["signhide"], case["cu_is_intra"])
    out = np.full(case["width"] ** 2, 77, np.int16)
    O.kvzo_rdoq(C.addressof(oc), case["lam"], C.addressof(ctx), ptr(case["coef"]), ptr(out), case["width"],
                case["type"], case["scan_idx"], case["tr_depth"])
    return out


def run_ref(case, ctx):
    R = kvzlibs.ref(1)
    rc = kvzlibs.RefQCfg(case["qp"], case["slice_is_intra"], case["signhide"], 1, 0, case["cu_is_intra"], case["lam"])
    out = np.full(case["width"] ** 2, 77, np.int16)
    coef = case["coef"].copy()
    R.ref_shim_rdoq_ctx(C.byref(rc), C.byref(ctx), ptr(coef), ptr(out), case["width"], case["type"],
                        case["scan_idx"], case["tr_depth"])
    return out


@needs_ref
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_rdoq_random_contexts(seed):
    rs = np.random.default_rng(100 + seed)
    n = changed = 0
    for case in rdoq_cases.cases(seed):
        ctx = kvzlibs.random_rdoq_ctx(rs)
        want = run_ref(case, ctx)
        got = run_oracle(case, ctx)
        assert np.array_equal(got, want), {k: v for k, v in case.items() if k != "coef"}
        n += 1
        changed += int(np.any(want != 0))
    assert n > 100 and changed > n // 3


@needs_ref
def test_rdoq_initial_contexts():
    """contexts as kvz_init_contexts (context.c:209-300) leaves them at the start of a slice"""
    R = kvzlibs.ref(1)
    for case in rdoq_cases.cases(9, per_shape=1):
        ctx = kvzlibs.RdoqCtx()
        R.ref_shim_init_rdoq_ctx(case["qp"], case["slice_is_intra"], C.byref(ctx))
        assert np.array_equal(run_oracle(case, ctx), run_ref(case, ctx)), {k: v for k, v in case.items() if k != "coef"}


@needs_ref
def test_rdoq_differs_from_plain_quant():
    """sanity: RDOQ is not the dead-zone quantiser (otherwise the comparison above proves little)"""
    O = kvzlibs.oracle()
    rs = np.random.default_rng(5)
    differ = 0
    for case in rdoq_cases.cases(4, per_shape=1, widths=(8, 16)):
        ctx = kvzlibs.random_rdoq_ctx(rs)
        oc = kvzlibs.OracleQCfg(case["qp"], 8, case["slice_is_intra"], case["signhide"], case["cu_is_intra"])
        plain = np.zeros(case["width"] ** 2, np.int16)
        O.kvzo_quant(C.byref(oc), ptr(case["coef"]), ptr(plain), case["width"], case["type"], case["scan_idx"])
        differ += int(not np.array_equal(plain, run_oracle(case, ctx)))
    assert differ > 5


@needs_ref
@pytest.mark.parametrize("rdoq_skip", [0, 1])
def test_quantize_residual_with_rdoq_vs_reference(rdoq_skip):
    """kvz_quantize_residual with cfg.rdoq_enable (quant-generic.c:211-221): transform -> kvz_rdoq -> reconstruction"""
    L = kvzlibs.ref(1)
    O = kvzlibs.oracle()
    O.kvzo_quantize_residual_rdoq_8.argtypes = [C.c_void_p, C.c_double, C.c_void_p] + [C.c_int] * 8 + [C.c_void_p] * 4
    rs = np.random.default_rng(70 + rdoq_skip)
    for n in (4, 8, 16, 32):
        for it in range(16):
            qp = int(rs.integers(10, 46))
            color = int(rs.integers(0, 3)) if n < 32 else 0
            intra = int(rs.integers(0, 2))
            trskip = int(n == 4 and rs.random() < 0.3)
            signhide = int(rs.integers(0, 2))
            scan = int(rs.integers(0, 3)) if n <= 8 else 0
            lam = float(0.57 * 2.0 ** ((qp - 12) / 3.0))
            stride = n + 8 * int(rs.integers(0, 3))
            ref_in = kvzlibs.aligned_copy(rs.integers(0, 256, (n, stride), dtype=np.uint8))
            noise = rs.integers(-int(rs.integers(1, 40)), 40, (n, stride))
            pred = kvzlibs.aligned_copy(np.clip(ref_in.astype(np.int64) + noise, 0, 255).astype(np.uint8))
            rc = kvzlibs.RefQCfg(qp, intra, signhide, 1, rdoq_skip, intra, lam)
            oc = kvzlibs.OracleQCfg(qp, 8, intra, signhide, intra)
            ctx = kvzlibs.RdoqCtx()
            L.ref_shim_init_rdoq_ctx(qp, intra, C.byref(ctx))
            rec_w = kvzlibs.aligned((n, stride), np.uint8); rec_g = kvzlibs.aligned((n, stride), np.uint8)
            co_w = kvzlibs.aligned((n * n,), np.int16); co_g = kvzlibs.aligned((n * n,), np.int16)
            hw = L.ref_shim_quantize_residual(C.byref(rc), n, color, scan, trskip, stride, stride,
                                              ptr(ref_in), ptr(pred), ptr(rec_w), ptr(co_w))
            hg = O.kvzo_quantize_residual_rdoq_8(C.addressof(oc), lam, C.addressof(ctx), rdoq_skip, 0, n, color, scan,
                                                 trskip, stride, stride, ptr(ref_in), ptr(pred), ptr(rec_g), ptr(co_g))
            assert hg == hw and np.array_equal(co_g, co_w), ("coeff", n, it, qp, color, intra, trskip)
            assert np.array_equal(rec_g[:, :n], rec_w[:, :n]), ("rec", n, it)
