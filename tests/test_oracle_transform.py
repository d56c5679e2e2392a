"""Pins the oracle's DCT/DST, quant/dequant and transform-skip restatement (SURVEY.md 8a
rows A13, A14) on the compiled reference (generic and AVX2), using the gradient blocks of
the reference's dct_tests.c plus seeded random residuals."""
import ctypes as C

import numpy as np
import pytest

import kat
import kvzlibs
from kvzlibs import ptr


def _blocks(n, rs, count=6):
    out = [kat.dct_gradient_block(n, 0), kat.dct_gradient_block(n, 1)]
    for _ in range(count):
        out.append(rs.integers(-255, 256, (n, n)).astype(np.int16))
    out.append(np.full((n, n), 255, np.int16))
    out.append(np.full((n, n), -255, np.int16))
    return out


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
@pytest.mark.parametrize("n,dst", [(4, 0), (4, 1), (8, 0), (16, 0), (32, 0)])
def test_forward_inverse_vs_reference(oracle_lib, ref_lib, cpuid, n, dst):
    L = kvzlibs.ref(cpuid)
    rs = np.random.default_rng(10 + n + dst)
    for blk in _blocks(n, rs):
        src = kvzlibs.aligned_copy(blk)
        want = kvzlibs.aligned((n, n), np.int16)
        got = kvzlibs.aligned((n, n), np.int16)
        # DST is selected for 4x4 intra luma only (strategies-dct.c:59-113)
        L.ref_shim_transform2d(ptr(src), ptr(want), n, 0, dst, 0)
        oracle_lib.kvzo_fwd_transform(n, dst, 8, ptr(src), ptr(got))
        assert np.array_equal(got, want), ("fwd", n, dst)
        # inverse on the forward coefficients
        back_w = kvzlibs.aligned((n, n), np.int16)
        back_g = kvzlibs.aligned((n, n), np.int16)
        L.ref_shim_transform2d(ptr(back_w), ptr(want), n, 0, dst, 1)
        oracle_lib.kvzo_inv_transform(n, dst, 8, ptr(want), ptr(back_g))
        assert np.array_equal(back_g, back_w), ("inv", n, dst)


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
@pytest.mark.parametrize("n", [4, 8, 16, 32])
def test_quant_dequant_vs_reference(oracle_lib, ref_lib, cpuid, n):
    L = kvzlibs.ref(cpuid)
    rs = np.random.default_rng(20 + n)
    for qp in (0, 17, 22, 27, 32, 37, 51):
        for intra_slice in (0, 1):
            for signhide in (0, 1):
                for type_ in (0, 2):
                    for scan in (0, 1, 2):
                        if scan and n > 8:
                            continue
                        if type_ == 2 and n == 32:
                            continue   # 4:2:0 has no 32x32 chroma TU (no scaling list for it)
                        scale = [40, 400, 4000][int(rs.integers(0, 3))]
                        coef = kvzlibs.aligned_copy((rs.standard_normal((n, n)) * scale).clip(-32000, 32000).astype(np.int16))
                        rc = kvzlibs.RefQCfg(qp, intra_slice, signhide, 0, 0, intra_slice, 0.0)
                        oc = kvzlibs.OracleQCfg(qp, 8, intra_slice, signhide, intra_slice)
                        want = kvzlibs.aligned((n, n), np.int16)
                        got = kvzlibs.aligned((n, n), np.int16)
                        L.ref_shim_quant(C.byref(rc), ptr(coef), ptr(want), n, type_, scan)
                        oracle_lib.kvzo_quant(C.byref(oc), ptr(coef), ptr(got), n, type_, scan)
                        assert np.array_equal(got, want), ("quant", n, qp, intra_slice, signhide, type_, scan)
                        dw = kvzlibs.aligned((n, n), np.int16)
                        dg = kvzlibs.aligned((n, n), np.int16)
                        L.ref_shim_dequant(C.byref(rc), ptr(want), ptr(dw), n, type_)
                        oracle_lib.kvzo_dequant(C.byref(oc), ptr(want), ptr(dg), n, type_)
                        assert np.array_equal(dg, dw), ("dequant", n, qp, type_)


def test_scan_order_is_a_permutation(oracle_lib):
    for lg in (2, 3, 4, 5):
        for scan in (0, 1, 2):
            out = np.zeros(1 << (2 * lg), np.uint32)
            oracle_lib.kvzo_scan_order(scan, lg, ptr(out))
            assert sorted(out.tolist()) == list(range(1 << (2 * lg)))
            assert out[0] == 0


def test_scaled_qp(oracle_lib):
    # transform.c:129-143 + kvz_g_chroma_scale (tables / HEVC Table 8-10)
    assert oracle_lib.kvzo_scaled_qp(0, 32, 0) == 32
    assert oracle_lib.kvzo_scaled_qp(2, 29, 0) == 29
    assert oracle_lib.kvzo_scaled_qp(2, 30, 0) == 29
    assert oracle_lib.kvzo_scaled_qp(2, 43, 0) == 37
    assert oracle_lib.kvzo_scaled_qp(2, 51, 0) == 45


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_quantize_residual_vs_reference(oracle_lib, ref_lib, cpuid):
    # the fused residual -> T -> Q -> Q^-1 -> T^-1 -> recon path (quant-generic.c:180-273), rdoq off
    L = kvzlibs.ref(cpuid)
    rs = np.random.default_rng(77)
    for n in (4, 8, 16, 32):
        for it in range(24):
            qp = int(rs.integers(10, 46))
            color = int(rs.integers(0, 3)) if n < 32 else 0
            intra = int(rs.integers(0, 2))
            trskip = int(n == 4 and rs.random() < 0.3)
            signhide = int(rs.integers(0, 2))
            scan = int(rs.integers(0, 3)) if n <= 8 else 0
            stride = n + 8 * int(rs.integers(0, 3))
            ref_in = kvzlibs.aligned_copy(rs.integers(0, 256, (n, stride), dtype=np.uint8))
            noise = rs.integers(-int(rs.integers(1, 40)), 40, (n, stride))
            pred = kvzlibs.aligned_copy(np.clip(ref_in.astype(np.int64) + noise, 0, 255).astype(np.uint8))
            rc = kvzlibs.RefQCfg(qp, intra, signhide, 0, 0, intra, 0.0)
            oc = kvzlibs.OracleQCfg(qp, 8, intra, signhide, intra)
            rec_w = kvzlibs.aligned((n, stride), np.uint8); rec_g = kvzlibs.aligned((n, stride), np.uint8)
            co_w = kvzlibs.aligned((n * n,), np.int16); co_g = kvzlibs.aligned((n * n,), np.int16)
            hw = L.ref_shim_quantize_residual(C.byref(rc), n, color, scan, trskip, stride, stride,
                                              ptr(ref_in), ptr(pred), ptr(rec_w), ptr(co_w))
            hg = oracle_lib.kvzo_quantize_residual_8(C.byref(oc), n, color, scan, trskip, stride, stride,
                                                     ptr(ref_in), ptr(pred), ptr(rec_g), ptr(co_g))
            assert hg == hw and np.array_equal(co_g, co_w), ("coeff", n, it, qp, color, intra, trskip)
            assert np.array_equal(rec_g[:, :n], rec_w[:, :n]), ("rec", n, it)
