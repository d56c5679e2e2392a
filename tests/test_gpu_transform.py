"""GPU parity: DCT/IDCT/DST, quant (with sign-bit hiding), dequant, coeff_abs_sum and the fused
kvz_quantize_residual against the oracle (pinned on the reference's generic and AVX2 strategies),
through the C ABI.  Bit-exact."""
import ctypes as C

import numpy as np
import pytest

import kat
import kvzlibs
from kvzlibs import ptr

pytestmark = pytest.mark.gpu


def _blocks(n, rs, count):
    out = [kat.dct_gradient_block(n, 0), kat.dct_gradient_block(n, 1), np.full((n, n), 255, np.int16),
           np.full((n, n), -255, np.int16), np.zeros((n, n), np.int16)]
    for _ in range(count):
        out.append(rs.integers(-255, 256, (n, n)).astype(np.int16))
    return np.stack(out)


@pytest.mark.parametrize("n,dst", [(4, 0), (4, 1), (8, 0), (16, 0), (32, 0)])
def test_forward_inverse(n, dst):
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(n + dst)
    src = _blocks(n, rs, 40)
    fwd = blocks.transform_blocks(torch.from_numpy(src).cuda(), n, inverse=False, dst4=bool(dst))
    inv = blocks.transform_blocks(fwd, n, inverse=True, dst4=bool(dst))
    # inverse of arbitrary (saturating) coefficients as well
    wild = rs.integers(-32768, 32768, src.shape).astype(np.int16)
    inv_wild = blocks.transform_blocks(torch.from_numpy(wild).cuda(), n, inverse=True, dst4=bool(dst)).cpu().numpy()
    fwd, inv = fwd.cpu().numpy(), inv.cpu().numpy()
    for i in range(len(src)):
        w = np.zeros((n, n), np.int16); b = np.zeros((n, n), np.int16); bw = np.zeros((n, n), np.int16)
        O.kvzo_fwd_transform(n, dst, 8, ptr(src[i]), ptr(w))
        O.kvzo_inv_transform(n, dst, 8, ptr(w), ptr(b))
        O.kvzo_inv_transform(n, dst, 8, ptr(wild[i]), ptr(bw))
        assert np.array_equal(fwd[i], w) and np.array_equal(inv[i], b) and np.array_equal(inv_wild[i], bw), i


@pytest.mark.parametrize("n", [4, 8, 16, 32])
def test_quant_dequant(n):
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(20 + n)
    for qp in (0, 17, 22, 27, 32, 37, 51):
        for intra in (0, 1):
            for signhide in (0, 1):
                for type_ in (0, 2):
                    scan = int(rs.integers(0, 3)) if n <= 8 else 0
                    scale = [40, 400, 4000][int(rs.integers(0, 3))]
                    coef = (rs.standard_normal((12, n, n)) * scale).clip(-32000, 32000).astype(np.int16)
                    coef[0] = 0
                    got = blocks.quant_blocks(torch.from_numpy(coef).cuda(), n, qp, type_, scan, intra, signhide)
                    deq = blocks.dequant_blocks(got, n, qp, type_).cpu().numpy()
                    sums = blocks.coeff_abs_sums(got, n * n).cpu().numpy()
                    got = got.cpu().numpy()
                    oc = kvzlibs.OracleQCfg(qp, 8, intra, signhide, intra)
                    for i in range(len(coef)):
                        w = np.zeros((n, n), np.int16); d = np.zeros((n, n), np.int16)
                        O.kvzo_quant(C.byref(oc), ptr(coef[i]), ptr(w), n, type_, scan)
                        O.kvzo_dequant(C.byref(oc), ptr(w), ptr(d), n, type_)
                        assert np.array_equal(got[i], w), ("quant", qp, intra, signhide, type_, scan, i)
                        assert np.array_equal(deq[i], d), ("dequant", qp, type_, i)
                        assert sums[i] == O.kvzo_coeff_abs_sum(ptr(w), n * n)


def test_coeff_abs_sum_known_answer():
    import torch
    from kvazaar_b200 import blocks
    c = kat.coeff_series()
    got = blocks.coeff_abs_sums(torch.from_numpy(c).cuda(), c.size).cpu().numpy()
    assert int(got[0]) == int(np.abs(c.astype(np.int64)).sum())


@pytest.mark.parametrize("bitdepth", [8, 10])
def test_fused_quantize_residual(bitdepth):
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(5)
    W, H = 256, 128
    f = kvzlibs.synth_luma(9, W, H, 2)
    ref, pred = f[1], f[0]
    pix = np.uint8
    if bitdepth > 8:
        pix = np.uint16
        ref = np.ascontiguousarray((ref.astype(np.uint16) << 2) | rs.integers(0, 4, ref.shape, dtype=np.uint16))
        pred = np.ascontiguousarray((pred.astype(np.uint16) << 2) | rs.integers(0, 4, pred.shape, dtype=np.uint16))
    dev = lambda a: torch.from_numpy(a.view(np.int16) if a.dtype == np.uint16 else a).cuda()
    fn = O.kvzo_quantize_residual_8 if bitdepth == 8 else O.kvzo_quantize_residual_10
    for intra_slice in (0, 1):
        for signhide in (0, 1):
            rows, coff = [], 0
            # 120 jobs of mixed sizes, then 70 that are all 4x4: groups of 32 jobs that are all 4x4 take the block-per-thread
            # kernel (one full group, one ragged one), the mixed groups the group kernel
            for k in range(190):
                n = int(rs.choice([4, 8, 16, 32])) if k < 120 else 4
                x, y = int(rs.integers(0, (W - n) // 4 + 1)) * 4, int(rs.integers(0, (H - n) // 4 + 1)) * 4
                color = int(rs.integers(0, 3)) if n < 32 else 0
                rows.append((y * W + x, y * W + x, y * W + x, coff, W, W, n, color,
                             int(rs.integers(0, 3)) if n <= 8 else 0, int(n == 4 and rs.random() < 0.3),
                             int(rs.integers(8, 48)), int(rs.integers(0, 2)), (0, 0)))
                coff += n * n
            jobs = np.array(rows, blocks.RESIDUAL_JOB_DT)
            # overlapping rec blocks would race: give every job its own rec plane row band instead
            d_rec = torch.zeros((len(jobs), H, W), dtype=torch.uint8 if bitdepth == 8 else torch.int16, device="cuda")
            jobs["rec_offset"] += (np.arange(len(jobs)) * H * W).astype(np.uint32)
            d_coeff = torch.zeros(coff, dtype=torch.int16, device="cuda")
            has = blocks.quantize_residuals(dev(ref), dev(pred), d_rec, d_coeff, jobs, intra_slice, signhide, bitdepth).cpu().numpy()
            rec, coeff = d_rec.cpu().numpy().view(pix), d_coeff.cpu().numpy()
            for i, jb in enumerate(jobs):
                n = int(jb["width"])
                oc = kvzlibs.OracleQCfg(int(jb["qp"]), bitdepth, intra_slice, signhide, int(jb["cu_is_intra"]))
                rw = np.zeros((n, n), pix); cw = np.zeros(n * n, np.int16)
                off = int(jb["ref_offset"])
                hw = fn(C.byref(oc), n, int(jb["color"]), int(jb["scan_idx"]),
                                                int(jb["use_trskip"]), W, n, ptr(ref.reshape(-1)[off:]),
                                                ptr(pred.reshape(-1)[off:]), ptr(rw), ptr(cw))
                y, x = off // W, off % W
                assert has[i] == hw, i
                assert np.array_equal(coeff[int(jb["coeff_offset"]):int(jb["coeff_offset"]) + n * n], cw), ("coeff", i, jb)
                assert np.array_equal(rec[i, y:y + n, x:x + n], rw), ("rec", i, jb)
