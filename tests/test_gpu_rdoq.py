"""A15 on the GPU: kvz_cuda_rdoq_blocks against the oracle (pinned on the reference, tests/test_oracle_rdoq.py)
and, where oracle/_ref travelled, against the compiled reference itself."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
import rdoq_cases
from kvzlibs import ptr
from test_oracle_rdoq import run_oracle, run_ref

pytestmark = pytest.mark.gpu


def gpu_rdoq(case_list, ctx_list, signhide):
    import torch
    from kvazaar_b200 import blocks
    jobs = np.zeros(len(case_list), blocks.RDOQ_JOB_DT)
    ctxs = np.zeros(len(ctx_list), blocks.RDOQ_CTX_DT)
    offs, total = [], 0
    for i, (cs, cx) in enumerate(zip(case_list, ctx_list)):
        ctxs[i] = np.frombuffer(bytes(cx), blocks.RDOQ_CTX_DT)[0]
        offs.append(total)
        jobs[i] = (total, i, cs["lam"], cs["width"], cs["type"], cs["scan_idx"], cs["tr_depth"], cs["qp"],
                   cs["cu_is_intra"], (0, 0))
        total += cs["width"] ** 2
    coef = np.concatenate([cs["coef"] for cs in case_list])
    out = blocks.rdoq_blocks(torch.from_numpy(coef).cuda(), ctxs, jobs, signhide=signhide).cpu().numpy()
    return [out[o:o + cs["width"] ** 2] for o, cs in zip(offs, case_list)]


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
@pytest.mark.parametrize("signhide", [0, 1])
def test_rdoq_matches_oracle(seed, signhide):
    rs = np.random.default_rng(300 + seed)
    case_list = [dict(cs, signhide=signhide) for cs in rdoq_cases.cases(seed)]     # signhide is per launch (cfg)
    ctx_list = [kvzlibs.random_rdoq_ctx(rs) for _ in case_list]
    got = gpu_rdoq(case_list, ctx_list, signhide)
    nonzero = 0
    for cs, cx, g in zip(case_list, ctx_list, got):
        want = run_oracle(cs, cx)
        assert np.array_equal(g, want), {k: v for k, v in cs.items() if k != "coef"}
        nonzero += int(np.any(want != 0))
    assert nonzero > len(case_list) // 3


@pytest.mark.skipif(not kvzlibs.have_ref(), reason="oracle/_ref not present")
def test_rdoq_matches_compiled_reference():
    R = kvzlibs.ref(1)
    case_list = [dict(cs, signhide=1) for cs in rdoq_cases.cases(11, per_shape=1)]
    ctx_list = []
    for cs in case_list:
        cx = kvzlibs.RdoqCtx()
        R.ref_shim_init_rdoq_ctx(cs["qp"], cs["slice_is_intra"], C.byref(cx))
        ctx_list.append(cx)
    got = gpu_rdoq(case_list, ctx_list, 1)
    for cs, cx, g in zip(case_list, ctx_list, got):
        assert np.array_equal(g, run_ref(cs, cx)), {k: v for k, v in cs.items() if k != "coef"}


def test_rdoq_frame_batch_properties():
    """A whole 1080p frame of 8x8 luma blocks (32400 blocks, one launch): RDOQ never raises a level above the
    rounded quantiser's, and an all-zero block stays zero."""
    import torch
    from kvazaar_b200 import blocks
    rs = np.random.default_rng(8)
    n = 32400
    yy, xx = np.mgrid[0:8, 0:8]
    coef = np.rint(rs.normal(0, 1, (n, 8, 8)) * (500.0 / (1 + xx + yy))).astype(np.int16)
    coef[::17] = 0
    jobs = np.zeros(n, blocks.RDOQ_JOB_DT)
    jobs["coef_offset"] = np.arange(n) * 64
    jobs["ctx_index"] = np.arange(n) % 7
    jobs["lambda"] = 40.0
    jobs["width"], jobs["qp"], jobs["cu_is_intra"] = 8, 32, 1
    ctxs = np.frombuffer(rs.integers(0, 128, 7 * 148, dtype=np.uint8).tobytes(), blocks.RDOQ_CTX_DT).copy()
    out = blocks.rdoq_blocks(torch.from_numpy(coef.reshape(-1)).cuda(), ctxs, jobs, signhide=False).cpu().numpy()
    out = out.reshape(n, 8, 8)
    q_bits = 14 + 32 // 6 + (15 - 8 - 3)
    top = (np.abs(coef.astype(np.int64)) * 20560 + (1 << (q_bits - 1))) >> q_bits        # qp 32 -> scale[2]
    assert np.all(np.abs(out.astype(np.int64)) <= top)
    assert np.all(np.sign(out) * np.sign(coef) >= 0)
    assert not out[::17].any() and out.any()


@pytest.mark.parametrize("bitdepth", [8, 10])
@pytest.mark.parametrize("rdoq_skip", [0, 1])
def test_fused_quantize_residual_with_rdoq(rdoq_skip, bitdepth):
    """kvz_cuda_quantize_residuals_rdoq == kvz_quantize_residual with cfg.rdoq_enable (oracle, pinned on the reference)"""
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    fn = O.kvzo_quantize_residual_rdoq_8 if bitdepth == 8 else O.kvzo_quantize_residual_rdoq_10
    fn.argtypes = [C.c_void_p, C.c_double, C.c_void_p] + [C.c_int] * 8 + [C.c_void_p] * 4
    rs = np.random.default_rng(15 + rdoq_skip)
    W, H = 256, 128
    f = kvzlibs.synth_luma(9, W, H, 2)
    ref, pred = f[1], f[0]
    pix = np.uint8
    if bitdepth > 8:
        pix = np.uint16
        ref = np.ascontiguousarray((ref.astype(np.uint16) << 2) | rs.integers(0, 4, ref.shape, dtype=np.uint16))
        pred = np.ascontiguousarray((pred.astype(np.uint16) << 2) | rs.integers(0, 4, pred.shape, dtype=np.uint16))
    dev = lambda a: torch.from_numpy(a.view(np.int16) if a.dtype == np.uint16 else a).cuda()
    for signhide in (0, 1):
        rows, rq, coff = [], [], 0
        n_ctx = 5
        ctx_list = [kvzlibs.random_rdoq_ctx(rs) for _ in range(n_ctx)]
        ctxs = np.zeros(n_ctx, blocks.RDOQ_CTX_DT)
        for i, cx in enumerate(ctx_list):
            ctxs[i] = np.frombuffer(bytes(cx), blocks.RDOQ_CTX_DT)[0]
        for _ in range(100):
            n = int(rs.choice([4, 8, 16, 32]))
            x, y = int(rs.integers(0, (W - n) // 4 + 1)) * 4, int(rs.integers(0, (H - n) // 4 + 1)) * 4
            color = int(rs.integers(0, 3)) if n < 32 else 0
            scan = int(rs.integers(0, 3)) if n <= 8 else 0
            qp, intra = int(rs.integers(8, 48)), int(rs.integers(0, 2))
            rows.append((y * W + x, y * W + x, y * W + x, coff, W, W, n, color, scan, int(n == 4 and rs.random() < 0.3),
                         qp, intra, (0, 0)))
            rq.append((coff, int(rs.integers(0, n_ctx)), float(0.57 * 2.0 ** ((qp - 12) / 3.0)), n, 0 if color == 0 else 2,
                       scan, int(rs.integers(0, 3)), qp, intra, (0, 0)))
            coff += n * n
        jobs = np.array(rows, blocks.RESIDUAL_JOB_DT)
        rjobs = np.array(rq, blocks.RDOQ_JOB_DT)
        d_rec = torch.zeros((len(jobs), H, W), dtype=torch.uint8 if bitdepth == 8 else torch.int16, device="cuda")
        jobs["rec_offset"] += (np.arange(len(jobs)) * H * W).astype(np.uint32)
        d_coeff = torch.zeros(coff, dtype=torch.int16, device="cuda")
        has = blocks.quantize_residuals_rdoq(dev(ref), dev(pred), d_rec, d_coeff,
                                             jobs, ctxs, rjobs, slice_is_intra=False, signhide=signhide,
                                             rdoq_skip=rdoq_skip, bitdepth=bitdepth).cpu().numpy()
        rec, coeff = d_rec.cpu().numpy().view(pix), d_coeff.cpu().numpy()
        for i, (jb, rj) in enumerate(zip(jobs, rjobs)):
            n = int(jb["width"])
            oc = kvzlibs.OracleQCfg(int(jb["qp"]), bitdepth, 0, signhide, int(jb["cu_is_intra"]))
            rw = np.zeros((n, n), pix); cw = np.zeros(n * n, np.int16)
            off = int(jb["ref_offset"])
            cx = ctx_list[int(rj["ctx_index"])]
            hw = fn(C.addressof(oc), float(rj["lambda"]), C.addressof(cx), rdoq_skip,
                                                 int(rj["tr_depth"]), n, int(jb["color"]), int(jb["scan_idx"]),
                                                 int(jb["use_trskip"]), W, n, ptr(ref.reshape(-1)[off:]),
                                                 ptr(pred.reshape(-1)[off:]), ptr(rw), ptr(cw))
            y, x = off // W, off % W
            assert has[i] == hw, i
            assert np.array_equal(coeff[int(jb["coeff_offset"]):int(jb["coeff_offset"]) + n * n], cw), ("coeff", i, jb)
            assert np.array_equal(rec[i, y:y + n, x:x + n], rw), ("rec", i, jb)
