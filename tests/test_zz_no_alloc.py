"""kvz_cuda_quantize_residuals_rdoq_noalloc (include/kvz_cuda.h): the fused residual coding with RDOQ whose scratch is the
library's built-in device array instead of a caller-supplied d_work, for callers that cannot call cudaMalloc.  It has no
counterpart in the reference; its outputs must equal kvz_cuda_quantize_residuals_rdoq's bit for bit (integer work).

Collected last on purpose: it must pass alone in a fresh process and after the others."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs

pytestmark = pytest.mark.gpu


def _batch(rs, W, H, n_blocks, n_ctx):
    from kvazaar_b200 import blocks
    rows, rq, coff = [], [], 0
    for _ in range(n_blocks):
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
    jobs["rec_offset"] += (np.arange(len(jobs)) * H * W).astype(np.uint32)
    ctxs = np.frombuffer(rs.integers(0, 128, n_ctx * 148, dtype=np.uint8).tobytes(), blocks.RDOQ_CTX_DT).copy()
    ctxs["reserved"] = 0
    return jobs, np.array(rq, blocks.RDOQ_JOB_DT), ctxs, coff


def _dev(a):
    import torch
    return torch.from_numpy(np.frombuffer(a.tobytes(), np.uint8).copy()).cuda()


def test_noalloc_entry_point_matches_existing_one_twice_in_a_row():
    import torch
    from kvazaar_b200 import blocks, capi
    capi.init(torch.cuda.current_device())
    W, H = 256, 128
    f = kvzlibs.synth_luma(21, W, H, 3)
    for call, (seed, signhide, rdoq_skip) in enumerate([(31, 1, 0), (32, 0, 1)]):     # two calls in a row, different batches:
        rs = np.random.default_rng(seed)                                              # the second finds the first's scratch
        jobs, rjobs, ctxs, coff = _batch(rs, W, H, 120, 5)
        d_ref, d_pred = torch.from_numpy(f[call + 1]).cuda(), torch.from_numpy(f[call]).cuda()
        max_width = int(jobs["width"].max())
        # existing entry point (caller-supplied d_work)
        rec_a = torch.zeros((len(jobs), H, W), dtype=torch.uint8, device="cuda")
        coeff_a = torch.zeros(coff, dtype=torch.int16, device="cuda")
        has_a = blocks.quantize_residuals_rdoq(d_ref, d_pred, rec_a, coeff_a, jobs, ctxs, rjobs, signhide=signhide,
                                               rdoq_skip=rdoq_skip)
        # new entry point: everything it touches is allocated by the caller beforehand
        rec_b = torch.zeros_like(rec_a)
        coeff_b = torch.zeros_like(coeff_a)
        has_b = torch.zeros(len(jobs), dtype=torch.uint8, device="cuda")
        d_jobs, d_rj, d_ctx = _dev(jobs), _dev(rjobs), _dev(ctxs)
        torch.cuda.synchronize()
        blocks.quantize_residuals_rdoq_noalloc(d_ref, d_pred, rec_b, coeff_b, d_jobs, len(jobs), has_b, d_ctx, d_rj, max_width,
                                               signhide=signhide, rdoq_skip=rdoq_skip)
        torch.cuda.synchronize()
        assert torch.equal(has_a, has_b), call
        assert torch.equal(coeff_a, coeff_b), call
        assert torch.equal(rec_a, rec_b), call
        assert int(has_a.sum()) > 10 and int((coeff_a != 0).sum()) > 100          # the comparison is not of zeros


def test_noalloc_entry_point_rejects_sizes_beyond_its_scratch():
    import torch
    from kvazaar_b200 import blocks, capi
    capi.init(torch.cuda.current_device())
    L = capi.lib()
    cfg = capi.QuantCfg(0, 8, 0, 0, 0)
    d = torch.zeros(64, dtype=torch.int16, device="cuda")
    p = C.c_void_p(d.data_ptr())
    ok = L.kvz_cuda_quantize_residuals_rdoq_noalloc(C.byref(cfg), p, p, p, p, blocks.NOALLOC_MAX_COEFFS + 1, p, 1, p, p, p,
                                                    8, 0, None)
    assert ok == 0
    msg = L.kvz_cuda_last_error().decode()
    assert "exceed the built-in scratch" in msg and str(blocks.NOALLOC_MAX_COEFFS) in msg
