"""GPU parity: kvz_cuda_coeff_bits (kvz_get_coeff_cost's CABAC path, rdo.c:158-197 + encode_coding_tree.c:103-345)
against the oracle, through the C ABI.  Bit-exact."""
import numpy as np
import pytest

import kvzlibs
import coeff_cases as cc
from test_oracle_coeff_cost import oracle_bits

pytestmark = pytest.mark.gpu


def _batch(rs, shapes, per_shape, n_ctx=16):
    from kvazaar_b200 import blocks
    ctxs = [kvzlibs.random_rdoq_ctx(rs) for _ in range(n_ctx)]
    ctx_np = np.frombuffer(b"".join(bytes(c) for c in ctxs), blocks.RDOQ_CTX_DT).copy()
    blks, jobs, off = [], [], 0
    for (w, t, s) in shapes:
        for b in cc.blocks(rs, w, per_shape):
            blks.append(b)
            jobs.append((off, int(rs.integers(0, n_ctx)), int(rs.integers(256, 511)), w, t, s))
            off += w * w
    j = np.zeros(len(jobs), blocks.COEFF_JOB_DT)
    for i, (o, c, r, w, t, s) in enumerate(jobs):
        j[i]["coef_offset"], j[i]["ctx_index"], j[i]["range"], j[i]["width"], j[i]["type"], j[i]["scan_idx"] = o, c, r, w, t, s
    flat = np.concatenate([b.reshape(-1) for b in blks]).astype(np.int16)
    return ctxs, ctx_np, blks, j, flat


@pytest.mark.parametrize("signhide,trskip", [(0, 0), (1, 0), (1, 1)])
def test_mixed_batch(signhide, trskip):
    import torch
    from kvazaar_b200 import blocks
    rs = np.random.default_rng(10 + signhide + 2 * trskip)
    ctxs, ctx_np, blks, jobs, flat = _batch(rs, cc.SHAPES, 24)
    got = blocks.coeff_bits(torch.from_numpy(flat).cuda(), ctx_np, jobs, signhide, trskip).cpu().numpy()
    for i, b in enumerate(blks):
        jb = jobs[i]
        want = oracle_bits(ctxs[int(jb["ctx_index"])], int(jb["range"]), b, int(jb["width"]), int(jb["type"]), int(jb["scan_idx"]),
                           signhide, trskip)
        assert int(got[i]) == want, (i, jb, int(got[i]), want)
    assert (got > 0).sum() > len(blks) * 0.9


def test_zero_blocks_and_single_widths():
    import torch
    from kvazaar_b200 import blocks
    rs = np.random.default_rng(4)
    for shapes in ([(4, 0, 0)], [(32, 0, 0)], [(8, 2, 0), (4, 2, 1)]):
        ctxs, ctx_np, blks, jobs, flat = _batch(rs, shapes, 12)
        flat[:blks[0].size] = 0                                       # first block all zero
        blks[0][:] = 0
        got = blocks.coeff_bits(torch.from_numpy(flat).cuda(), ctx_np, jobs, True, False).cpu().numpy()
        assert got[0] == 0
        for i, b in enumerate(blks):
            jb = jobs[i]
            assert int(got[i]) == oracle_bits(ctxs[int(jb["ctx_index"])], int(jb["range"]), b, int(jb["width"]), int(jb["type"]),
                                              int(jb["scan_idx"]), 1, 0)


def test_frame_of_quantised_blocks():
    """every 8x8 block of a 416x240 residual through transform + quant on device, then its bit cost"""
    import torch
    from kvazaar_b200 import blocks
    rs = np.random.default_rng(6)
    n, W, H = 8, 416, 240
    cnt = (W // n) * (H // n)
    resid = torch.from_numpy(rs.integers(-40, 41, cnt * n * n).astype(np.int16)).cuda()
    lv = blocks.quant_blocks(blocks.transform_blocks(resid, n), n, 27, signhide=True)
    ctx = kvzlibs.random_rdoq_ctx(rs)
    ctx_np = np.frombuffer(bytes(ctx), blocks.RDOQ_CTX_DT).copy()
    jobs = np.zeros(cnt, blocks.COEFF_JOB_DT)
    jobs["coef_offset"] = np.arange(cnt, dtype=np.uint32) * n * n
    jobs["range"], jobs["width"] = 300, n
    got = blocks.coeff_bits(lv, ctx_np, jobs, True, False).cpu().numpy()
    lv_h = lv.cpu().numpy().reshape(cnt, n, n)
    for i in range(0, cnt, 7):
        assert int(got[i]) == oracle_bits(ctx, 300, np.ascontiguousarray(lv_h[i]), n, 0, 0, 1, 0)
    assert got.max() > 0
