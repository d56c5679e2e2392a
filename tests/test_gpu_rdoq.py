"""A15 on the GPU: kvz_cuda_rdoq_blocks against the oracle (pinned on the reference, tests/test_oracle_rdoq.py)
and, where oracle/_ref travelled, against the compiled reference itself."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
import rdoq_cases
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
