"""Golden vectors produced by the UNMODIFIED reference (tests/golden/make_golden.py, run where
/root/reference and oracle/_ref exist) checked against the oracle on CPU and, on the GPU box,
against the CUDA path through the C ABI.  Neither test needs the reference at run time."""
import ctypes as C
import os

import numpy as np
import pytest

import kvzlibs
import me_oracle
from kvzlibs import ptr

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors.npz"))
W, H, RANGE, NREFS, LAMBDA, FME = 128, 72, 8, 2, 23, 4          # as in make_golden.py
ME_CFG = {"free": {}, "owfwpp": {"owf": 1, "wpp": 1}, "mvc4": {"mv_constraint": 4}}


def _me_inputs():
    from kvazaar_b200.frame import ME_JOB_DT, ME_PU_DT
    pus = G["me_pus"].view(ME_PU_DT).copy()
    jobs = G["me_jobs"].view(ME_JOB_DT).copy()
    cur = np.ascontiguousarray(G["me_cur"])
    refs = [np.ascontiguousarray(r) for r in G["me_refs"]]
    return pus, jobs, cur, refs


def _me_want(name):
    from kvazaar_b200.frame import ME_RESULT_DT
    return G["me_res_" + name].view(ME_RESULT_DT)


# ---------------------------------------------------------------- CPU: oracle vs golden
@pytest.mark.parametrize("name", list(ME_CFG))
def test_oracle_motion_search_matches_reference_vectors(oracle_lib, name):
    pus, jobs, cur, refs = _me_inputs()
    got = me_oracle.oracle_search(pus, jobs, cur, refs, W, H, RANGE, FME, LAMBDA, **ME_CFG[name])
    me_oracle.assert_results_equal(got, _me_want(name), pus)


def test_oracle_block_costs_match_reference_vectors(oracle_lib):
    a, b = G["cost_a"], G["cost_b"]
    for i, n in enumerate((4, 8, 16, 32, 64)):
        x, y = np.ascontiguousarray(a[:n, :n]), np.ascontiguousarray(b[:n, :n])
        assert oracle_lib.kvzo_sad_nxn_8(n, ptr(x), ptr(y)) == int(G["cost_sad_nxn"][i])
        assert oracle_lib.kvzo_satd_nxn_8(n, ptr(x), ptr(y)) == int(G["cost_satd_nxn"][i])


@pytest.mark.parametrize("n,dst", [(4, 0), (4, 1), (8, 0), (16, 0), (32, 0)])
def test_oracle_transform_quant_match_reference_vectors(oracle_lib, n, dst):
    k = f"tr_{n}_{dst}_"
    src = np.ascontiguousarray(G[k + "in"])
    coef = np.zeros((n, n), np.int16); back = np.zeros((n, n), np.int16)
    oracle_lib.kvzo_fwd_transform(n, dst, 8, ptr(src), ptr(coef))
    assert np.array_equal(coef, G[k + "coef"])
    oracle_lib.kvzo_inv_transform(n, dst, 8, ptr(coef), ptr(back))
    assert np.array_equal(back, G[k + "back"])
    oc = kvzlibs.OracleQCfg(27, 8, 0, 1, 0)
    q = np.zeros((n, n), np.int16); dq = np.zeros((n, n), np.int16)
    oracle_lib.kvzo_quant(C.byref(oc), ptr(coef), ptr(q), n, 0, 0)
    assert np.array_equal(q, G[k + "q"])
    oracle_lib.kvzo_dequant(C.byref(oc), ptr(q), ptr(dq), n, 0)
    assert np.array_equal(dq, G[k + "dq"])


# ---------------------------------------------------------------- GPU: CUDA path vs golden
@pytest.mark.gpu
@pytest.mark.parametrize("name", list(ME_CFG))
@pytest.mark.parametrize("mode", ["windows", "box", "on_demand"])
def test_cuda_motion_search_matches_reference_vectors(name, mode):
    pus, jobs, cur, refs = _me_inputs()
    kw = dict(ME_CFG[name])
    if mode == "box":
        kw["box"] = (-2 * RANGE, -2 * RANGE, 4 * RANGE + 1, 4 * RANGE + 1)
    got = me_oracle.gpu_search(pus, jobs, cur, refs, W, H, RANGE, FME, LAMBDA, use_maps=mode != "on_demand", **kw)
    me_oracle.assert_results_equal(got, _me_want(name), pus)


@pytest.mark.gpu
def test_cuda_block_costs_match_reference_vectors():
    import torch
    from kvazaar_b200 import blocks
    a, b = G["cost_a"], G["cost_b"]
    jobs = np.zeros(10, blocks.BLOCK_JOB_DT)
    for i, n in enumerate((4, 8, 16, 32, 64)):
        for j, kind in enumerate((blocks.COST_SAD_NXN, blocks.COST_SATD_NXN)):
            jb = jobs[2 * i + j]
            jb["a_offset"] = jb["b_offset"] = 0
            jb["a_stride"] = jb["b_stride"] = 64
            jb["w"] = jb["h"] = n
            jb["kind"] = kind
    out = blocks.block_costs(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), jobs).cpu().numpy()
    assert np.array_equal(out[0::2], G["cost_sad_nxn"])
    assert np.array_equal(out[1::2], G["cost_satd_nxn"])


@pytest.mark.gpu
@pytest.mark.parametrize("n,dst", [(4, 0), (4, 1), (8, 0), (16, 0), (32, 0)])
def test_cuda_transform_quant_match_reference_vectors(n, dst):
    import torch
    from kvazaar_b200 import blocks
    k = f"tr_{n}_{dst}_"
    src = torch.from_numpy(np.ascontiguousarray(G[k + "in"])[None]).cuda()
    coef = blocks.transform_blocks(src, n, inverse=False, dst4=bool(dst))
    assert np.array_equal(coef.cpu().numpy()[0], G[k + "coef"])
    back = blocks.transform_blocks(coef, n, inverse=True, dst4=bool(dst))
    assert np.array_equal(back.cpu().numpy()[0], G[k + "back"])
    q = blocks.quant_blocks(coef, n, 27, type_=0, scan_idx=0, slice_is_intra=False, signhide=True)
    assert np.array_equal(q.cpu().numpy()[0], G[k + "q"])
    dq = blocks.dequant_blocks(q, n, 27, type_=0)
    assert np.array_equal(dq.cpu().numpy()[0], G[k + "dq"])


# ---- A15: RDOQ vectors of the compiled reference (tests/golden/make_golden_rdoq.py) ----
RDOQ_PARAMS = ("width", "type", "scan_idx", "cu_is_intra", "tr_depth", "qp", "signhide", "slice_is_intra")


def _rdoq_cases():
    import kvzlibs as kl
    v = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rdoq_vectors.npz"))
    at = 0
    for prm, lam, ctx in zip(v["params"], v["lam"], v["ctx"]):
        case = dict(zip(RDOQ_PARAMS, (int(x) for x in prm)))
        nn = case["width"] ** 2
        case["lam"], case["coef"] = float(lam), v["coef"][at:at + nn].copy()
        yield case, kl.RdoqCtx.from_buffer_copy(ctx.tobytes()), v["levels"][at:at + nn]
        at += nn


def test_oracle_rdoq_matches_reference_vectors(oracle_lib):
    from test_oracle_rdoq import run_oracle
    n = 0
    for case, ctx, want in _rdoq_cases():
        assert np.array_equal(run_oracle(case, ctx), want), {k: v for k, v in case.items() if k != "coef"}
        n += 1
    assert n > 60


@pytest.mark.gpu
def test_cuda_rdoq_matches_reference_vectors():
    from test_gpu_rdoq import gpu_rdoq
    for sh in (0, 1):
        sel = [(c, x, w) for c, x, w in _rdoq_cases() if c["signhide"] == sh]
        got = gpu_rdoq([c for c, _, _ in sel], [x for _, x, _ in sel], sh)
        for (c, _, want), g in zip(sel, got):
            assert np.array_equal(g, want), {k: v for k, v in c.items() if k != "coef"}
