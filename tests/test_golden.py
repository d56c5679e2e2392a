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


# ---- deblocking, SAO, coefficient bits, merge / AMVP candidates, checksum: vectors of the compiled reference
# ---- (tests/golden/make_golden_filters.py) ----
def _fv():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "filter_vectors.npz"))


DEBLOCK_CASES = [dict(size=(136, 72), qp=32), dict(size=(200, 136), qp=37, slice_b=1),
                 dict(size=(136, 136), qp=27, cu_qp=1, beta=2, tc=-1)]


def _deblock_case(v, i):
    import deblock_oracle as do
    case = DEBLOCK_CASES[i]
    w, h = case["size"]
    scus = np.ascontiguousarray(v[f"db{i}_scus"]).view(do.SCU_DT).reshape(h // 4, w // 4)
    planes = [np.ascontiguousarray(v[f"db{i}_in{k}"]) for k in range(3)]
    want = [v[f"db{i}_out{k}"] for k in range(3)]
    return case, scus, planes, want


@pytest.mark.parametrize("i", [0, 1, 2])
def test_oracle_deblock_matches_reference_vectors(oracle_lib, i):
    import deblock_oracle as do
    from test_oracle_deblock import oracle_deblock
    case, scus, planes, want = _deblock_case(_fv(), i)
    cfg = do.make_cfg(case["size"][0], case["size"][1], qp=case["qp"], cu_qp=case.get("cu_qp", 0), beta=case.get("beta", 0),
                      tc=case.get("tc", 0), slice_b=case.get("slice_b", 0))
    got = oracle_deblock(cfg, scus, planes)
    for k in range(3):
        assert np.array_equal(got[k], want[k]), k


def _sao_case(v):
    import sao_oracle as so
    orig = [np.ascontiguousarray(v[f"sao_orig{k}"]) for k in range(3)]
    rec = [np.ascontiguousarray(v[f"sao_rec{k}"]) for k in range(3)]
    luma = np.ascontiguousarray(v["sao_luma_params"]).view(so.SAO_PARAMS_DT).reshape(-1)
    chroma = np.ascontiguousarray(v["sao_chroma_params"]).view(so.SAO_PARAMS_DT).reshape(-1)
    return orig, rec, luma, chroma, v["sao_edge_stats"], [v[f"sao_out{k}"] for k in range(3)]


def test_oracle_sao_matches_reference_vectors(oracle_lib):
    import sao_oracle as so
    orig, rec, luma, chroma, edge, want = _sao_case(_fv())
    st = np.concatenate(so.oracle_ctu_stats(orig, rec))
    assert np.array_equal(st["edge"], edge)
    got = so.oracle_reconstruct(rec, luma, chroma)
    for k in range(3):
        assert np.array_equal(got[k], want[k]), k


def _coeff_cases(v):
    at = 0
    for prm, ctx, bits in zip(v["cb_params"], v["cb_ctx"], v["cb_bits"]):
        w, t, s, rng, sh, ts = (int(x) for x in prm)
        blk = np.ascontiguousarray(v["cb_coef"][at:at + w * w].reshape(w, w))
        at += w * w
        yield (w, t, s, rng, sh, ts), kvzlibs.RdoqCtx.from_buffer_copy(ctx.tobytes()), blk, int(bits)


def test_oracle_coeff_bits_match_reference_vectors(oracle_lib):
    from test_oracle_coeff_cost import oracle_bits
    n = 0
    for (w, t, s, rng, sh, ts), ctx, blk, bits in _coeff_cases(_fv()):
        assert oracle_bits(ctx, rng, blk, w, t, s, sh, ts) == bits, (w, t, s, rng, sh, ts)
        n += 1
    assert n == 72


class _GoldenScene:
    """the bookkeeping cand_oracle.Scene carries, rebuilt from the fixture"""

    def __init__(self, v, i):
        import deblock_oracle as do
        m = v[f"cand{i}_meta"]
        self.width, self.height, self.poc = int(m[0]), int(m[1]), int(m[2])
        self.slice_b, self.tmvp, self.n_refs = bool(m[3]), bool(m[4]), int(m[5])
        self.scus = np.ascontiguousarray(v[f"cand{i}_scus"]).view(do.SCU_DT).reshape(self.height // 4, self.width // 4)
        self.ref_scus = [np.ascontiguousarray(r).view(do.SCU_DT).reshape(self.height // 4, self.width // 4) for r in v[f"cand{i}_ref_scus"]]
        self.ref_pocs = [int(x) for x in v[f"cand{i}_ref_pocs"]]
        self.col_ref_pocs = [[int(x) for x in row] for row in v[f"cand{i}_col_ref_pocs"]]
        self.lists = [[int(x) for x in v[f"cand{i}_list0"]], [int(x) for x in v[f"cand{i}_list1"]]]
        self.pus = [tuple(int(x) for x in p) for p in v[f"cand{i}_pus"]]
        self.merge, self.amvp = v[f"cand{i}_merge"], v[f"cand{i}_amvp"]

    def frame(self):
        import cand_oracle as co
        return co.Scene.frame(self)


@pytest.mark.parametrize("i", [0, 1])
def test_oracle_candidates_match_reference_vectors(oracle_lib, i):
    import cand_oracle as co
    from test_oracle_cand import oracle_merge, oracle_mv_cand
    sc = _GoldenScene(_fv(), i)
    f = sc.frame()
    for k, pu in enumerate(sc.pus):
        n, got = oracle_merge(f, pu)
        want = np.ascontiguousarray(sc.merge[k]).view(co.MERGE_DT).reshape(5)
        for name in ("dir", "ref", "mv"):
            assert np.array_equal(got[name], want[name]), (pu, name)
        for l in range(2):
            for idx in range(len(sc.lists[l])):
                assert np.array_equal(oracle_mv_cand(f, pu, l, idx), sc.amvp[k][l * 4 + idx]), (pu, l, idx)


def test_oracle_checksum_matches_reference_vector(oracle_lib):
    from test_oracle_picture_stats import oracle_checksum
    v = _fv()
    assert oracle_checksum(np.ascontiguousarray(v["ck_plane"])) == int(v["ck_sum"][0])


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("i", [0, 1, 2])
def test_cuda_deblock_matches_reference_vectors(i):
    from kvazaar_b200 import blocks
    case, scus, planes, want = _deblock_case(_fv(), i)
    d = [_dev(p) for p in planes]
    blocks.deblock_picture(d, scus.view(blocks.SCU_DT), case["qp"], case.get("cu_qp", 0), case.get("beta", 0), case.get("tc", 0),
                           case.get("slice_b", 0), [[j for j in range(16)], [(j + 1) % 4 for j in range(16)]])
    for k in range(3):
        assert np.array_equal(d[k].cpu().numpy(), want[k]), k


@pytest.mark.gpu
def test_cuda_sao_matches_reference_vectors():
    import torch
    from kvazaar_b200 import blocks
    orig, rec, luma, chroma, edge, want = _sao_case(_fv())
    st = blocks.sao_ctu_stats([_dev(p) for p in orig], [_dev(p) for p in rec]).cpu().numpy().view(blocks.SAO_STATS_DT)
    assert np.array_equal(st["edge"], edge)
    d_in = [_dev(p) for p in rec]
    d_out = [torch.zeros_like(t) for t in d_in]
    blocks.sao_reconstruct(d_in, d_out, luma.view(blocks.SAO_PARAMS_DT), chroma.view(blocks.SAO_PARAMS_DT))
    for k in range(3):
        assert np.array_equal(d_out[k].cpu().numpy(), want[k]), k


@pytest.mark.gpu
def test_cuda_coeff_bits_match_reference_vectors():
    from kvazaar_b200 import blocks
    cases = list(_coeff_cases(_fv()))
    for sh, ts in ((0, 0), (1, 0), (0, 1), (1, 1)):
        sel = [c for c in cases if c[0][4] == sh and c[0][5] == ts]
        ctx_np = np.frombuffer(b"".join(bytes(c[1]) for c in sel), blocks.RDOQ_CTX_DT).copy()
        jobs = np.zeros(len(sel), blocks.COEFF_JOB_DT)
        off = 0
        for k, ((w, t, s, rng, _, _), _, blk, _) in enumerate(sel):
            jobs[k]["coef_offset"], jobs[k]["ctx_index"], jobs[k]["range"] = off, k, rng
            jobs[k]["width"], jobs[k]["type"], jobs[k]["scan_idx"] = w, t, s
            off += w * w
        flat = np.concatenate([c[2].reshape(-1) for c in sel]).astype(np.int16)
        got = blocks.coeff_bits(_dev(flat), ctx_np, jobs, sh, ts).cpu().numpy()
        assert got.tolist() == [c[3] for c in sel]


@pytest.mark.gpu
@pytest.mark.parametrize("i", [0, 1])
def test_cuda_candidates_match_reference_vectors(i):
    import cand_oracle as co
    from test_gpu_inter_cand import _records, _gpu_fill
    sc = _GoldenScene(_fv(), i)
    P, J, flags, refs = _records(sc, sc.pus)
    got_p, got_j = _gpu_fill(sc, P, J, flags)
    for k, pu in enumerate(sc.pus):
        want = np.ascontiguousarray(sc.merge[k]).view(co.MERGE_DT).reshape(5)
        for name in ("dir", "ref", "mv"):
            assert np.array_equal(got_p[k]["merge_cand"][name], want[name]), (pu, name)
        for q, r in enumerate(refs):
            l, idx = (0, sc.lists[0].index(r)) if r in sc.lists[0] and (r not in sc.lists[1] or sc.lists[0].index(r) <= sc.lists[1].index(r)) \
                else (1, sc.lists[1].index(r))
            assert np.array_equal(got_j[k * len(refs) + q]["mv_cand"], sc.amvp[k][l * 4 + idx]), (pu, r)


@pytest.mark.gpu
def test_cuda_checksum_matches_reference_vector():
    from kvazaar_b200 import blocks
    v = _fv()
    got = blocks.picture_checksum([_dev(v["ck_plane"])]).cpu().numpy().view(np.uint32)
    assert int(got[0]) == int(v["ck_sum"][0])
