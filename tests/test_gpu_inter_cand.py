"""GPU parity: kvz_cuda_inter_candidates (merge list, AMVP predictors, colocated start vector: inter.c:713-1446,
search_inter.c:1142-1206) against the oracle, through the C ABI.  Bit-exact; and the records it fills drive
kvz_cuda_me_search to the same decisions as host-derived candidates."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
import cand_oracle as co
from test_oracle_cand import oracle_merge, oracle_mv_cand, SCENES

pytestmark = pytest.mark.gpu


def _records(sc, pus):
    from kvazaar_b200.frame import ME_PU_DT, ME_JOB_DT
    refs = sorted(set(sc.lists[0] + sc.lists[1]))
    P = np.zeros(len(pus), ME_PU_DT)
    J = np.zeros(len(pus) * len(refs), ME_JOB_DT)
    flags = np.zeros(len(pus), np.uint8)
    for i, (x, y, w, h, a1, b1) in enumerate(pus):
        P[i]["x"], P[i]["y"], P[i]["w"], P[i]["h"] = x, y, w, h
        P[i]["first_job"], P[i]["num_jobs"] = i * len(refs), len(refs)
        flags[i] = (0 if a1 else 1) | (0 if b1 else 2)
        for q, r in enumerate(refs):
            J[i * len(refs) + q]["ref_idx"] = r
    return P, J, flags, refs


def _oracle_fill(sc, f, pus, P, J, refs):
    O = kvzlibs.oracle()
    for i, pu in enumerate(pus):
        n, mc = oracle_merge(f, pu)
        P[i]["num_merge_cand"] = n
        for k in range(5):
            P[i]["merge_cand"][k]["dir"], P[i]["merge_cand"][k]["ref"], P[i]["merge_cand"][k]["mv"] = mc[k]["dir"], mc[k]["ref"], mc[k]["mv"]
        for q, r in enumerate(refs):
            l, idx = (0, sc.lists[0].index(r)) if r in sc.lists[0] and (r not in sc.lists[1] or sc.lists[0].index(r) <= sc.lists[1].index(r)) \
                else (1, sc.lists[1].index(r))
            J[i * len(refs) + q]["mv_cand"] = oracle_mv_cand(f, pu, l, idx)
            ex = np.zeros(2, np.int16)
            O.kvzo_colocated_start_mv(C.byref(f), *pu[:4], r, ex.ctypes.data_as(C.c_void_p))
            J[i * len(refs) + q]["extra_mv"] = ex


def _gpu_fill(sc, P, J, flags):
    import torch
    from kvazaar_b200.frame import inter_candidates
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
    d_p, d_j = dev(P), dev(J)
    inter_candidates(sc, dev(sc.scus), [dev(r) for r in sc.ref_scus], d_p, len(P), d_j, dev(flags))
    torch.cuda.synchronize()
    return d_p.cpu().numpy().view(P.dtype), d_j.cpu().numpy().view(J.dtype)


@pytest.mark.parametrize("scene", SCENES, ids=[str(i) for i in range(len(SCENES))])
def test_candidates_match_oracle(scene):
    kw = {k: v for k, v in scene.items() if k != "size"}
    rs = np.random.default_rng(scene["size"][1] + len(kw))
    sc = co.Scene(rs, scene["size"][0], scene["size"][1], **kw)
    f = sc.frame()
    pus = sc.pus(rs, 400)
    P, J, flags, refs = _records(sc, pus)
    want_p, want_j = P.copy(), J.copy()
    _oracle_fill(sc, f, pus, want_p, want_j, refs)
    got_p, got_j = _gpu_fill(sc, P, J, flags)
    assert np.array_equal(got_p["num_merge_cand"], want_p["num_merge_cand"])
    for name in ("dir", "ref", "mv"):
        bad = np.nonzero((got_p["merge_cand"][name] != want_p["merge_cand"][name]).reshape(len(P), -1).any(axis=1))[0]
        assert len(bad) == 0, (name, pus[bad[0]], got_p[bad[0]]["merge_cand"], want_p[bad[0]]["merge_cand"])
    for name in ("mv_cand", "extra_mv"):
        bad = np.nonzero((got_j[name] != want_j[name]).reshape(len(J), -1).any(axis=1))[0]
        assert len(bad) == 0, (name, pus[bad[0] // len(refs)], got_j[bad[0]], want_j[bad[0]])
    assert (want_j["extra_mv"] != 0).any() and (want_j["mv_cand"] != 0).any()


def test_frame_of_pus_feeds_me_search():
    """1080p: all 2Nx2N PUs of the frame through the candidate kernel; the filled records are a valid
    kvz_cuda_me_search input (sampled against the oracle) and every list has five candidates"""
    rs = np.random.default_rng(77)
    sc = co.Scene(rs, 1920, 1080 - 1080 % 8, slice_b=False, n_refs=4)
    pus = []
    for size in (8, 16, 32, 64):
        for y in range(0, sc.height - size + 1, size):
            for x in range(0, sc.width - size + 1, size):
                pus.append((x, y, size, size, 1, 1))
    P, J, flags, refs = _records(sc, pus)
    got_p, got_j = _gpu_fill(sc, P, J, flags)
    assert (got_p["num_merge_cand"] == 5).all()
    f = sc.frame()
    pick = rs.choice(len(pus), 300, replace=False)
    sub = [pus[i] for i in pick]
    sp, sj, _, _ = _records(sc, sub)
    _oracle_fill(sc, f, sub, sp, sj, refs)
    for k, i in enumerate(pick):
        assert got_p[i]["merge_cand"].tobytes() == sp[k]["merge_cand"].tobytes()
        assert np.array_equal(got_j[i * len(refs):(i + 1) * len(refs)]["mv_cand"], sj[k * len(refs):(k + 1) * len(refs)]["mv_cand"])


def test_device_candidates_drive_me_search_to_the_oracle_decisions():
    """A18 -> A3-A5 composed on the device: candidates derived by kvz_cuda_inter_candidates, then kvz_cuda_me_search
    on those records, against the oracle's candidate derivation followed by the oracle's search"""
    import me_oracle
    W, H, n_refs, rng = 200, 136, 2, 8
    rs = np.random.default_rng(91)
    sc = co.Scene(rs, W, H, slice_b=False, n_refs=n_refs, p_intra=0.2)
    f = sc.frame()
    pus = [p for p in sc.pus(rs, 160) if p[2] % 8 == 0 and p[3] % 8 == 0][:60]
    P, J, flags, refs = _records(sc, pus)
    want_p, want_j = P.copy(), J.copy()
    _oracle_fill(sc, f, pus, want_p, want_j, refs)
    got_p, got_j = _gpu_fill(sc, P, J, flags)
    frames = kvzlibs.synth_luma(17, W, H, n_refs + 1)
    cur, planes = frames[n_refs], [frames[n_refs - 1 - i] for i in range(n_refs)]
    ref_LX = [[i % n_refs for i in range(16)], [0] * 16]
    got = me_oracle.gpu_search(got_p, got_j, cur, planes, W, H, rng, 4, 20, use_maps=False, ref_LX=ref_LX)
    want = me_oracle.oracle_search(want_p, want_j, cur, planes, W, H, rng, 4, 20, ref_LX=ref_LX)
    me_oracle.assert_results_equal(got, want, want_p)
    assert got["merged"].any() or (got["mv"] != 0).any()
