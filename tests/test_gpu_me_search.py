"""GPU parity: kvz_cuda_me_search (integer full search + fractional refinement + merge/AMVP
bookkeeping, search_inter.c:886-1297) against the oracle, through the C ABI.  Bit-exact."""
import numpy as np
import pytest

import kvzlibs
import me_oracle

pytestmark = pytest.mark.gpu


def _run(width, height, rng, fme, n_refs, count, seed, use_maps=True, sizes=(8, 16, 32, 64), centres=(), box=None, **cfg):
    rs = np.random.default_rng(seed)
    f = kvzlibs.synth_luma(seed, width, height, n_refs + 1)
    cur, refs = f[n_refs], [f[n_refs - 1 - i] for i in range(n_refs)]
    pus, jobs = me_oracle.random_pus(rs, width, height, n_refs, count, sizes=sizes, mv_span=rng)
    if centres:
        # make the extra_mv / first merge candidate of most PUs land on the cached window centres
        for i in range(count):
            c = centres[i % len(centres)]
            if i % 4 != 3:
                jobs["extra_mv"][i * n_refs:(i + 1) * n_refs] = (c[0] * 4 + 1, c[1] * 4 + 2)
            if pus[i]["num_merge_cand"] > 0 and pus[i]["merge_cand"][0]["dir"] != 3:
                pus[i]["merge_cand"][0]["mv"][:] = (c[0] * 4, c[1] * 4)
    lam = int(rs.integers(3, 45))
    got = me_oracle.gpu_search(pus, jobs, cur, refs, width, height, rng, fme, lam, use_maps=use_maps,
                               centres=centres, box=box, **cfg)
    want = me_oracle.oracle_search(pus, jobs, cur, refs, width, height, rng, fme, lam, **cfg)
    me_oracle.assert_results_equal(got, want, pus)
    return got


@pytest.mark.parametrize("fme", [0, 1, 2, 3, 4])
def test_small_frame_range8(fme):
    _run(200, 136, 8, fme, 2, 60, 100 + fme)


def test_range16_three_refs():
    got = _run(256, 192, 16, 4, 3, 48, 7)
    assert got["took"].reshape(-1, 3)[:, 0].all()        # the first reference always lowers MAX_INT


def test_owf_wpp_restriction():
    _run(256, 192, 16, 4, 2, 48, 8, owf=1, wpp=1)
    _run(256, 192, 8, 2, 2, 32, 9, owf=1, wpp=1, sao_type=0)


def test_mv_constraint_frame_and_tile_margin():
    _run(200, 136, 8, 4, 2, 40, 10, mv_constraint=4)
    _run(200, 136, 8, 4, 1, 24, 11, mv_constraint=1)


def test_without_dense_maps():
    # zero window evaluated on demand as well (what non-2Nx2N PUs use)
    _run(200, 136, 8, 4, 2, 24, 12, use_maps=False)


@pytest.mark.parametrize("algo", [1, 2, 3])
@pytest.mark.parametrize("early", [0, 1, 2])
def test_heuristic_integer_search(algo, early):
    # hexagon_search / diamond_search / tz_search (search_inter.c:595-883) with the early_terminate variants
    _run(256, 192, 16, 4, 2, 64, 40 + 3 * algo + early, use_maps=False, ime_search=algo, me_early_termination=early)


def test_heuristic_integer_search_step_limit_and_restrictions():
    _run(256, 192, 16, 2, 3, 48, 50, use_maps=False, ime_search=1, me_max_steps=2)
    _run(256, 192, 16, 0, 2, 48, 51, use_maps=False, ime_search=2, me_max_steps=1)
    _run(256, 192, 16, 4, 2, 48, 52, use_maps=False, ime_search=1, owf=1, wpp=1)
    _run(200, 136, 8, 4, 2, 40, 53, use_maps=False, ime_search=2, mv_constraint=4, me_early_termination=2)
    _run(256, 192, 16, 4, 2, 48, 55, use_maps=False, ime_search=3, owf=1, wpp=1)
    _run(200, 136, 8, 2, 2, 40, 56, use_maps=False, ime_search=3, mv_constraint=4)
    # dense maps present: the heuristic searches ignore them
    _run(200, 136, 8, 4, 2, 24, 54, use_maps=True, ime_search=1)


def test_cached_window_centres():
    # windows round extra_mv / merge candidates served from centred dense maps, the rest on demand
    _run(256, 192, 16, 4, 2, 48, 14, centres=[(5, 3), (-7, 10)])
    _run(200, 136, 8, 2, 2, 40, 15, centres=[(-13, -6)])


def test_white_noise():
    rs = np.random.default_rng(99)
    w, h, n_refs, rng = 192, 128, 2, 16
    cur = rs.integers(0, 256, (h, w), dtype=np.uint8)
    refs = [rs.integers(0, 256, (h, w), dtype=np.uint8) for _ in range(n_refs)]
    pus, jobs = me_oracle.random_pus(rs, w, h, n_refs, 40, mv_span=rng)
    got = me_oracle.gpu_search(pus, jobs, cur, refs, w, h, rng, 4, 20)
    want = me_oracle.oracle_search(pus, jobs, cur, refs, w, h, rng, 4, 20)
    me_oracle.assert_results_equal(got, want, pus)


def test_1080p_full32_sampled():
    # BASELINE config 2 geometry; oracle on a sample of PUs
    _run(1920, 1080, 32, 4, 2, 24, 13)


def test_box_map_sets_union_of_windows():
    # one box per reference that covers the zero window and the windows round the cached centres: jobs
    # whose windows all lie inside it take the single-pass union scan (first covering window owns a
    # displacement), the others the per-window path
    _run(256, 192, 16, 4, 2, 64, 21, centres=[(5, 3), (-7, 10)], box=(-16 - 7, -16, 16 + 5 + 16 + 7 + 1, 16 + 10 + 16 + 1))
    _run(256, 192, 8, 4, 3, 64, 22, centres=[(3, -2)], box=(-8, -10, 8 + 3 + 8 + 1, 8 + 10 + 1))


def test_box_map_sets_with_restrictions_and_ties():
    # flat content makes equal costs frequent: the (cost, window, position) order has to hold in the union scan
    import kvzlibs as kl
    width, height, rng, n_refs = 200, 136, 8, 2
    rs = np.random.default_rng(5)
    f = kl.synth_luma(23, width, height, n_refs + 1)
    f = [(x // 64) * 64 for x in f]                       # 4 grey levels: many ties
    cur, refs = f[n_refs], [f[n_refs - 1 - i] for i in range(n_refs)]
    pus, jobs = me_oracle.random_pus(rs, width, height, n_refs, 48, mv_span=rng)
    for kw in ({}, {"owf": 1, "wpp": 1}, {"mv_constraint": 4}):
        got = me_oracle.gpu_search(pus, jobs, cur, refs, width, height, rng, 4, 20, box=(-16, -16, 48, 33), **kw)
        want = me_oracle.oracle_search(pus, jobs, cur, refs, width, height, rng, 4, 20, **kw)
        me_oracle.assert_results_equal(got, want, pus)


def test_pipelined_host_steps_equal_serial():
    """MePipeline.step_from_host_pipelined (copies of neighbouring steps overlapped with the kernels, two buffer
    sets) returns the results of MePipeline.step_from_host for every step, also when the inputs change per step."""
    import torch
    from kvazaar_b200.pipeline import MePipeline
    w, h, n_refs = 416, 240, 2
    f = kvzlibs.synth_luma(31, w, h, 6)
    pipe = MePipeline(w, h, 16, n_refs)
    steps = [(torch.from_numpy(f[k + 2]).pin_memory(), [torch.from_numpy(f[k + 1 - i]).pin_memory() for i in range(n_refs)])
             for k in range(4)]
    want = []
    for cur, refs in steps:
        pipe.step_from_host(cur, refs)
        want.append(pipe.host_results.clone())
    got = []
    for cur, refs in steps:
        r = pipe.step_from_host_pipelined(cur, refs)
        if r is not None:
            got.append(r.clone())
    got.append(pipe.flush_pipelined().clone())
    assert len(got) == len(want)
    for k, (a, b) in enumerate(zip(got, want)):
        assert torch.equal(a, b), k
    assert not torch.equal(want[0], want[1])
