"""GPU parity: kvz_cuda_me_search_fused -- the one-CTA-per-PU, one-launch form of the motion decision that
kvz_cuda_enc_search_pu (include/kvz_cuda_enc.h) calls while a host walks its CTUs -- against the oracle, through the
C ABI.  Covers its three integer-search paths: all windows inside a box map set (the warp scan), a window inside a
box but the job not (map reads by the CTA), a window outside every box (evaluated from the copy of the reference
region in shared memory, including windows far outside the picture).  Bit-exact."""
import numpy as np
import pytest

import kvzlibs
import me_oracle

pytestmark = pytest.mark.gpu


def _run(width, height, rng, fme, n_refs, count, seed, box=None, use_maps=True, sizes=(8, 16, 32, 64), far_prob=0.05,
         mv_span=None, pad=None, **cfg):
    rs = np.random.default_rng(seed)
    f = kvzlibs.synth_luma(seed, width, height, n_refs + 1)
    cur, refs = f[n_refs], [f[n_refs - 1 - i] for i in range(n_refs)]
    pus, jobs = me_oracle.random_pus(rs, width, height, n_refs, count, sizes=sizes, mv_span=mv_span or rng, far_prob=far_prob)
    lam = int(rs.integers(3, 45))
    got = me_oracle.gpu_search(pus, jobs, cur, refs, width, height, rng, fme, lam, use_maps=use_maps, box=box, fused=True, pad=pad,
                               **cfg)
    want = me_oracle.oracle_search(pus, jobs, cur, refs, width, height, rng, fme, lam, **cfg)
    me_oracle.assert_results_equal(got, want, pus)
    return got


@pytest.mark.parametrize("fme", [0, 2, 4])
def test_no_maps_every_window_from_shared_memory(fme):
    _run(200, 136, 8, fme, 2, 60, 200 + fme, use_maps=False)


def test_range16_three_refs_no_maps():
    _run(256, 192, 16, 4, 3, 48, 207, use_maps=False)


def test_box_holds_some_windows():
    # small box: most jobs have at least one window outside (CTA path: map reads + shared-memory windows mixed)
    _run(256, 192, 16, 4, 2, 64, 221, box=(-16, -16, 40, 33))
    # wide box: most jobs entirely inside (warp scan), a few outside
    _run(256, 192, 16, 4, 2, 64, 222, box=(-40, -40, 88, 81), mv_span=20)
    _run(256, 192, 8, 3, 3, 64, 223, box=(-24, -20, 56, 45), mv_span=12)


def test_windows_far_outside_the_picture():
    # merge candidates up to 300 samples away: the copied region is clamped per sample like image.c:320-452
    _run(200, 136, 8, 4, 2, 48, 230, use_maps=False, far_prob=0.5)
    _run(200, 136, 16, 2, 2, 48, 231, box=(-16, -16, 40, 33), far_prob=0.5)


def test_restrictions_and_ties():
    width, height, rng, n_refs = 200, 136, 8, 2
    rs = np.random.default_rng(6)
    f = kvzlibs.synth_luma(24, width, height, n_refs + 1)
    f = [(x // 64) * 64 for x in f]                       # 4 grey levels: many equal costs
    cur, refs = f[n_refs], [f[n_refs - 1 - i] for i in range(n_refs)]
    pus, jobs = me_oracle.random_pus(rs, width, height, n_refs, 48, mv_span=rng)
    for kw in ({}, {"owf": 1, "wpp": 1}, {"mv_constraint": 4}):
        for box in (None, (-16, -16, 48, 33)):
            got = me_oracle.gpu_search(pus, jobs, cur, refs, width, height, rng, 4, 20, box=box, use_maps=box is not None,
                                       fused=True, **kw)
            want = me_oracle.oracle_search(pus, jobs, cur, refs, width, height, rng, 4, 20, **kw)
            me_oracle.assert_results_equal(got, want, pus)


@pytest.mark.parametrize("algo", [1, 2, 3])
def test_heuristic_searches(algo):
    _run(256, 192, 16, 4, 2, 64, 240 + algo, use_maps=False, ime_search=algo, me_early_termination=algo % 3)


def test_1080p_full32_sampled():
    _run(1920, 1080, 32, 4, 2, 24, 213, box=(-48, -40, 104, 81), mv_span=40)
    _run(1920, 1080, 32, 4, 1, 12, 214, use_maps=False, sizes=(32, 64), mv_span=40)


def test_full64_window_region_fits_shared_memory():
    _run(512, 320, 64, 2, 1, 12, 215, use_maps=False, sizes=(16, 64), mv_span=64)


def test_equals_the_three_launch_form_on_a_batch():
    """the same PUs through kvz_cuda_me_search and kvz_cuda_me_search_fused"""
    width, height, rng, n_refs = 416, 240, 16, 3
    rs = np.random.default_rng(77)
    f = kvzlibs.synth_luma(77, width, height, n_refs + 1)
    cur, refs = f[n_refs], [f[n_refs - 1 - i] for i in range(n_refs)]
    pus, jobs = me_oracle.random_pus(rs, width, height, n_refs, 200, mv_span=24)
    a = me_oracle.gpu_search(pus, jobs, cur, refs, width, height, rng, 4, 27, box=(-32, -24, 72, 57))
    b = me_oracle.gpu_search(pus, jobs, cur, refs, width, height, rng, 4, 27, box=(-32, -24, 72, 57), fused=True)
    me_oracle.assert_results_equal(b, a, pus)


@pytest.mark.parametrize("fused", [False, True])
def test_minimum_border_with_vectors_far_outside(fused):
    """kvz_cuda_me_search accepts reference borders from 84 samples (64 + the 20 the refinement's window adds): with
    64-wide PUs and merge candidates hundreds of samples outside the picture the staged windows must still equal the
    per-sample clamp of kvz_get_extended_block"""
    width, height, rng, n_refs = 256, 192, 8, 2
    rs = np.random.default_rng(91)
    f = kvzlibs.synth_luma(91, width, height, n_refs + 1)
    cur, refs = f[n_refs], [f[n_refs - 1 - i] for i in range(n_refs)]
    pus, jobs = me_oracle.random_pus(rs, width, height, n_refs, 40, sizes=(32, 64), mv_span=rng, far_prob=0.6)
    got = me_oracle.gpu_search(pus, jobs, cur, refs, width, height, rng, 4, 23, use_maps=False, fused=fused, pad=84)
    want = me_oracle.oracle_search(pus, jobs, cur, refs, width, height, rng, 4, 23)
    me_oracle.assert_results_equal(got, want, pus)


def test_border_below_84_is_refused():
    from kvazaar_b200 import capi
    with pytest.raises(capi.KvzCudaError, match="border must be at least 84"):
        _run(200, 136, 8, 2, 1, 4, 92, use_maps=False, pad=80)
