"""GPU parity: the encoder-session face (include/kvz_cuda_enc.h) driven directly -- host pictures in, one
search_pu_inter reference loop per call out -- against the oracle: the blocking kvz_cuda_enc_search_pu, and the split
form kvz_cuda_enc_search_begin / _end with several PUs of one CTU per launch (what kvazaar_b200/host/ uses to launch
PUs ahead of the CU walk), including tickets that expire.  Bit-exact."""
import numpy as np
import pytest

import kvzlibs
import me_oracle

pytestmark = pytest.mark.gpu

W, H, RNG, FME, NREFS = 192, 128, 16, 4, 2


def _setup(seed, count, **cfg):
    from kvazaar_b200.session import Session
    rs = np.random.default_rng(seed)
    f = kvzlibs.synth_luma(seed, W, H, NREFS + 1)
    cur, refs = np.ascontiguousarray(f[NREFS]), [np.ascontiguousarray(f[NREFS - 1 - i]) for i in range(NREFS)]
    pus, jobs = me_oracle.random_pus(rs, W, H, NREFS, count, mv_span=RNG, far_prob=0.03)
    lam = int(rs.integers(3, 45))
    want = me_oracle.oracle_search(pus, jobs, cur, refs, W, H, RNG, FME, lam, **cfg)
    s = Session(W, H, search_range=RNG, fme_level=FME, **cfg)
    hints = np.zeros(2 * 64, np.int16)
    handle = s.frame_begin(1, cur, refs, hints=hints)
    return s, handle, pus, jobs, lam, want


def test_blocking_search_pu_equals_oracle():
    s, handle, pus, jobs, lam, want = _setup(11, 40)
    try:
        got = np.concatenate([s.search_pu(handle, lam, pus[i], jobs[i * NREFS:(i + 1) * NREFS]) for i in range(len(pus))])
    finally:
        s.close()
    me_oracle.assert_results_equal(got, want, pus)


def test_begin_end_several_pus_per_launch_equals_oracle():
    s, handle, pus, jobs, lam, want = _setup(12, 120)
    try:
        by_ctu = {}
        for i, pu in enumerate(pus):
            by_ctu.setdefault((int(pu["x"]) >> 6, int(pu["y"]) >> 6), []).append(i)
        got = np.zeros_like(want)
        pending = []
        for idx in by_ctu.values():
            for k in range(0, len(idx), 5):                       # launches of 1..5 PUs, several in flight at once
                grp = idx[k:k + 5]
                t = s.search_begin(handle, lam, pus[grp], [jobs[i * NREFS:(i + 1) * NREFS] for i in grp])
                pending.append((t, grp))
                if len(pending) == 6:
                    for t0, g0 in reversed(pending):               # collected out of order
                        for j in reversed(range(len(g0))):
                            got[g0[j] * NREFS:(g0[j] + 1) * NREFS] = s.search_end(t0, j, NREFS)
                    pending = []
        for t0, g0 in pending:
            for j in range(len(g0)):
                got[g0[j] * NREFS:(g0[j] + 1) * NREFS] = s.search_end(t0, j, NREFS)
    finally:
        s.close()
    me_oracle.assert_results_equal(got, want, pus)


def test_tickets_expire_and_bad_launches_are_refused():
    s, handle, pus, jobs, lam, want = _setup(13, 40)
    try:
        first = s.search_begin(handle, lam, pus[:1], [jobs[:NREFS]])
        got0 = s.search_end(first, 0, NREFS)
        assert s.search_ready(first, 0) == 1 and s.search_ready(first, 1) == -1
        me_oracle.assert_results_equal(got0, want[:NREFS])
        for i in range(64):                                        # the thread's 64 slots go round once
            k = i % len(pus)
            s.search_end(s.search_begin(handle, lam, pus[k:k + 1], [jobs[k * NREFS:(k + 1) * NREFS]]), 0, NREFS)
        with pytest.raises(RuntimeError, match="expired"):
            s.search_end(first, 0, NREFS)
        assert s.search_ready(first, 0) == -1
        a = next(i for i in range(len(pus)) if (int(pus[i]["x"]) >> 6, int(pus[i]["y"]) >> 6) != (int(pus[0]["x"]) >> 6, int(pus[0]["y"]) >> 6))
        with pytest.raises(RuntimeError, match="one CTU"):
            s.search_begin(handle, lam, pus[[0, a]], [jobs[:NREFS], jobs[a * NREFS:(a + 1) * NREFS]])
        with pytest.raises(RuntimeError, match="1..8"):
            s.search_begin(handle, lam, pus[:9], [jobs[i * NREFS:(i + 1) * NREFS] for i in range(9)])
        with pytest.raises(RuntimeError, match="handle"):
            s.search_begin(handle + 7, lam, pus[:1], [jobs[:NREFS]])
    finally:
        s.close()
