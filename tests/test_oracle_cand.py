"""Oracle pinning for the inter candidate derivation: kvzo_inter_merge_cand /
kvzo_inter_mv_cand against the compiled reference's kvz_inter_get_merge_cand (inter.c:1314-1446) and
kvz_inter_get_mv_cand (inter.c:1209-1228) run on a working lcu_t cut from the same SCU arrays."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
import cand_oracle as co

needs_ref = pytest.mark.skipif(not kvzlibs.have_ref(), reason="compiled reference not present")


def oracle_merge(f, pu):
    O = kvzlibs.oracle()
    out = np.zeros(5, co.MERGE_DT)
    n = O.kvzo_inter_merge_cand(C.byref(f), *pu, out.ctypes.data_as(C.c_void_p))
    return n, out


def oracle_mv_cand(f, pu, reflist, lx_idx):
    O = kvzlibs.oracle()
    out = np.zeros((2, 2), np.int16)
    O.kvzo_inter_mv_cand(C.byref(f), *pu[:4], reflist, lx_idx, out.ctypes.data_as(C.c_void_p))
    return out


SCENES = [dict(size=(200, 136)), dict(size=(200, 136), slice_b=True, n_refs=4), dict(size=(136, 72), tmvp=False),
          dict(size=(264, 200), slice_b=True, n_refs=2, p_intra=0.05), dict(size=(192, 128), n_refs=1, poc=1),
          dict(size=(192, 128), slice_b=True, n_refs=3, p_intra=0.6)]


@needs_ref
@pytest.mark.parametrize("scene", SCENES, ids=[str(i) for i in range(len(SCENES))])
def test_candidates_match_reference(scene):
    R = kvzlibs.ref(1)
    kw = {k: v for k, v in scene.items() if k != "size"}
    rs = np.random.default_rng(scene["size"][0] + len(kw))
    sc = co.Scene(rs, scene["size"][0], scene["size"][1], **kw)
    f = sc.frame()
    distinct = set()
    for pu in sc.pus(rs, 150):
        n, got = oracle_merge(f, pu)
        want = np.zeros(5, co.MERGE_DT)
        wn = R.ref_shim_merge_cand(C.byref(f), *pu, want.ctypes.data_as(C.c_void_p))
        assert n == wn == 5
        for name in ("dir", "ref", "mv"):
            assert np.array_equal(got[name], want[name]), (pu, name, got, want)
        distinct.add(got.tobytes())
        for l in range(2):
            for idx in range(len(sc.lists[l])):
                a = oracle_mv_cand(f, pu, l, idx)
                b = np.zeros((2, 2), np.int16)
                R.ref_shim_mv_cand(C.byref(f), *pu[:4], l, idx, b.ctypes.data_as(C.c_void_p))
                assert np.array_equal(a, b), (pu, l, idx, a, b)
    assert len(distinct) > 25
