"""Oracle pinning for the deblocking filter: kvzo_deblock_picture against the compiled
reference's kvz_filter_deblock_lcu (filter.c:770-779) run over every LCU of a fake frame."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
import deblock_oracle as do

needs_ref = pytest.mark.skipif(not kvzlibs.have_ref(), reason="compiled reference not present")


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def oracle_deblock(cfg, scus, planes, bitdepth=8):
    O = kvzlibs.oracle()
    fn = O.kvzo_deblock_picture_8 if bitdepth == 8 else O.kvzo_deblock_picture_10
    out = [np.ascontiguousarray(p).copy() for p in planes]
    s = np.ascontiguousarray(scus)
    fn(C.byref(cfg), _vp(s), _vp(out[0]), _vp(out[1]), _vp(out[2]))
    return out


def ref_deblock(cfg, scus, planes):
    R = kvzlibs.ref(1)
    out = [np.ascontiguousarray(p).copy() for p in planes]
    s = np.ascontiguousarray(scus)
    R.ref_shim_deblock_picture(C.byref(cfg), _vp(s), _vp(out[0]), _vp(out[1]), _vp(out[2]))
    return out


CASES = [
    dict(size=(64, 64)), dict(size=(136, 72)), dict(size=(200, 136), qp=37), dict(size=(256, 128), qp=22, blockiness=5),
    dict(size=(200, 136), slice_b=1), dict(size=(264, 200), slice_b=1, p_intra=0.05, qp=40),
    dict(size=(200, 136), cu_qp=1, qp_spread=6), dict(size=(200, 136), beta=3, tc=-2), dict(size=(200, 136), beta=-6, tc=6, qp=45),
    dict(size=(136, 136), chroma=0), dict(size=(192, 128), p_intra=1.0), dict(size=(192, 128), p_intra=0.0, qp=51),
]


@needs_ref
@pytest.mark.parametrize("case", CASES, ids=[str(i) for i in range(len(CASES))])
def test_picture_matches_reference(case):
    w, h = case["size"]
    rs = np.random.default_rng(w * 3 + h + case.get("qp", 0))
    cfg = do.make_cfg(w, h, qp=case.get("qp", 32), cu_qp=case.get("cu_qp", 0), beta=case.get("beta", 0), tc=case.get("tc", 0),
                      slice_b=case.get("slice_b", 0), chroma=case.get("chroma", 1))
    scus = do.random_scus(rs, w, h, slice_b=bool(case.get("slice_b", 0)), p_intra=case.get("p_intra", 0.3),
                          qp=case.get("qp", 32), qp_spread=case.get("qp_spread", 0))
    planes = do.picture(rs, w, h, blockiness=case.get("blockiness", 12))
    got, want = oracle_deblock(cfg, scus, planes), ref_deblock(cfg, scus, planes)
    for k in range(3):
        assert np.array_equal(got[k], want[k]), k
    assert not np.array_equal(got[0], planes[0])
    if not case.get("chroma", 1):
        assert np.array_equal(got[1], planes[1])
    elif case.get("p_intra", 0.3) > 0:
        assert not np.array_equal(got[1], planes[1])
