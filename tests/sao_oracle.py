"""Shared helpers of the SAO tests: ctypes views of the oracle's SAO functions and the picture-level
restatement (per CTU: the statistics sao_search_edge_sao / sao_search_band_sao gather, sao.c:337-455; the
reconstruction encoder_sao_reconstruct amounts to over a whole picture, encoderstate.c:245-441)."""
import ctypes as C

import numpy as np

import kvzlibs

SAO_PARAMS_DT = np.dtype([("type", "i1"), ("eo_class", "i1"), ("band_position", "i1", (2,)), ("offsets", "i1", (10,)),
                          ("reserved", "i1", (2,))])
SAO_STATS_DT = np.dtype([("edge", "<i4", (4, 2, 5)), ("band", "<i4", (2, 32))])
assert SAO_PARAMS_DT.itemsize == 16 and SAO_STATS_DT.itemsize == 416


class SaoInfo(C.Structure):
    _fields_ = [("type", C.c_int32), ("eo_class", C.c_int32), ("band_position", C.c_int32 * 2), ("offsets", C.c_int32 * 10)]


def info_of(p):
    s = SaoInfo()
    s.type, s.eo_class = int(p["type"]), int(p["eo_class"])
    for i in range(2):
        s.band_position[i] = int(p["band_position"][i])
    for i in range(10):
        s.offsets[i] = int(p["offsets"][i])
    return s


def sfx(bitdepth):
    return "8" if bitdepth == 8 else "10"


def _pp(a):
    return a.ctypes.data_as(C.c_void_p)


def ints(v):
    return (C.c_int * len(v))(*[int(x) for x in v])


def random_params(rs, n, bitdepth=8, p_none=0.2):
    """n random kvz_cuda_sao_params: offsets follow the signs the syntax can code for the edge type
    (sao.c:373-379), |offset| <= SAO_ABS_OFFSET_MAX = (1 << (min(bitdepth, 10) - 5)) - 1"""
    m = (1 << (min(bitdepth, 10) - 5)) - 1
    out = np.zeros(n, SAO_PARAMS_DT)
    out["type"] = rs.choice([0, 1, 2], n, p=[p_none, (1 - p_none) / 2, (1 - p_none) / 2])
    out["eo_class"] = rs.integers(0, 4, n)
    out["band_position"] = rs.integers(0, 29, (n, 2))
    off = rs.integers(-m, m + 1, (n, 10))
    edge = out["type"] == 2
    off[edge, 1:3] = np.abs(off[edge, 1:3]); off[edge, 6:8] = np.abs(off[edge, 6:8])
    off[edge, 3:5] = -np.abs(off[edge, 3:5]); off[edge, 8:10] = -np.abs(off[edge, 8:10])
    off[:, 0] = 0; off[:, 5] = 0
    out["offsets"] = off
    return out


def ctu_blocks(plane, ctu):
    """(index, y0, x0, h, w) of the CTUs of a plane, raster order"""
    h, w = plane.shape
    k = 0
    for y0 in range(0, h, ctu):
        for x0 in range(0, w, ctu):
            yield k, y0, x0, min(ctu, h - y0), min(ctu, w - x0)
            k += 1


def oracle_ctu_stats(orig_planes, rec_planes, bitdepth=8):
    """per plane an array of SAO_STATS_DT, one record per CTU (64 luma / 32 chroma samples wide)"""
    O = kvzlibs.oracle()
    edge = getattr(O, "kvzo_calc_sao_edge_dir_" + sfx(bitdepth))
    band = getattr(O, "kvzo_calc_sao_bands_" + sfx(bitdepth))
    out = []
    for k, (o, r) in enumerate(zip(orig_planes, rec_planes)):
        ctu = 64 if k == 0 else 32
        n = ((o.shape[0] + ctu - 1) // ctu) * ((o.shape[1] + ctu - 1) // ctu)
        st = np.zeros(n, SAO_STATS_DT)
        for i, y0, x0, h, w in ctu_blocks(o, ctu):
            ob = np.ascontiguousarray(o[y0:y0 + h, x0:x0 + w])
            rb = np.ascontiguousarray(r[y0:y0 + h, x0:x0 + w])
            for e in range(4):
                acc = (C.c_int * 10)()
                edge(_pp(ob), _pp(rb), e, w, h, acc)
                st[i]["edge"][e] = np.array(acc[:]).reshape(2, 5)
            acc = (C.c_int * 64)()
            band(_pp(ob), _pp(rb), w, h, acc)
            st[i]["band"] = np.array(acc[:]).reshape(2, 32)
        out.append(st)
    return out


def oracle_reconstruct(pre_planes, luma_params, chroma_params, bitdepth=8):
    """every CTU of every plane through kvzo_sao_reconstruct with the parameters of that CTU"""
    O = kvzlibs.oracle()
    fn = getattr(O, "kvzo_sao_reconstruct_" + sfx(bitdepth))
    outs = []
    for k, pre in enumerate(pre_planes):
        pre = np.ascontiguousarray(pre)
        out = pre.copy()
        ctu = 64 if k == 0 else 32
        params = luma_params if k == 0 else chroma_params
        for i, y0, x0, h, w in ctu_blocks(pre, ctu):
            s = info_of(params[i])
            fn(_pp(pre), _pp(out), pre.shape[1], pre.shape[0], x0, y0, w, h, C.byref(s), k)
        outs.append(out)
    return outs
