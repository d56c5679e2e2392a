"""ctypes mirror of include/kvz_cuda_enc.h -- the encoder-session face of libkvz_cuda.so (host pictures in, one
search_pu_inter reference loop per call out).  The reference encoder reaches this face through kvazaar_b200/host/;
this module exists for tests and tools that drive it without an encoder.  Nothing here computes."""
import ctypes as C

import numpy as np

from . import capi
from .frame import ME_JOB_DT, ME_PU_DT, ME_RESULT_DT

MAX_REFS = 16


class EncCfg(C.Structure):
    """kvz_cuda_enc_cfg (include/kvz_cuda_enc.h)"""
    _fields_ = [(n, C.c_int32) for n in (
        "device", "width", "height", "bitdepth", "ime_search", "range", "me_early_termination", "me_max_steps",
        "fme_level", "owf", "wpp", "sao_type", "deblock_enable", "mv_constraint", "max_ref_lcu_right",
        "max_ref_lcu_down", "max_map_mbytes")]


class EncRef(C.Structure):
    """kvz_cuda_enc_ref"""
    _fields_ = [("luma", C.c_void_p), ("stride", C.c_int32), ("poc", C.c_int32), ("key", C.c_uint64),
                ("mv_hints", C.c_void_p), ("n_mv_hints", C.c_int32), ("in_flight", C.c_int32)]


class EncFrame(C.Structure):
    """kvz_cuda_enc_frame"""
    _fields_ = [("frame_id", C.c_uint64), ("src_luma", C.c_void_p), ("src_stride", C.c_int32), ("src_key", C.c_uint64),
                ("src_pts", C.c_int64), ("tile_x", C.c_int32), ("tile_y", C.c_int32), ("tile_w", C.c_int32),
                ("tile_h", C.c_int32), ("ref_LX", (C.c_uint8 * 16) * 2), ("n_refs", C.c_int32), ("refs", EncRef * MAX_REFS)]


def _err():
    msg = capi.lib().kvz_cuda_last_error()
    return msg.decode() if msg else "unknown error"


def _lib():
    L = capi.lib()
    if not getattr(L, "_enc_bound", False):
        L.kvz_cuda_enc_open.restype = C.c_void_p
        L.kvz_cuda_enc_open.argtypes = [C.POINTER(EncCfg)]
        L.kvz_cuda_enc_close.restype = None
        L.kvz_cuda_enc_close.argtypes = [C.c_void_p]
        L.kvz_cuda_enc_frame_begin.restype = C.c_int
        L.kvz_cuda_enc_frame_begin.argtypes = [C.c_void_p, C.POINTER(EncFrame)]
        L.kvz_cuda_enc_search_pu.restype = C.c_int
        L.kvz_cuda_enc_search_pu.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.kvz_cuda_enc_search_begin.restype = C.c_int
        L.kvz_cuda_enc_search_begin.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]
        L.kvz_cuda_enc_search_end.restype = C.c_int
        L.kvz_cuda_enc_search_end.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.kvz_cuda_enc_search_ready.restype = C.c_int
        L.kvz_cuda_enc_search_ready.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L._enc_bound = True
    return L


class Session:
    """kvz_cuda_enc_open .. kvz_cuda_enc_close; pictures are numpy uint8 planes in host memory."""

    def __init__(self, width, height, search_range=32, fme_level=4, ime_search=0, device=0, **cfg):
        self.L = _lib()
        c = EncCfg(device=device, width=width, height=height, bitdepth=8, ime_search=ime_search, range=search_range,
                   me_early_termination=cfg.pop("me_early_termination", 0), me_max_steps=cfg.pop("me_max_steps", -1),
                   fme_level=fme_level, owf=cfg.pop("owf", 0), wpp=cfg.pop("wpp", 0), sao_type=cfg.pop("sao_type", 3),
                   deblock_enable=cfg.pop("deblock_enable", 1), mv_constraint=cfg.pop("mv_constraint", 0),
                   max_ref_lcu_right=cfg.pop("max_ref_lcu_right", 1), max_ref_lcu_down=cfg.pop("max_ref_lcu_down", 1),
                   max_map_mbytes=cfg.pop("max_map_mbytes", 0))
        assert not cfg, cfg
        self.width, self.height = width, height
        self.h = self.L.kvz_cuda_enc_open(C.byref(c))
        if not self.h:
            raise RuntimeError("kvz_cuda_enc_open: " + _err())
        self._keep = []

    def close(self):
        if self.h:
            self.L.kvz_cuda_enc_close(self.h)
            self.h = None

    def frame_begin(self, frame_num, cur, refs, pocs=None, hints=None):
        """cur / refs: contiguous uint8 (height, width) planes; returns the frame handle"""
        f = EncFrame()
        f.frame_id = frame_num << 16
        f.src_luma, f.src_stride, f.src_key, f.src_pts = cur.ctypes.data, cur.strides[0], cur.ctypes.data, -1
        f.tile_x = f.tile_y = 0
        f.tile_w, f.tile_h = self.width, self.height
        f.n_refs = len(refs)
        for i, r in enumerate(refs):
            f.ref_LX[0][i] = i
            f.refs[i].luma, f.refs[i].stride = r.ctypes.data, r.strides[0]
            f.refs[i].poc = pocs[i] if pocs else -1 - i
            f.refs[i].key = r.ctypes.data
            if hints is not None:
                f.refs[i].mv_hints, f.refs[i].n_mv_hints = hints.ctypes.data, len(hints) // 2
        self._keep.append((cur, refs, hints))
        handle = self.L.kvz_cuda_enc_frame_begin(self.h, C.byref(f))
        if handle < 0:
            raise RuntimeError("kvz_cuda_enc_frame_begin: " + _err())
        return handle

    def search_pu(self, handle, lambda_sqrt_i, pu, jobs):
        """one blocking search; pu: ME_PU_DT record, jobs: its ME_JOB_DT records -> ME_RESULT_DT array"""
        pu = np.array([pu], ME_PU_DT)
        jobs = np.ascontiguousarray(jobs, ME_JOB_DT)
        res = np.zeros(len(jobs), ME_RESULT_DT)
        if not self.L.kvz_cuda_enc_search_pu(self.h, handle, lambda_sqrt_i, pu.ctypes.data, jobs.ctypes.data, res.ctypes.data):
            raise RuntimeError("kvz_cuda_enc_search_pu: " + _err())
        return res

    def search_begin(self, handle, lambda_sqrt_i, pus, jobs_per_pu):
        """1..8 PUs of one CTU in one launch; jobs_per_pu[i] = the job records of pus[i]; returns the ticket"""
        pus = np.ascontiguousarray(pus, ME_PU_DT)
        padded = np.zeros(len(pus) * MAX_REFS, ME_JOB_DT)
        for i, jb in enumerate(jobs_per_pu):
            padded[i * MAX_REFS:i * MAX_REFS + len(jb)] = jb
        ticket = C.c_int(-1)
        if not self.L.kvz_cuda_enc_search_begin(self.h, handle, lambda_sqrt_i, len(pus), pus.ctypes.data, padded.ctypes.data,
                                                C.byref(ticket)):
            raise RuntimeError("kvz_cuda_enc_search_begin: " + _err())
        return ticket.value

    def search_ready(self, ticket, pu_index):
        return self.L.kvz_cuda_enc_search_ready(self.h, ticket, pu_index)

    def search_end(self, ticket, pu_index, n_jobs):
        res = np.zeros(n_jobs, ME_RESULT_DT)
        if not self.L.kvz_cuda_enc_search_end(self.h, ticket, pu_index, res.ctypes.data):
            raise RuntimeError("kvz_cuda_enc_search_end: " + _err())
        return res
