"""ctypes bindings used by the tests (and bench/smoke) for the CHECKERS only:

* ``oracle()``  -- oracle/libkvz_oracle.so, our CPU restatement (test infrastructure)
* ``ref()``     -- oracle/_ref/libkvazaar_ref.so, the unmodified reference compiled
                   from /root/reference (present in this container and, because
                   oracle/_ref/ travels with gpurun snapshots, on the GPU box)

Nothing in kvazaar_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libkvz_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libkvazaar_ref.so")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "kvazaar")


class MergeCand(C.Structure):
    _fields_ = [("dir", C.c_uint8), ("ref", C.c_uint8 * 2), ("mv", (C.c_int16 * 2) * 2)]


class OracleMeJob(C.Structure):
    """kvzo_me_job (oracle/kvz_oracle.h)"""
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("origin_x", C.c_int32), ("origin_y", C.c_int32), ("pu_w", C.c_int32), ("pu_h", C.c_int32),
        ("tile_off_x", C.c_int32), ("tile_off_y", C.c_int32),
        ("mv_cand", (C.c_int16 * 2) * 2),
        ("num_merge_cand", C.c_int32),
        ("merge_cand", MergeCand * 5),
        ("ref_LX", (C.c_uint8 * 16) * 2),
        ("ref_idx", C.c_int32),
        ("extra_mv", C.c_int32 * 2),
        ("lambda_sqrt_i", C.c_int32),
        ("owf", C.c_int32), ("wpp", C.c_int32), ("sao_type", C.c_int32),
        ("deblock_enable", C.c_int32), ("mv_constraint", C.c_int32),
        ("max_ref_lcu_right", C.c_int32), ("max_ref_lcu_down", C.c_int32),
        ("fme_level", C.c_int32), ("me_early_termination", C.c_int32), ("me_max_steps", C.c_int32),
        ("ime_algorithm", C.c_int32),
        ("best_mv", C.c_int32 * 2),
        ("best_cost", C.c_uint32), ("best_bitcost", C.c_uint32),
    ]


class RefMeJob(C.Structure):
    """ref_me_job (oracle/ref_shim.c)"""
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("pic_y", C.c_void_p), ("ref_y", C.c_void_p),
        ("origin_x", C.c_int32), ("origin_y", C.c_int32), ("pu_w", C.c_int32), ("pu_h", C.c_int32),
        ("tile_off_x", C.c_int32), ("tile_off_y", C.c_int32),
        ("mv_cand", (C.c_int16 * 2) * 2),
        ("num_merge_cand", C.c_int32),
        ("merge_dir", C.c_uint8 * 5),
        ("merge_ref", (C.c_uint8 * 2) * 5),
        ("merge_mv", ((C.c_int16 * 2) * 2) * 5),
        ("ref_LX", (C.c_uint8 * 16) * 2),
        ("ref_idx", C.c_int32),
        ("extra_mv", C.c_int32 * 2),
        ("lambda_sqrt", C.c_double),
        ("owf", C.c_int32), ("wpp", C.c_int32), ("sao_type", C.c_int32),
        ("deblock_enable", C.c_int32), ("mv_constraint", C.c_int32),
        ("max_ref_lcu_right", C.c_int32), ("max_ref_lcu_down", C.c_int32),
        ("fme_level", C.c_int32), ("me_early_termination", C.c_int32), ("me_max_steps", C.c_int32),
        ("ime_algorithm", C.c_int32),
        ("best_mv", C.c_int32 * 2),
        ("best_cost", C.c_uint32), ("best_bitcost", C.c_uint32),
    ]


class OracleQCfg(C.Structure):
    _fields_ = [("qp", C.c_int32), ("bitdepth", C.c_int32), ("slice_is_intra", C.c_int32),
                ("signhide", C.c_int32), ("cu_is_intra", C.c_int32)]


class RefQCfg(C.Structure):
    _fields_ = [("qp", C.c_int32), ("slice_is_intra", C.c_int32), ("signhide", C.c_int32),
                ("rdoq", C.c_int32), ("rdoq_skip", C.c_int32), ("cu_is_intra", C.c_int32),
                ("lambda_", C.c_double)]


class RdoqCtx(C.Structure):
    """kvzo_rdoq_ctx == ref_rdoq_ctx == kvz_cuda_rdoq_ctx: the CABAC context states kvz_rdoq reads"""
    _fields_ = [("cbf_luma", C.c_uint8 * 4), ("cbf_chroma", C.c_uint8 * 4), ("sig_group", C.c_uint8 * 4),
                ("sig_luma", C.c_uint8 * 27), ("sig_chroma", C.c_uint8 * 15),
                ("last_y_luma", C.c_uint8 * 15), ("last_y_chroma", C.c_uint8 * 15),
                ("last_x_luma", C.c_uint8 * 15), ("last_x_chroma", C.c_uint8 * 15),
                ("one_luma", C.c_uint8 * 16), ("one_chroma", C.c_uint8 * 8),
                ("abs_luma", C.c_uint8 * 4), ("abs_chroma", C.c_uint8 * 2),
                ("root_cbf", C.c_uint8), ("trskip_luma", C.c_uint8), ("trskip_chroma", C.c_uint8), ("reserved", C.c_uint8)]


assert C.sizeof(RdoqCtx) == 148


def random_rdoq_ctx(rs):
    """context states drawn uniformly from the 128 (state, MPS) pairs"""
    raw = rs.integers(0, 128, 148, dtype=np.uint8)
    raw[147:] = 0
    return RdoqCtx.from_buffer_copy(raw.tobytes())


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


_oracle = None
_ref = None
_ref_cpuid = None


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"])


def oracle():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        L = C.CDLL(ORACLE_SO)
        for n in ("kvzo_reg_sad", "kvzo_image_sad", "kvzo_sad_nxn", "kvzo_satd_nxn",
                  "kvzo_satd_any_size", "kvzo_image_satd", "kvzo_ssd"):
            for s in ("_8", "_10"):
                getattr(L, n + s).restype = C.c_uint
        L.kvzo_mvd_cost.restype = C.c_uint32
        L.kvzo_coeff_abs_sum.restype = C.c_uint32
        _oracle = L
    return _oracle


def have_ref():
    return os.path.exists(REF_SO)


def ref(cpuid=1):
    """Load the compiled reference; cpuid=1 selects AVX2 strategies, 0 generic."""
    global _ref, _ref_cpuid
    if _ref is None or _ref_cpuid != cpuid:
        L = _ref if _ref is not None else C.CDLL(REF_SO)
        # strategyselector prints its report to stderr; silence it
        devnull = os.open(os.devnull, os.O_WRONLY)
        saved = os.dup(2)
        os.dup2(devnull, 2)
        try:
            ok = L.ref_shim_init(cpuid)
        finally:
            os.dup2(saved, 2)
            os.close(devnull)
            os.close(saved)
        assert ok, "kvz_strategyselector_init failed"
        for n in ("ref_shim_image_calc_sad", "ref_shim_image_calc_satd"):
            getattr(L, n).restype = C.c_uint
        L.ref_shim_mvd_cost.restype = C.c_uint32
        _ref = L
        _ref_cpuid = cpuid
    return _ref


def ref_strategy(name, restype, argtypes, cpuid=1):
    """Fetch one of the reference's global strategy function pointers
    (e.g. 'kvz_reg_sad', strategies-picture.c:32-63) as a callable."""
    L = ref(cpuid)
    addr = C.c_void_p.in_dll(L, name).value
    return C.CFUNCTYPE(restype, *argtypes)(addr)


def job_pair(**kw):
    """Build matching (OracleMeJob, RefMeJob) from keyword settings."""
    oj, rj = OracleMeJob(), RefMeJob()
    lam = kw.pop("lambda_sqrt", 20.0)
    merge = kw.pop("merge", [])            # list of (dir, ref0, ref1, mv00, mv01, mv10, mv11)
    mv_cand = kw.pop("mv_cand", [[0, 0], [0, 0]])
    ref_LX = kw.pop("ref_LX", None)
    extra_mv = kw.pop("extra_mv", (0, 0))
    best_mv = kw.pop("best_mv", (0, 0))
    for j in (oj, rj):
        for k, v in kw.items():
            setattr(j, k, v)
        for a in range(2):
            for b in range(2):
                j.mv_cand[a][b] = mv_cand[a][b]
        j.num_merge_cand = len(merge)
        j.extra_mv[0], j.extra_mv[1] = extra_mv
        j.best_mv[0], j.best_mv[1] = best_mv
        for l in range(2):
            for i in range(16):
                j.ref_LX[l][i] = ref_LX[l][i] if ref_LX is not None else i
    oj.lambda_sqrt_i = int(lam + 0.5)
    rj.lambda_sqrt = lam
    for i, (d, r0, r1, a, b, c, e) in enumerate(merge):
        oj.merge_cand[i].dir = d
        oj.merge_cand[i].ref[0], oj.merge_cand[i].ref[1] = r0, r1
        oj.merge_cand[i].mv[0][0], oj.merge_cand[i].mv[0][1] = a, b
        oj.merge_cand[i].mv[1][0], oj.merge_cand[i].mv[1][1] = c, e
        rj.merge_dir[i] = d
        rj.merge_ref[i][0], rj.merge_ref[i][1] = r0, r1
        rj.merge_mv[i][0][0], rj.merge_mv[i][0][1] = a, b
        rj.merge_mv[i][1][0], rj.merge_mv[i][1][1] = c, e
    return oj, rj


def synth_luma(seed, w, h, frames, bitdepth=8, motion=(5, 3)):
    """Seeded synthetic luma clip of SURVEY.md 8(d): band-limited texture,
    global translation of `motion` px/frame, +-4 noise."""
    rng = np.random.default_rng(seed)
    mx, my = motion
    pad_w = w + abs(mx) * frames + 16
    pad_h = h + abs(my) * frames + 16
    base = rng.integers(0, 1 << bitdepth, size=(pad_h // 8 + 16, pad_w // 8 + 16)).astype(np.int64)
    tex = np.kron(base, np.ones((8, 8), dtype=np.int64))
    for _ in range(2):
        p = np.pad(tex, 1, mode="edge")
        tex = (p[1:-1, 1:-1] + p[:-2, 1:-1] + p[2:, 1:-1] + p[1:-1, :-2] + p[1:-1, 2:]) // 5
    out = []
    dt = np.uint8 if bitdepth == 8 else np.uint16
    for f in range(frames):
        ox, oy = 8 + mx * f, 8 + my * f
        win = tex[oy:oy + h, ox:ox + w] + rng.integers(-4, 5, size=(h, w))
        out.append(np.clip(win, 0, (1 << bitdepth) - 1).astype(dt))
    return out


def aligned(shape, dtype, align=64):
    """numpy array whose data pointer is `align`-byte aligned (the reference's
    AVX2 strategies use aligned loads on coefficient buffers)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape))
    raw = np.zeros(n * dtype.itemsize + align, dtype=np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n * dtype.itemsize].view(dtype).reshape(shape)


def aligned_copy(a, align=64):
    out = aligned(a.shape, a.dtype, align)
    out[...] = a
    return out
