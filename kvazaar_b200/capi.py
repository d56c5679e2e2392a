"""ctypes bindings of include/kvz_cuda.h (one Python function per C entry point)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class KvzCudaError(RuntimeError):
    pass


def lib_path():
    return os.path.join(_HERE, "libkvz_cuda.so")


class ImeDesc(C.Structure):
    """kvz_cuda_ime_desc"""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("bitdepth", C.c_int32),
                ("range", C.c_int32),
                ("tile_x", C.c_int32), ("tile_y", C.c_int32), ("tile_w", C.c_int32), ("tile_h", C.c_int32),
                ("cur_pitch", C.c_int32), ("ref_pitch", C.c_int32), ("ref_pad", C.c_int32),
                ("center_x", C.c_int32), ("center_y", C.c_int32)]


MAX_REFS = 16


MAX_MAPSETS = 4
IME_MAX_MULTI = 8


class ImeBoxDesc(C.Structure):
    """kvz_cuda_ime_box_desc"""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("bitdepth", C.c_int32),
                ("tile_x", C.c_int32), ("tile_y", C.c_int32), ("tile_w", C.c_int32), ("tile_h", C.c_int32),
                ("cur_pitch", C.c_int32), ("ref_pitch", C.c_int32), ("ref_pad", C.c_int32),
                ("x0", C.c_int32), ("y0", C.c_int32), ("n_dx", C.c_int32), ("n_dy", C.c_int32)]


class Plane(C.Structure):
    """kvz_cuda_plane"""
    _fields_ = [("d_data", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32), ("pitch", C.c_int32)]


class CandFrame(C.Structure):
    """kvz_cuda_cand_frame"""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("in_width", C.c_int32), ("in_height", C.c_int32),
                ("tile_offset_x", C.c_int32), ("tile_offset_y", C.c_int32), ("poc", C.c_int32), ("slice_b", C.c_int32),
                ("tmvp_enable", C.c_int32), ("n_refs", C.c_int32), ("ref_pocs", C.c_int32 * 16),
                ("col_ref_pocs", (C.c_int32 * 16) * 2), ("ref_LX", (C.c_uint8 * 16) * 2), ("ref_LX_size", C.c_uint8 * 2),
                ("reserved", C.c_uint8 * 2), ("ref_scu_pitch", C.c_int32), ("d_scus", C.c_void_p),
                ("d_ref_scus", C.c_void_p * 16)]


class DeblockCfg(C.Structure):
    """kvz_cuda_deblock_cfg"""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("qp", C.c_int32), ("cu_qp", C.c_int32),
                ("beta_offset_div2", C.c_int32), ("tc_offset_div2", C.c_int32), ("slice_b", C.c_int32),
                ("reserved", C.c_int32), ("ref_LX", (C.c_uint8 * 16) * 2)]


class ImeBoxRef(C.Structure):
    """kvz_cuda_ime_box_ref"""
    _fields_ = [("desc", ImeBoxDesc), ("d_ref_padded", C.c_void_p), ("d_sad8", C.c_void_p), ("d_sad16", C.c_void_p),
                ("d_sad32", C.c_void_p), ("d_sad64", C.c_void_p)]


class MeMapSet(C.Structure):
    """kvz_cuda_me_mapset"""
    _fields_ = [("center_x", C.c_int32), ("center_y", C.c_int32), ("n_dx", C.c_int32), ("n_dy", C.c_int32),
                ("d_sad8", C.c_void_p), ("d_sad16", C.c_void_p), ("d_sad32", C.c_void_p), ("d_sad64", C.c_void_p)]


class MeRef(C.Structure):
    """kvz_cuda_me_ref"""
    _fields_ = [("d_ref_padded", C.c_void_p), ("ref_pitch", C.c_int32), ("ref_pad", C.c_int32),
                ("n_mapsets", C.c_int32), ("reserved", C.c_int32), ("mapsets", MeMapSet * MAX_MAPSETS)]


class MeFrame(C.Structure):
    """kvz_cuda_me_frame"""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("bitdepth", C.c_int32), ("range", C.c_int32),
                ("tile_x", C.c_int32), ("tile_y", C.c_int32), ("tile_w", C.c_int32), ("tile_h", C.c_int32),
                ("d_cur", C.c_void_p), ("cur_pitch", C.c_int32), ("lambda_sqrt_i", C.c_int32),
                ("fme_level", C.c_int32), ("owf", C.c_int32), ("wpp", C.c_int32), ("sao_type", C.c_int32),
                ("deblock_enable", C.c_int32), ("mv_constraint", C.c_int32),
                ("max_ref_lcu_right", C.c_int32), ("max_ref_lcu_down", C.c_int32),
                ("ref_LX", (C.c_uint8 * 16) * 2), ("n_refs", C.c_int32),
                ("refs", MeRef * MAX_REFS),
                ("ime_search", C.c_int32), ("me_early_termination", C.c_int32), ("me_max_steps", C.c_int32)]


class QuantCfg(C.Structure):
    """kvz_cuda_quant_cfg"""
    _fields_ = [("qp", C.c_int32), ("bitdepth", C.c_int32), ("slice_is_intra", C.c_int32),
                ("signhide", C.c_int32), ("cu_is_intra", C.c_int32)]


# name -> (restype, argtypes); every symbol include/kvz_cuda.h declares
_VP = C.c_void_p
SIGNATURES = {
    "kvz_cuda_init": (C.c_int, [C.c_int]),
    "kvz_cuda_shutdown": (None, []),
    "kvz_cuda_last_error": (C.c_char_p, []),
    "kvz_cuda_device_count": (C.c_int, []),
    "kvz_cuda_launch_count": (C.c_ulonglong, []),
    "kvz_cuda_measure_int_pipe_peak": (C.c_int, [C.POINTER(C.c_double)]),
    "kvz_cuda_malloc": (_VP, [C.c_size_t]),
    "kvz_cuda_free": (None, [_VP]),
    "kvz_cuda_malloc_host": (_VP, [C.c_size_t]),
    "kvz_cuda_free_host": (None, [_VP]),
    "kvz_cuda_memcpy_h2d": (C.c_int, [_VP, _VP, C.c_size_t, _VP]),
    "kvz_cuda_memcpy_d2h": (C.c_int, [_VP, _VP, C.c_size_t, _VP]),
    "kvz_cuda_memset": (C.c_int, [_VP, C.c_int, C.c_size_t, _VP]),
    "kvz_cuda_stream_sync": (C.c_int, [_VP]),
    "kvz_cuda_pad_plane": (C.c_int, [_VP, C.c_int, C.c_int, C.c_int, _VP, C.c_int, C.c_int, C.c_int, _VP]),
    "kvz_cuda_ime_required_pad": (C.c_int, [C.c_int]),
    "kvz_cuda_ime_map_pitch": (C.c_int, [C.c_int]),
    "kvz_cuda_ime_map_bytes": (C.c_size_t, [C.POINTER(ImeDesc), C.c_int]),
    "kvz_cuda_ime_sad_maps": (C.c_int, [C.POINTER(ImeDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "kvz_cuda_ime_box_map_bytes": (C.c_size_t, [C.POINTER(ImeBoxDesc), C.c_int]),
    "kvz_cuda_ime_sad_box_maps": (C.c_int, [C.POINTER(ImeBoxDesc), _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "kvz_cuda_ime_sad_box_maps_multi": (C.c_int, [C.POINTER(ImeBoxRef), C.c_int, _VP, _VP]),
    "kvz_cuda_block_costs": (C.c_int, [_VP, _VP, _VP, C.c_int, C.c_int, _VP, _VP]),
    "kvz_cuda_picture_checksum": (C.c_int, [_VP, C.c_int, C.c_int, _VP, _VP]),
    "kvz_cuda_picture_sse": (C.c_int, [_VP, _VP, C.c_int, C.c_int, _VP, _VP]),
    "kvz_cuda_sao_ctu_stats": (C.c_int, [_VP, _VP, C.c_int, C.c_int, _VP, _VP]),
    "kvz_cuda_sao_reconstruct": (C.c_int, [_VP, _VP, C.c_int, C.c_int, _VP, _VP, _VP]),
    "kvz_cuda_deblock_picture": (C.c_int, [_VP, _VP, _VP, C.c_int, C.c_int, _VP]),
    "kvz_cuda_coeff_bits": (C.c_int, [_VP, _VP, _VP, C.c_int, C.c_int, C.c_int, C.c_int, _VP, _VP]),
    "kvz_cuda_inter_candidates": (C.c_int, [_VP, _VP, C.c_int, _VP, _VP, _VP]),
    "kvz_cuda_image_costs": (C.c_int, [_VP, C.c_int, _VP, C.c_int, C.c_int, C.c_int, _VP, C.c_int, C.c_int, _VP, _VP]),
    "kvz_cuda_transform_blocks": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, _VP, _VP, C.c_int, _VP]),
    "kvz_cuda_quant_blocks": (C.c_int, [C.POINTER(QuantCfg), C.c_int, C.c_int, C.c_int, _VP, _VP, C.c_int, _VP]),
    "kvz_cuda_dequant_blocks": (C.c_int, [C.POINTER(QuantCfg), C.c_int, C.c_int, _VP, _VP, C.c_int, _VP]),
    "kvz_cuda_coeff_abs_sums": (C.c_int, [_VP, C.c_int, C.c_int, _VP, _VP]),
    "kvz_cuda_quantize_residuals": (C.c_int, [C.POINTER(QuantCfg), _VP, _VP, _VP, _VP, _VP, C.c_int, _VP, _VP]),
    "kvz_cuda_rdoq_blocks": (C.c_int, [C.POINTER(QuantCfg), _VP, _VP, _VP, _VP, C.c_int, C.c_int, _VP]),
    "kvz_cuda_quantize_residuals_rdoq": (C.c_int, [C.POINTER(QuantCfg), _VP, _VP, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP,
                                                   C.c_int, _VP, C.c_int, _VP]),
    "kvz_cuda_quantize_residuals_rdoq_noalloc": (C.c_int, [C.POINTER(QuantCfg), _VP, _VP, _VP, _VP, C.c_size_t, _VP, C.c_int,
                                                           _VP, _VP, _VP, C.c_int, C.c_int, _VP]),
    "kvz_cuda_interp_blocks": (C.c_int, [_VP, C.c_int, C.c_int, C.c_int, C.c_int, _VP, C.c_int, _VP, _VP, _VP]),
    "kvz_cuda_bipred_blend": (C.c_int, [_VP, _VP, C.c_size_t, C.c_int, _VP, _VP]),
    "kvz_cuda_intra_mode_costs": (C.c_int, [_VP, _VP, _VP, C.c_int, C.c_int, _VP, _VP, _VP]),
    "kvz_cuda_me_search": (C.c_int, [C.POINTER(MeFrame), _VP, C.c_int, _VP, _VP, _VP]),
    "kvz_cuda_me_search_fused": (C.c_int, [C.POINTER(MeFrame), _VP, C.c_int, _VP, _VP, _VP]),
}


def lib():
    """Load libkvz_cuda.so (built in-tree by __graft_entry__.build())."""
    global _LIB
    if _LIB is None:
        path = lib_path()
        if not os.path.exists(path):
            raise KvzCudaError(
                f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        L = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(ok, what):
    if not ok:
        msg = lib().kvz_cuda_last_error()
        raise KvzCudaError(f"{what} failed: {msg.decode() if msg else 'unknown error'}")


_inited = None


def init(device=0):
    """kvz_cuda_init(); raises if no B200 is present."""
    global _inited
    if _inited != device:
        check(lib().kvz_cuda_init(device), "kvz_cuda_init")
        _inited = device
    return device


def launch_count():
    return int(lib().kvz_cuda_launch_count())


def measure_int_pipe_peak():
    """pixel absolute differences / s of a register-only VABSDIFF4 loop on this device"""
    v = C.c_double(0.0)
    check(lib().kvz_cuda_measure_int_pipe_peak(C.byref(v)), "kvz_cuda_measure_int_pipe_peak")
    return v.value


def ime_required_pad(rng):
    return int(lib().kvz_cuda_ime_required_pad(rng))


def ime_map_pitch(rng):
    return int(lib().kvz_cuda_ime_map_pitch(rng))
