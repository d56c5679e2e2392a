"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU
and exports every symbol include/kvz_cuda.h declares; compute calls fail loudly
(no CPU fallback) when there is no device."""
import os
import re

import pytest

import kvazaar_b200 as kb
from kvazaar_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "kvz_cuda.h")).read()
    return sorted(set(re.findall(r"KVZ_CUDA_API[^;(]*?\b(kvz_cuda_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = kb.lib()
    names = _declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/kvz_cuda.h but not exported"
        assert n in capi.SIGNATURES, f"{n} has no ctypes signature in kvazaar_b200/capi.py"


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    L = kb.lib()
    assert L.kvz_cuda_init(0) == 0
    assert b"no CUDA device" in L.kvz_cuda_last_error()
    with pytest.raises(kb.KvzCudaError):
        capi.init(0)


def test_map_geometry_helpers():
    assert kb.ime_map_pitch(32) == 80 and kb.ime_map_pitch(64) == 144 and kb.ime_map_pitch(8) == 32
    assert kb.ime_required_pad(32) == 128 and kb.ime_required_pad(64) == 160
    d = capi.ImeDesc(width=1920, height=1080, bitdepth=8, range=32, tile_x=0, tile_y=0,
                     tile_w=1920, tile_h=1080)
    import ctypes as C
    assert kb.lib().kvz_cuda_ime_map_bytes(C.byref(d), 0) == 510 * 64 * 66 * 80 * 2   # K + 1 rows per PU: row K is the column dx = +R
    assert kb.lib().kvz_cuda_ime_map_bytes(C.byref(d), 3) == 510 * 66 * 80 * 4


def _declared_strategy_symbols():
    text = open(os.path.join(ROOT, "include", "kvz_cuda_strategies.h")).read()
    return sorted(set(re.findall(r"KVZ_CUDA_API[^;(]*?\b(kvz_(?:cuda|strategy_register)_\w+)\s*\(", text)))


def test_per_call_face_exports_every_strategy_symbol():
    L = kb.lib()
    names = _declared_strategy_symbols()
    assert len(names) >= 50
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/kvz_cuda_strategies.h but not exported"
    # every strategy type of the hot path has a per-call function (STRATEGIES_*_EXPORTS of the reference)
    for t in ["reg_sad", "satd_any_size", "satd_any_size_quad", "pixels_calc_ssd", "coeff_abs_sum", "angular_pred",
              "intra_pred_planar", "sample_quarterpel_luma", "sample_octpel_chroma", "filter_hpel_blocks_hor_ver_luma",
              "filter_qpel_blocks_diag_luma", "dct_32x32", "idct_4x4", "fast_forward_dst_4x4"] + \
             [f"{k}_{n}x{n}{d}" for k in ("sad", "satd") for n in (4, 8, 16, 32, 64) for d in ("", "_dual")]:
        assert hasattr(L, "kvz_cuda_" + t), t


def test_registration_contract_without_device(ref_lib):
    """kvz_strategy_register_*_cuda(opaque, bitdepth, kvz_strategyselector_register): at another bit depth
    nothing is registered and 1 is returned (like avx2/picture-avx2.c:1230); without a device 0 is
    returned so that kvz_strategyselector_init fails encoder_open (strategyselector.c:54-95)."""
    import ctypes as C
    import torch
    L = kb.lib()

    class StrategyList(C.Structure):      # strategy_list_t (strategyselector.h:47-53)
        _fields_ = [("count", C.c_uint), ("allocated", C.c_uint), ("strategies", C.c_void_p)]

    reg = C.cast(ref_lib.kvz_strategyselector_register, C.c_void_p)
    for fam in ("picture", "dct", "quant", "intra", "ipol"):
        fn = getattr(L, f"kvz_strategy_register_{fam}_cuda_with")
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_uint8, C.c_void_p]
        lst = StrategyList()
        assert fn(C.byref(lst), 10, reg) == 1 and lst.count == 0
        if not torch.cuda.is_available():
            assert fn(C.byref(lst), 8, reg) == 0 and lst.count == 0
            assert b"no CUDA device" in L.kvz_cuda_last_error()
