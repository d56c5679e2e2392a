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
    assert kb.lib().kvz_cuda_ime_map_bytes(C.byref(d), 0) == 510 * 64 * 65 * 80 * 2
    assert kb.lib().kvz_cuda_ime_map_bytes(C.byref(d), 3) == 510 * 65 * 80 * 4
