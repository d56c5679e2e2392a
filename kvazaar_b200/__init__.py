"""kvazaar_b200 -- B200-native inner encode loop for the Kvazaar HEVC encoder.

The product is ``libkvz_cuda.so`` (hand-written CUDA for sm_100a behind the C ABI
of ``include/kvz_cuda.h``).  This package is only the thin Python host layer the
tests and the benchmark use: ctypes bindings over that C ABI, with torch used for
device memory and streams.  There is no CPU fallback: importing works anywhere,
but every compute call raises :class:`KvzCudaError` without a B200.
"""
from .capi import (KvzCudaError, lib, lib_path, init, launch_count, ImeDesc,  # noqa: F401
                   ime_required_pad, ime_map_pitch)
from .capi import measure_int_pipe_peak  # noqa: F401
from .frame import PaddedPlane, ImeSearch, ImeBoxSearch  # noqa: F401
from .pipeline import MePipeline  # noqa: F401

__all__ = ["KvzCudaError", "lib", "lib_path", "init", "launch_count", "ImeDesc",
           "ime_required_pad", "ime_map_pitch", "PaddedPlane", "ImeSearch", "ImeBoxSearch", "MePipeline", "measure_int_pipe_peak"]
