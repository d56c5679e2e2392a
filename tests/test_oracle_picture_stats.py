"""Oracle pinning for the picture-level reductions: kvzo_array_checksum against the
compiled reference's kvz_array_checksum strategy (strategies-nal.h:42-47: generic and the AVX2-era `generic8`
variants selected by the strategy selector) and a hand-computed vector; kvzo_plane_sse against the reference's
kvz_pixels_calc_ssd strategy (the same sum on square blocks) and numpy."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
from kvzlibs import ptr


def oracle_checksum(a, bitdepth=8):
    O = kvzlibs.oracle()
    fn = O.kvzo_array_checksum_8 if bitdepth == 8 else O.kvzo_array_checksum_10
    fn.restype = C.c_uint32
    fn.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    assert a.strides[1] == a.itemsize
    return int(fn(a.ctypes.data, a.shape[0], a.shape[1], a.strides[0] // a.itemsize))


def oracle_sse(a, b, bitdepth=8):
    O = kvzlibs.oracle()
    fn = O.kvzo_plane_sse_8 if bitdepth == 8 else O.kvzo_plane_sse_10
    fn.restype = C.c_uint64
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    return int(fn(a.ctypes.data, b.ctypes.data, a.shape[1], a.shape[0], a.strides[0] // a.itemsize,
                  b.strides[0] // b.itemsize))


def numpy_checksum(a, bitdepth=8):
    h, w = a.shape
    y, x = np.mgrid[0:h, 0:w]
    mask = ((x & 0xff) ^ (y & 0xff) ^ (x >> 8) ^ (y >> 8)) & 0xff
    s = int((((a.astype(np.int64)) & 0xff) ^ mask).sum())
    if bitdepth > 8:
        s += int((((a.astype(np.int64) >> 8) & 0xff) ^ mask).sum())
    return s & 0xffffffff


def test_checksum_known_answer():
    # 2x2 plane: masks are [[0, 1], [1, 0]]
    a = np.array([[10, 20], [30, 40]], np.uint8)
    assert oracle_checksum(a) == (10 ^ 0) + (20 ^ 1) + (30 ^ 1) + (40 ^ 0)


@pytest.mark.parametrize("shape", [(1, 1), (7, 13), (64, 64), (136, 200), (300, 520), (1080, 1920)])
def test_checksum_matches_numpy(shape):
    rs = np.random.default_rng(shape[0] * 7 + shape[1])
    a = rs.integers(0, 256, shape, dtype=np.uint8)
    assert oracle_checksum(a) == numpy_checksum(a)
    b = rs.integers(0, 1024, shape, dtype=np.uint16)
    assert oracle_checksum(b, 10) == numpy_checksum(b, 10)


@pytest.mark.skipif(not kvzlibs.have_ref(), reason="compiled reference not present")
@pytest.mark.parametrize("cpuid", [0, 1])
@pytest.mark.parametrize("shape", [(1, 1), (7, 13), (64, 64), (136, 200), (270, 520), (540, 960), (1080, 1920)])
def test_checksum_matches_reference(shape, cpuid):
    rs = np.random.default_rng(shape[0] * 3 + shape[1])
    a = rs.integers(0, 256, shape, dtype=np.uint8)
    fn = kvzlibs.ref_strategy("kvz_array_checksum", None,
                              [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_ubyte), C.c_uint8], cpuid=cpuid)
    out = (C.c_ubyte * 16)()
    fn(a.ctypes.data, shape[0], shape[1], shape[1], out, 8)
    want = (out[0] << 24) | (out[1] << 16) | (out[2] << 8) | out[3]
    assert oracle_checksum(a) == want


@pytest.mark.parametrize("shape", [(1, 1), (5, 9), (136, 200), (1080, 1920)])
def test_sse_matches_numpy(shape):
    rs = np.random.default_rng(shape[0] + shape[1])
    a, b = rs.integers(0, 256, (2,) + shape, dtype=np.uint8)
    assert oracle_sse(a, b) == int(((a.astype(np.int64) - b) ** 2).sum())
    a, b = rs.integers(0, 1024, (2,) + shape, dtype=np.uint16)
    assert oracle_sse(a, b, 10) == int(((a.astype(np.int64) - b) ** 2).sum())
    # strided views
    big = rs.integers(0, 256, (shape[0], shape[1] + 24), dtype=np.uint8)
    v, w = big[:, 3:3 + shape[1]], big[:, 11:11 + shape[1]]
    assert oracle_sse(v, w) == int(((v.astype(np.int64) - w) ** 2).sum())


@pytest.mark.skipif(not kvzlibs.have_ref(), reason="compiled reference not present")
@pytest.mark.parametrize("n", [8, 16, 32, 64])
def test_sse_matches_reference_ssd(n):
    rs = np.random.default_rng(n)
    a, b = rs.integers(0, 256, (2, n, n), dtype=np.uint8)
    fn = kvzlibs.ref_strategy("kvz_pixels_calc_ssd", C.c_uint, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int])
    assert oracle_sse(a, b) == int(fn(a.ctypes.data, b.ctypes.data, n, n, n))
