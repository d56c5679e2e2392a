"""Known-answer vectors of the reference's own unit tests, restated as data.

Sources (all under /root/reference/tests/):
  * sad_tests.c:37-57,126-258   17 closed-form kvz_image_calc_sad values (edge clamp)
  * sad_tests.c:85-103,373-387  reg_sad inputs (i*i/32+i)%255 vs (i*i/16+i)%255, 18 shapes
  * satd_tests.c:45-96,111-160  SATD goldens for N = 4..64
  * coeff_sum_tests.c:27-49     arithmetic-series coefficient sum
  * dct_tests.c:56-127          clipped radial-gradient blocks (generic == every strategy)
The CPU tests pin the oracle on them, the GPU tests pin the CUDA path on them.
"""
import numpy as np

# ---- sad_tests.c:37-57 ----
_REF8 = np.array([
    1, 2, 2, 2, 2, 2, 2, 3,
    4, 5, 5, 5, 5, 5, 5, 6,
    4, 5, 5, 5, 5, 5, 5, 6,
    4, 5, 5, 5, 5, 5, 5, 6,
    4, 5, 5, 5, 5, 5, 5, 6,
    4, 5, 5, 5, 5, 5, 5, 6,
    4, 5, 5, 5, 5, 5, 5, 6,
    7, 8, 8, 8, 8, 8, 8, 9], dtype=np.int64).reshape(8, 8)


def sad_edge_pictures(dtype=np.uint8):
    """(pic, ref) 8x8 pictures of sad_tests.c:78-84 (pattern + 48)."""
    pic = np.full((8, 8), 1 + 48, dtype=dtype)
    ref = (_REF8 + 48).astype(dtype)
    return pic, ref


# (ref_x, ref_y) -> expected kvz_image_calc_sad(pic, ref, 0, 0, ref_x, ref_y, 8, 8)
SAD_EDGE_CASES = [
    ((-3, -3), 1 * (4 * 4) + (2 + 4) * (4 * 4) + 5 * (4 * 4) - 64),
    ((0, -3), (1 + 3) * 4 + 2 * (6 * 4) + (4 + 6) * 4 + 5 * (6 * 4) - 64),
    ((3, -3), 3 * (4 * 4) + (2 + 6) * (4 * 4) + 5 * (4 * 4) - 64),
    ((-3, 0), (1 + 7) * 4 + 4 * (6 * 4) + (2 + 8) * 4 + 5 * (6 * 4) - 64),
    ((0, 0), (1 + 3 + 7 + 9) + (2 + 4 + 6 + 8) * 6 + 5 * (6 * 6) - 64),
    ((3, 0), (3 + 9) * 4 + 6 * (4 * 6) + (2 + 8) * 4 + 5 * (6 * 4) - 64),
    ((-3, 3), 7 * (4 * 4) + (4 + 8) * (4 * 4) + 5 * (4 * 4) - 64),
    ((0, 3), (7 + 9) * 4 + 8 * (6 * 4) + (4 + 6) * 4 + 5 * (6 * 4) - 64),
    ((3, 3), 9 * (4 * 4) + (6 + 8) * (4 * 4) + 5 * (4 * 4) - 64),
    ((-10, -10), 1 * (8 * 8) - 64),
    ((0, -10), (1 + 3) * 8 + 2 * (6 * 8) - 64),
    ((10, -10), 3 * (8 * 8) - 64),
    ((-10, 0), (1 + 7) * 8 + 4 * (6 * 8) - 64),
    ((10, 0), (3 + 9) * 8 + 6 * (6 * 8) - 64),
    ((-10, 10), 7 * (8 * 8) - 64),
    ((0, 10), (7 + 9) * 8 + 8 * (6 * 8) - 64),
    ((10, 10), 9 * (8 * 8) - 64),
]
assert len(SAD_EDGE_CASES) == 17

# sad_tests.c:373-387 (w, h) incl. SMP/AMP shapes
REG_SAD_SHAPES = [(8, 8), (16, 16), (32, 32), (64, 64), (64, 63), (1, 1),
                  (8, 16), (16, 8), (16, 32), (32, 16), (32, 64), (64, 32),
                  (48, 16), (16, 48), (24, 16), (16, 24), (12, 4), (4, 12)]


def reg_sad_pictures(dtype=np.uint8):
    """sad_tests.c:85-96: 64x64 big_pic / big_ref"""
    i = np.arange(64 * 64, dtype=np.int64)
    pic = ((i * i // 32 + i) % 255).astype(dtype).reshape(64, 64)
    ref = ((i * i // 16 + i) % 255).astype(dtype).reshape(64, 64)
    return pic, ref


# ---- satd_tests.c ----
SATD_BLACK_WHITE = {4: 2040, 8: 4080, 16: 16320, 32: 65280, 64: 261120}
SATD_CHECKERS = dict(SATD_BLACK_WHITE)
SATD_GRADIENT = {4: 3140, 8: 9004, 16: 20481, 32: 67262, 64: 258672}


def satd_buffers(test, n, dtype=np.uint8):
    """satd_tests.c:45-96: test 0 black/white, 1 checkers, 2 radial gradient."""
    lw = int(np.log2(n))
    i = np.arange(n * n, dtype=np.int64)
    if test == 0:
        a = np.zeros(n * n, dtype=np.int64)
        b = np.full(n * n, 255, dtype=np.int64)
    elif test == 1:
        a = 255 * ((((i >> lw) % 2) + (i % 2)) % 2)
        b = (a + 1) % 2
    else:
        col, row = i % n, i // n
        r = np.floor(np.sqrt((row * row + col * col).astype(np.float64))).astype(np.int64)
        a = 255 // (r + 1)
        b = 255 - 255 // (r + 1)
    return a.astype(dtype), b.astype(dtype)


SATD_GOLDEN = {0: SATD_BLACK_WHITE, 1: SATD_CHECKERS, 2: SATD_GRADIENT}


# ---- dct_tests.c:56-68: gradient block clipped to the legal residual range ----
def dct_gradient_block(n, variant=0):
    """Radial-gradient residual block like dct_tests.c's inputs: values
    255/(r+1) and the mirrored 255-255/(r+1) minus 128 (variant 1)."""
    i = np.arange(n * n, dtype=np.int64)
    col, row = i % n, i // n
    r = np.floor(np.sqrt((row * row + col * col).astype(np.float64))).astype(np.int64)
    v = 255 // (r + 1)
    if variant:
        v = 255 - v
    return (v - 128 * variant).astype(np.int16).reshape(n, n)


# ---- coeff_sum_tests.c:27-49: coeffs = -k..k style arithmetic series ----
def coeff_series(n=64 * 64, lo=-1000):
    return (np.arange(n, dtype=np.int64) + lo).astype(np.int16)
