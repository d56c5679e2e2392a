"""Shared cases of the coefficient bit-cost tests: quantised-coefficient-like blocks (sparse, energy near DC,
a few large levels), every width / type / scan the reference uses."""
import numpy as np

# (width, type, scan_mode): hor / ver scans only for 4x4 and 8x8 luma / 4x4 chroma intra blocks
SHAPES = [(4, 0, 0), (4, 0, 1), (4, 0, 2), (4, 2, 0), (4, 2, 1), (8, 0, 0), (8, 0, 1), (8, 0, 2), (8, 2, 0),
          (16, 0, 0), (16, 2, 0), (32, 0, 0)]


def random_block(rs, width, density, big=0.05, amp=3):
    """levels with an exponentially decaying density away from DC"""
    y, x = np.mgrid[0:width, 0:width]
    p = density * np.exp(-(x + y) / (width * 0.35))
    mask = rs.random((width, width)) < p
    lv = rs.integers(1, amp + 1, (width, width))
    lv = np.where(rs.random((width, width)) < big, lv * rs.integers(2, 40, (width, width)), lv)
    sign = np.where(rs.random((width, width)) < 0.5, -1, 1)
    return (mask * lv * sign).astype(np.int16)


def blocks(rs, width, n):
    out = []
    for i in range(n):
        kind = i % 6
        if kind == 0:
            b = np.zeros((width, width), np.int16)
            b[rs.integers(0, width), rs.integers(0, width)] = rs.integers(1, 4) * rs.choice([-1, 1])    # a single level
        elif kind == 1:
            b = random_block(rs, width, 0.3)
        elif kind == 2:
            b = random_block(rs, width, 0.9, big=0.2, amp=6)
        elif kind == 3:
            b = rs.integers(-2, 3, (width, width)).astype(np.int16)                                    # dense
        elif kind == 4:
            b = random_block(rs, width, 0.6)
            b[-1, -1] = 1                                                                               # last position at the far corner
        else:
            b = rs.integers(-3000, 3001, (width, width)).astype(np.int16)                               # escape codes everywhere
        out.append(np.ascontiguousarray(b))
    return out
