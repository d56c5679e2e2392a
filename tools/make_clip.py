"""Writes the seeded synthetic yuv420p clips of SURVEY.md 8(d): luma = kvzlibs.synth_luma (band-limited texture,
global motion (+5,+3) px/frame, +-4 noise; seed 99 = white noise), U = Y[::2,::2]//2 + 2^(BD-2), V = 2^BD-1-U.
usage: python tools/make_clip.py OUT.yuv WIDTH HEIGHT FRAMES SEED [BITDEPTH]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import kvzlibs  # noqa: E402


def luma_frames(seed, w, h, frames, bitdepth=8):
    """seed 99: the hostile clip of SURVEY.md 8(d) -- uniform white noise, a new field every frame (no early-out ever
    fires, motion vectors are arbitrary); any other seed: kvzlibs.synth_luma"""
    if seed == 99:
        rng = np.random.default_rng(99)
        dt = np.uint8 if bitdepth == 8 else np.uint16
        return [rng.integers(0, 1 << bitdepth, (h, w)).astype(dt) for _ in range(frames)]
    return kvzlibs.synth_luma(seed, w, h, frames, bitdepth)


def clip_planes(seed, w, h, frames, bitdepth=8):
    for y in luma_frames(seed, w, h, frames, bitdepth):
        u = (y[::2, ::2] // 2 + (1 << (bitdepth - 2))).astype(y.dtype)
        v = ((1 << bitdepth) - 1 - u).astype(y.dtype)
        yield y, u, v


def write_clip(path, seed, w, h, frames, bitdepth=8):
    with open(path, "wb") as f:
        for y, u, v in clip_planes(seed, w, h, frames, bitdepth):
            f.write(y.tobytes()); f.write(u.tobytes()); f.write(v.tobytes())


if __name__ == "__main__":
    out, w, h, n, seed = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    write_clip(out, seed, w, h, n, int(sys.argv[6]) if len(sys.argv) > 6 else 8)
