"""GPU parity: luma / chroma fractional interpolation (pixel, 14-bit and FME-candidate variants)
and the bi-prediction blend against the oracle, through the C ABI.  Bit-exact."""
import numpy as np
import pytest

import kvzlibs
from kvzlibs import ptr

pytestmark = pytest.mark.gpu


def _run(W, H, ref, count, seed, hostile=False, bitdepth=8):
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(seed)
    rows, off8, off16 = [], 0, 0
    for i in range(count):
        chroma = int(rs.integers(0, 2))
        w, h = [(8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (8, 16), (4, 4), (4, 8), (32, 8)][int(rs.integers(0, 9))]
        if chroma:
            w, h = min(w, 32), min(h, 32)
        elif w < 8 or h < 8:
            w, h = 8, 8
        mode = int(rs.integers(0, 3)) if not chroma else int(rs.integers(0, 2))
        x, y = int(rs.integers(0, (W - w) // 4 + 1)) * 4, int(rs.integers(0, (H - h) // 4 + 1)) * 4
        span = 130 if rs.random() < 0.8 else 4 * max(W, H)
        mvx, mvy = int(rs.integers(-span, span + 1)), int(rs.integers(-span, span + 1))
        if mode == 1:
            rows.append((x, y, mvx, mvy, w, h, chroma, mode, (0, 0), off16, w)); off16 += w * h
        else:
            rows.append((x, y, mvx, mvy, w, h, chroma, mode, (0, 0), off8, w)); off8 += w * h
    jobs = np.array(rows, blocks.INTERP_JOB_DT)
    sfx = "8" if bitdepth == 8 else "10"
    pix = np.uint8 if bitdepth == 8 else np.uint16
    d8 = torch.zeros(max(off8, 1), dtype=torch.uint8 if bitdepth == 8 else torch.int16, device="cuda")
    d16 = torch.zeros(max(off16, 1), dtype=torch.int16, device="cuda")
    blocks.interp_blocks(torch.from_numpy(ref if bitdepth == 8 else ref.view(np.int16)).cuda(), jobs, d8, d16, bitdepth)
    g8, g16 = d8.cpu().numpy().view(pix), d16.cpu().numpy()
    for jb in jobs:
        w, h, o = int(jb["w"]), int(jb["h"]), int(jb["dst_offset"])
        args = (ptr(ref), W, H, int(jb["x"]), int(jb["y"]), int(jb["mv_x"]), int(jb["mv_y"]), w, h)
        if jb["mode"] == 1:
            want = np.zeros((h, w), np.int16)
            getattr(O, ("kvzo_interp_chroma14_" if jb["chroma"] else "kvzo_interp_luma14_") + sfx)(*args, ptr(want), w)
            got = g16[o:o + w * h].reshape(h, w)
        elif jb["mode"] == 0:
            want = np.zeros((h, w), pix)
            getattr(O, ("kvzo_interp_chroma_" if jb["chroma"] else "kvzo_interp_luma_") + sfx)(*args, ptr(want), w)
            got = g8[o:o + w * h].reshape(h, w)
        else:
            # FME candidate = 14-bit (int16-wrapped) value rounded to a pixel, ipol-generic.c:285-300
            w14 = np.zeros((h, w), np.int16)
            getattr(O, "kvzo_interp_luma14_" + sfx)(*args, ptr(w14), w)
            sh = 14 - bitdepth
            want = np.clip((w14.astype(np.int32) + (1 << (sh - 1))) >> sh, 0, (1 << bitdepth) - 1).astype(pix)
            got = g8[o:o + w * h].reshape(h, w)
        assert np.array_equal(got, want), jb


def test_interp_textured():
    W, H = 160, 96
    _run(W, H, kvzlibs.synth_luma(21, W, H, 1)[0], 150, 1)


def test_interp_10bit():
    W, H = 160, 96
    ref = (kvzlibs.synth_luma(23, W, H, 1)[0].astype(np.uint16) << 2) | np.random.default_rng(5).integers(0, 4, (H, W), dtype=np.uint16)
    _run(W, H, np.ascontiguousarray(ref), 150, 6, bitdepth=10)
    hostile = np.zeros((64, 128), np.uint16)
    hostile[:, ::2] = 1023
    hostile[::2, :] ^= 1023
    _run(128, 64, hostile, 120, 7, bitdepth=10)


def test_interp_hostile_pattern():
    # alternating 0/255 columns and rows: the pattern that drives the int16 intermediates to wrap
    W, H = 128, 64
    ref = np.zeros((H, W), np.uint8)
    ref[:, ::2] = 255
    ref[::2, :] ^= 255
    _run(W, H, ref, 120, 2)
    rs = np.random.default_rng(3)
    _run(W, H, (rs.integers(0, 2, (H, W)) * 255).astype(np.uint8), 120, 4)


def test_fme_candidate_blocks_match_oracle_filters():
    # the 4 blocks one FME step emits (kvzo_fme_blocks) == 4 interp jobs in FME mode
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    W, H = 160, 96
    ref = kvzlibs.synth_luma(22, W, H, 1)[0]
    cross = [(-1, 0), (1, 0), (0, -1), (0, 1)]
    diag = [(-1, -1), (1, -1), (-1, 1), (1, 1)]
    rs = np.random.default_rng(6)
    for (w, h) in [(8, 8), (16, 16), (32, 32), (64, 64)]:
        for step in range(4):
            hx, hy = (int(rs.integers(-1, 2)), int(rs.integers(-1, 2))) if step >= 2 else (0, 0)
            x0, y0 = int(rs.integers(-20, W)), int(rs.integers(-20, H))
            pat = diag if step & 1 else cross
            rows = []
            for k in range(4):
                q = (2 * pat[k][0], 2 * pat[k][1]) if step < 2 else (2 * hx + pat[k][0], 2 * hy + pat[k][1])
                rows.append((x0, y0, q[0], q[1], w, h, 0, 2, (0, 0), k * 64 * 64, 64))
            d8 = torch.zeros(4 * 64 * 64, dtype=torch.uint8, device="cuda")
            blocks.interp_blocks(torch.from_numpy(ref).cuda(), np.array(rows, blocks.INTERP_JOB_DT), d8, None)
            want = np.zeros((4, 64, 64), np.uint8)
            O.kvzo_fme_blocks_8(ptr(ref), W, H, x0, y0, w, h, step, hx, hy, ptr(want))
            got = d8.cpu().numpy().reshape(4, 64, 64)
            assert np.array_equal(got[:, :h, :w], want[:, :h, :w]), (w, h, step)


def test_bipred_blend():
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(8)
    s0 = rs.integers(-3000, 19000, 64 * 64 * 3).astype(np.int16)
    s1 = rs.integers(-3000, 19000, 64 * 64 * 3).astype(np.int16)
    got = blocks.bipred_blend(torch.from_numpy(s0).cuda(), torch.from_numpy(s1).cuda()).cpu().numpy()
    want = np.zeros(s0.size, np.uint8)
    O.kvzo_bipred_blend_8(ptr(s0), ptr(s1), s0.size, ptr(want))
    assert np.array_equal(got, want)
    got10 = blocks.bipred_blend(torch.from_numpy(s0).cuda(), torch.from_numpy(s1).cuda(), 10).cpu().numpy().view(np.uint16)
    want10 = np.zeros(s0.size, np.uint16)
    O.kvzo_bipred_blend_10(ptr(s0), ptr(s1), s0.size, ptr(want10))
    assert np.array_equal(got10, want10)
