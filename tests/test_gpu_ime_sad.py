"""GPU parity: dense full-search SAD maps (kvz_cuda_ime_sad_maps) vs the oracle's
kvz_image_calc_sad restatement, through the C ABI.  Bit-exact (integer work)."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
from kvzlibs import ptr

pytestmark = pytest.mark.gpu


def _maps_vs_oracle(width, height, rng_, pic, ref, sample_pus=None, seed=0, center=(0, 0)):
    import torch
    import kvazaar_b200 as kb
    O = kvzlibs.oracle()
    srch = kb.ImeSearch(width, height, rng_, max_center=max(abs(center[0]), abs(center[1])))
    cur_d = torch.from_numpy(pic).cuda()
    ref_d = torch.from_numpy(ref).cuda()
    padded = kb.PaddedPlane(width, height, srch.pad).fill_from(ref_d)
    maps = [m.cpu().numpy() for m in srch.run(cur_d, padded, center=center)]
    torch.cuda.synchronize()
    k = 2 * rng_ + 1
    rs = np.random.default_rng(seed)
    checked = 0
    for level in range(4):
        size = 8 << level
        pus = [(x, y) for y in range(0, height - size + 1, size) for x in range(0, width - size + 1, size)]
        if sample_pus is not None and len(pus) > sample_pus:
            idx = rs.choice(len(pus), sample_pus, replace=False)
            # always keep the corners: they exercise the edge clamp on both axes
            pus = [pus[0], pus[-1]] + [pus[i] for i in idx]
        for (x, y) in pus:
            want = np.zeros(k * k, dtype=np.uint32)
            if center == (0, 0):
                O.kvzo_dense_sad_map_8(ptr(pic), ptr(ref), width, height, x, y, size, size, rng_, ptr(want))
            else:
                for dy in range(-rng_, rng_ + 1):
                    for dx in range(-rng_, rng_ + 1):
                        want[(dy + rng_) * k + dx + rng_] = O.kvzo_image_sad_8(
                            ptr(pic), ptr(ref), width, height, x, y, x + center[0] + dx, y + center[1] + dy, size, size)
            row = srch.pu_index(level, x, y)
            got = kb.ImeSearch.window(maps[level][row]).astype(np.uint32) & (0xFFFF if level < 2 else 0xFFFFFFFF)
            assert np.array_equal(got, want.reshape(k, k)), (level, x, y)
            checked += 1
    return checked


@pytest.mark.parametrize("rng_", [8, 16, 32])
def test_small_frame_all_pus(rng_):
    # 200x136: partial CTUs on both axes (200 = 3*64+8, 136 = 2*64+8)
    w, h = 200, 136
    f = kvzlibs.synth_luma(7, w, h, 2)
    n = _maps_vs_oracle(w, h, rng_, f[1], f[0])
    assert n > 400


@pytest.mark.parametrize("center", [(5, 3), (-10, 7), (20, -12), (-1, -30)])
def test_centred_window(center):
    # the windows search_mv_full opens round extra_mv / merge candidates (search_inter.c:898-955)
    w, h = 136, 72
    f = kvzlibs.synth_luma(17, w, h, 2)
    _maps_vs_oracle(w, h, 8, f[1], f[0], sample_pus=10, seed=3, center=center)


def test_white_noise_full64():
    w, h = 128, 64
    rs = np.random.default_rng(99)
    pic = rs.integers(0, 256, (h, w), dtype=np.uint8)
    ref = rs.integers(0, 256, (h, w), dtype=np.uint8)
    _maps_vs_oracle(w, h, 64, pic, ref)


def test_extremes_no_overflow():
    # 0 vs 255 everywhere: SAD8 = 16320, SAD16 = 65280 (uint16 ceiling), SAD64 = 1044480
    w, h = 64, 64
    pic = np.zeros((h, w), dtype=np.uint8)
    ref = np.full((h, w), 255, dtype=np.uint8)
    _maps_vs_oracle(w, h, 8, pic, ref)


def test_1080p_sampled_pus():
    # BASELINE config 2 geometry; oracle on a sample of PUs (incl. the last, 56-row CTU row)
    w, h = 1920, 1080
    f = kvzlibs.synth_luma(2, w, h, 2)
    _maps_vs_oracle(w, h, 32, f[1], f[0], sample_pus=24, seed=5)


def test_hierarchy_property_1080p():
    """Size-independent property at full size: every 16/32/64 cost is the exact sum
    of its four children at the same displacement."""
    import torch
    import kvazaar_b200 as kb
    w, h, r = 1920, 1080, 32
    f = kvzlibs.synth_luma(3, w, h, 2)
    srch = kb.ImeSearch(w, h, r)
    padded = kb.PaddedPlane(w, h, srch.pad).fill_from(torch.from_numpy(f[0]).cuda())
    maps = srch.run(torch.from_numpy(f[1]).cuda(), padded)
    k = 2 * r + 1
    n_ctu = srch.ctus_x * srch.ctus_y
    # only CTUs fully inside the picture carry all four levels
    full = [cy * srch.ctus_x + cx for cy in range(h // 64) for cx in range(w // 64)]
    full_t = torch.tensor(full, device="cuda")
    prev = (kb.ImeSearch.window(maps[0].view(n_ctu, 8, 8, k + 1, -1)[full_t]).to(torch.int64) & 0xFFFF)
    for level in range(1, 4):
        n = 8 >> level
        cur = kb.ImeSearch.window(maps[level].view(n_ctu, n, n, k + 1, -1)[full_t]).to(torch.int64)
        cur = cur & (0xFFFF if level < 2 else 0xFFFFFFFF)
        s = prev.view(len(full), n, 2, n, 2, k, k).sum(dim=(2, 4))
        assert torch.equal(s, cur), level
        prev = cur


@pytest.mark.parametrize("box", [(-8, -8, 32, 17), (-20, 3, 48, 25), (5, -30, 16, 41), (-9, -7, 24, 12), (-3, -3, 8, 7),
                                 (-12, -10, 40, 30), (0, 0, 56, 1), (-16, -5, 72, 19)])
def test_box_maps_small_frame(box):
    """kvz_cuda_ime_sad_box_maps: every entry of a rectangular displacement box vs the oracle's clamped SAD."""
    import torch
    import kvazaar_b200 as kb
    w, h = 200, 136
    f = kvzlibs.synth_luma(17, w, h, 2)
    O = kvzlibs.oracle()
    srch = kb.ImeBoxSearch(w, h, box)
    x0, y0, n_dx, n_dy = srch.box
    padded = kb.PaddedPlane(w, h, srch.pad).fill_from(torch.from_numpy(f[0]).cuda())
    srch.run(torch.from_numpy(f[1]).cuda(), padded)
    torch.cuda.synchronize()
    maps = [srch.row_major(level) for level in range(4)]
    rs = np.random.default_rng(3)
    for level in range(4):
        size = 8 << level
        pus = [(x, y) for y in range(0, h - size + 1, size) for x in range(0, w - size + 1, size)]
        if len(pus) > 24:
            pus = [pus[0], pus[-1]] + [pus[i] for i in rs.choice(len(pus), 22, replace=False)]
        for (x, y) in pus:
            got = maps[level][srch.pu_index(level, x, y)].astype(np.uint32) & (0xFFFF if level < 2 else 0xFFFFFFFF)
            want = np.zeros((n_dy, n_dx), np.uint32)
            for i in range(n_dy):
                for j in range(n_dx):
                    want[i, j] = O.kvzo_image_sad_8(ptr(f[1]), ptr(f[0]), w, h, x, y, x + x0 + j, y + y0 + i, size, size)
            assert np.array_equal(got, want), (level, x, y)


def test_box_maps_multi_equals_single():
    """kvz_cuda_ime_sad_box_maps_multi (all references in one launch, heavy CTAs first) writes exactly the maps
    of one kvz_cuda_ime_sad_box_maps call per reference: full and half last chunks, different n_dy per box."""
    import torch
    import kvazaar_b200 as kb
    w, h = 328, 200                       # partial CTUs on both edges
    f = kvzlibs.synth_luma(23, w, h, 6)
    boxes = [(-8, -8, 32, 17), (-20, 3, 40, 25), (5, -30, 16, 41), (-9, -7, 24, 12), (-16, -5, 72, 19)]
    cur = torch.from_numpy(f[5]).cuda()
    single = [kb.ImeBoxSearch(w, h, b) for b in boxes]
    multi = [kb.ImeBoxSearch(w, h, b) for b in boxes]
    pad = max(s.pad for s in single)
    padded = [kb.PaddedPlane(w, h, pad).fill_from(torch.from_numpy(f[r]).cuda()) for r in range(len(boxes))]
    for s, p in zip(single, padded):
        s.run(cur, p)
    for s in multi:
        for m in s.maps:
            m.fill_(-1)
    kb.ImeBoxSearch.run_many(multi, cur, padded)
    torch.cuda.synchronize()
    n_ctu = ((w + 63) // 64) * ((h + 63) // 64)
    for r, (a, b) in enumerate(zip(single, multi)):
        for level in range(4):
            assert a.maps[level].numel() == n_ctu * (8 >> level) ** 2 * a.box[2] * a.box[3]
            assert torch.equal(a.maps[level], b.maps[level]), (r, level)


@pytest.mark.parametrize("name", ["x-15_72x19", "x-1_16x41", "five", "one_177"])
def test_tma_and_cp_async_staging_write_the_same_maps(name):
    """kvz_cuda_ime_sad_box_maps_multi stages its reference tile rows with one TMA load per CTA (cp.async.bulk.tensor,
    the default) or with cp.async words (KVZ_CUDA_SAD_TMA=0): every map entry must be the same, for window starts at
    any byte alignment (TMA itself only starts at multiples of 16 bytes).  The switch is read once per process, so
    each staging runs in a process of its own (tools/tma_probe.py); the in-process tests above check the default
    against the oracle."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    digests = []
    for tma in ("1", "0"):
        p = subprocess.run([sys.executable, os.path.join(root, "tools", "tma_probe.py"), name], capture_output=True, text=True,
                           timeout=300, env=dict(os.environ, KVZ_CUDA_SAD_TMA=tma))
        assert p.returncode == 0, p.stderr[-600:]
        digests.append(p.stdout.strip().splitlines()[-1].split("sha1")[-1].strip())
    assert len(digests[0]) == 40 and digests[0] == digests[1]


@pytest.mark.parametrize("box", [(-8, -8, 32, 17), (-20, 3, 40, 9), (5, -12, 16, 21), (-3, -3, 8, 7)])
def test_box_maps_10bit(box):
    """bitdepth 10 (uint16 samples): kvz_cuda_ime_sad_box_maps vs the oracle's kvz_image_calc_sad of the 10-bit build
    (clamped reference, >> (BD - 8)), all four PU levels, partial CTUs, full and half last chunks, extreme samples."""
    import torch
    import kvazaar_b200 as kb
    w, h = 200, 136
    rs = np.random.default_rng(41)
    f8 = np.stack(kvzlibs.synth_luma(19, w, h, 2)).astype(np.uint16)
    f = (f8 * 4 + rs.integers(0, 4, f8.shape)).astype(np.uint16)           # 10-bit texture
    f[0, :16, :16], f[1, :16, :16] = 1023, 0                                # PIXEL_MAX vs 0
    O = kvzlibs.oracle()
    srch = kb.ImeBoxSearch(w, h, box, bitdepth=10)
    x0, y0, n_dx, n_dy = srch.box
    padded = kb.PaddedPlane(w, h, srch.pad, bitdepth=10).fill_from(torch.from_numpy(f[0].view(np.int16)).cuda())
    srch.run(torch.from_numpy(f[1].view(np.int16)).cuda(), padded)
    torch.cuda.synchronize()
    maps = [srch.row_major(level) for level in range(4)]
    for level in range(4):
        size = 8 << level
        pus = [(x, y) for y in range(0, h - size + 1, size) for x in range(0, w - size + 1, size)]
        if len(pus) > 14:
            pus = [pus[0], pus[-1]] + [pus[i] for i in rs.choice(len(pus), 12, replace=False)]
        for (x, y) in pus:
            got = maps[level][srch.pu_index(level, x, y)].astype(np.uint32) & (0xFFFF if level < 2 else 0xFFFFFFFF)
            want = np.zeros((n_dy, n_dx), np.uint32)
            for i in range(n_dy):
                for j in range(n_dx):
                    want[i, j] = O.kvzo_image_sad_10(ptr(f[1]), ptr(f[0]), w, h, x, y, x + x0 + j, y + y0 + i, size, size)
            assert np.array_equal(got, want), (level, x, y)
