"""GPU parity: batched block costs (reg_sad, sad/satd NxN, satd_any_size[_quad], SSD) and the
picture-level clamped SAD / SATD, against the reference's known-answer vectors and the oracle."""
import numpy as np
import pytest

import kat
import kvzlibs
from kvzlibs import ptr

pytestmark = pytest.mark.gpu


def _gpu_block(a, b, jobs):
    from kvazaar_b200 import blocks
    out = blocks.block_costs(blocks._dev(a), blocks._dev(b), jobs)
    return out.cpu().numpy().astype(np.uint32)


def test_satd_and_sad_known_answers():
    from kvazaar_b200 import blocks
    for test in (0, 1, 2):
        for n in (4, 8, 16, 32, 64):
            a, b = kat.satd_buffers(test, n)
            jobs = np.zeros(2, blocks.BLOCK_JOB_DT)
            jobs["a_stride"] = jobs["b_stride"] = jobs["w"] = jobs["h"] = n
            jobs["kind"] = blocks.COST_SATD_NXN
            got = _gpu_block(a, b, jobs[:1])[0]
            got_sym = _gpu_block(b, a, jobs[:1])[0]
            assert got == got_sym == kat.SATD_GOLDEN[test][n], (test, n)


def test_reg_sad_shapes_known_answers():
    from kvazaar_b200 import blocks
    pic, ref = kat.reg_sad_pictures()
    jobs = np.zeros(len(kat.REG_SAD_SHAPES), blocks.BLOCK_JOB_DT)
    jobs["a_stride"] = jobs["b_stride"] = 64
    jobs["kind"] = blocks.COST_REG_SAD
    for i, (w, h) in enumerate(kat.REG_SAD_SHAPES):
        jobs[i]["w"], jobs[i]["h"] = w, h
    got = _gpu_block(pic, ref, jobs)
    want = [int(np.abs(pic[:h, :w].astype(np.int64) - ref[:h, :w].astype(np.int64)).sum()) for (w, h) in kat.REG_SAD_SHAPES]
    assert got.tolist() == want
    # 0 vs PIXEL_MAX (sad_tests.c:294-316)
    z, m = np.zeros((64, 64), np.uint8), np.full((64, 64), 255, np.uint8)
    j = jobs[3:4].copy()
    assert _gpu_block(z, m, j)[0] == 64 * 64 * 255


def test_image_sad_edge_known_answers():
    import torch
    from kvazaar_b200 import blocks
    pic, ref = kat.sad_edge_pictures()
    jobs = np.zeros(len(kat.SAD_EDGE_CASES), blocks.IMAGE_JOB_DT)
    jobs["w"] = jobs["h"] = 8
    for i, ((rx, ry), _) in enumerate(kat.SAD_EDGE_CASES):
        jobs[i]["ref_x"], jobs[i]["ref_y"] = rx, ry
    got = blocks.image_costs(torch.from_numpy(pic).cuda(), torch.from_numpy(ref).cuda(), jobs).cpu().numpy()
    assert got.tolist() == [want for _, want in kat.SAD_EDGE_CASES]


def test_random_blocks_vs_oracle():
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(1)
    a = rs.integers(0, 256, (80, 96), dtype=np.uint8)
    b = rs.integers(0, 256, (72, 112), dtype=np.uint8)
    shapes = [(8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (8, 16), (64, 32), (12, 16), (16, 12), (4, 8),
              (8, 4), (24, 32), (4, 4), (48, 64), (64, 16), (12, 12), (20, 4)]
    jobs, want = [], []
    for (w, h) in shapes:
        for kind in (blocks.COST_REG_SAD, blocks.COST_SATD_ANY, blocks.COST_SATD_QUAD, blocks.COST_SSD):
            ox, oy = int(rs.integers(0, 8)), int(rs.integers(0, 8))
            ao, bo = oy * 96 + ox, oy * 112 + ox
            jobs.append((ao, bo, 96, 112, w, h, kind))
            av, bv = a.reshape(-1)[ao:], b.reshape(-1)[bo:]
            if kind == blocks.COST_REG_SAD:
                want.append(O.kvzo_reg_sad_8(ptr(av), ptr(bv), w, h, 96, 112))
            elif kind == blocks.COST_SATD_ANY:
                want.append(O.kvzo_satd_any_size_8(w, h, ptr(av), 96, ptr(bv), 112))
            elif kind == blocks.COST_SATD_QUAD:
                import ctypes as C
                costs = (C.c_uint * 4)()
                vp = (C.c_void_p * 4)(*[bv.ctypes.data] * 4)
                O.kvzo_satd_any_size_quad_8(w, h, vp, 112, ptr(av), 96, costs)
                want.append(costs[0])
            else:
                want.append(O.kvzo_ssd_8(ptr(av), ptr(bv), 96, 112, w))
    got = _gpu_block(a, b, np.array(jobs, blocks.BLOCK_JOB_DT))
    assert got.tolist() == want
    # contiguous NxN
    for n in (4, 8, 16, 32, 64):
        x = rs.integers(0, 256, (6, n * n), dtype=np.uint8)
        y = rs.integers(0, 256, (6, n * n), dtype=np.uint8)
        jobs = np.zeros(12, blocks.BLOCK_JOB_DT)
        jobs["a_stride"] = jobs["b_stride"] = jobs["w"] = jobs["h"] = n
        jobs["a_offset"] = jobs["b_offset"] = np.tile(np.arange(6) * n * n, 2)
        jobs["kind"][:6], jobs["kind"][6:] = blocks.COST_SAD_NXN, blocks.COST_SATD_NXN
        got = _gpu_block(x, y, jobs)
        want = [O.kvzo_sad_nxn_8(n, ptr(x[i]), ptr(y[i])) for i in range(6)] + \
               [O.kvzo_satd_nxn_8(n, ptr(x[i]), ptr(y[i])) for i in range(6)]
        assert got.tolist() == want, n


@pytest.mark.parametrize("n,count", [(4, 200), (8, 200), (4, 32), (8, 33), (4, 31)])
def test_batches_of_small_blocks_vs_oracle(n, count):
    """Batches that are all 4x4 (one job per thread) or all 8x8 (one per 8 lanes), strided and contiguous, SAD / SATD /
    reg_sad mixed, with a ragged tail that takes the warp-per-job path; and a batch with one larger block in the middle."""
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(100 + n + count)
    a = rs.integers(0, 256, (96, 160), dtype=np.uint8)
    b = rs.integers(0, 256, (96, 176), dtype=np.uint8)
    for spoil in (False, True):
        jobs, want = [], []
        for i in range(count):
            kind = (blocks.COST_REG_SAD, blocks.COST_SAD_NXN, blocks.COST_SATD_NXN)[int(rs.integers(0, 3))]
            w = h = n
            if spoil and i == 40 % count:
                w = h = 16
            ox, oy = int(rs.integers(0, 140)), int(rs.integers(0, 70))
            ao, bo = oy * 160 + ox, oy * 176 + ox
            av, bv = a.reshape(-1)[ao:], b.reshape(-1)[bo:]
            if kind == blocks.COST_SATD_NXN:
                ca = np.ascontiguousarray(a[oy:oy + h, ox:ox + w]).reshape(-1)
                cb = np.ascontiguousarray(b[oy:oy + h, ox:ox + w]).reshape(-1)
                want.append(O.kvzo_satd_nxn_8(w, ptr(ca), ptr(cb)))
            else:
                want.append(O.kvzo_reg_sad_8(ptr(av), ptr(bv), w, h, 160, 176))
            jobs.append((ao, bo, 160, 176, w, h, kind))
        got = _gpu_block(a, b, np.array(jobs, blocks.BLOCK_JOB_DT))
        assert got.tolist() == want, (n, count, spoil)


def test_image_costs_vs_oracle_1080p():
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    W, H = 1920, 1080
    f = kvzlibs.synth_luma(5, W, H, 2)
    pic, ref = f[1], f[0]
    rs = np.random.default_rng(2)
    rows, want = [], []
    for (bw, bh) in [(8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (8, 16), (32, 16), (24, 32), (12, 16), (16, 4)]:
        for _ in range(12):
            px = int(rs.integers(0, (W - bw) // 4 + 1)) * 4
            py = int(rs.integers(0, (H - bh) // 4 + 1)) * 4
            far = rs.random() < 0.3
            rx = px + int(rs.integers(-90, 91)) if not far else int(rs.integers(-200, W + 200))
            ry = py + int(rs.integers(-90, 91)) if not far else int(rs.integers(-200, H + 200))
            for kind in (blocks.IMAGE_SAD, blocks.IMAGE_SATD):
                rows.append((px, py, rx, ry, bw, bh, kind))
                fn = O.kvzo_image_sad_8 if kind == blocks.IMAGE_SAD else O.kvzo_image_satd_8
                want.append(fn(ptr(pic), ptr(ref), W, H, px, py, rx, ry, bw, bh))
    got = blocks.image_costs(torch.from_numpy(pic).cuda(), torch.from_numpy(ref).cuda(),
                             np.array(rows, blocks.IMAGE_JOB_DT)).cpu().numpy().astype(np.uint32)
    assert got.tolist() == want


def test_10bit_blocks_and_image_costs_vs_oracle():
    """bitdepth 10 (uint16 samples, KVZ_BIT_DEPTH=10 build of the reference): block costs incl. the >> (BD-8)
    of the NxN / any-size / SSD flavours, and the clamped picture-level SAD / SATD (shifted twice, image.c:507-512)"""
    import torch
    from kvazaar_b200 import blocks
    O = kvzlibs.oracle()
    rs = np.random.default_rng(10)
    a = rs.integers(0, 1024, (80, 96), dtype=np.uint16)
    b = rs.integers(0, 1024, (72, 112), dtype=np.uint16)
    a[:8, :8], b[:8, :8] = 0, 1023                                  # extremes: 0 vs PIXEL_MAX
    shapes = [(8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (64, 32), (12, 16), (4, 8), (24, 32), (4, 4), (48, 64)]
    jobs, want = [], []
    for (w, h) in shapes:
        for kind in (blocks.COST_REG_SAD, blocks.COST_SATD_ANY, blocks.COST_SSD):
            ox, oy = (0, 0) if (w, h) == (8, 8) else (int(rs.integers(0, 8)), int(rs.integers(0, 8)))
            ao, bo = oy * 96 + ox, oy * 112 + ox
            jobs.append((ao, bo, 96, 112, w, h, kind))
            av, bv = a.reshape(-1)[ao:], b.reshape(-1)[bo:]
            if kind == blocks.COST_REG_SAD:
                want.append(O.kvzo_reg_sad_10(ptr(av), ptr(bv), w, h, 96, 112))
            elif kind == blocks.COST_SATD_ANY:
                want.append(O.kvzo_satd_any_size_10(w, h, ptr(av), 96, ptr(bv), 112))
            else:
                want.append(O.kvzo_ssd_10(ptr(av), ptr(bv), 96, 112, w))
    got = blocks.block_costs(torch.from_numpy(a.view(np.int16)).cuda(), torch.from_numpy(b.view(np.int16)).cuda(),
                             np.array(jobs, blocks.BLOCK_JOB_DT), bitdepth=10).cpu().numpy().astype(np.uint32)
    assert got.tolist() == want
    for n in (4, 8, 16, 32, 64):
        x = rs.integers(0, 1024, (4, n * n), dtype=np.uint16)
        y = rs.integers(0, 1024, (4, n * n), dtype=np.uint16)
        jb = np.zeros(8, blocks.BLOCK_JOB_DT)
        jb["a_stride"] = jb["b_stride"] = jb["w"] = jb["h"] = n
        jb["a_offset"] = jb["b_offset"] = np.tile(np.arange(4) * n * n, 2)
        jb["kind"][:4], jb["kind"][4:] = blocks.COST_SAD_NXN, blocks.COST_SATD_NXN
        got = blocks.block_costs(torch.from_numpy(x.view(np.int16)).cuda(), torch.from_numpy(y.view(np.int16)).cuda(), jb,
                                 bitdepth=10).cpu().numpy().astype(np.uint32)
        want = [O.kvzo_sad_nxn_10(n, ptr(x[i]), ptr(y[i])) for i in range(4)] + \
               [O.kvzo_satd_nxn_10(n, ptr(x[i]), ptr(y[i])) for i in range(4)]
        assert got.tolist() == want, n
    # picture-level costs with the reference block partly / wholly outside the frame
    W, H = 96, 72
    pic = rs.integers(0, 1024, (H, W), dtype=np.uint16)
    ref = rs.integers(0, 1024, (H, W), dtype=np.uint16)
    rows = []
    for _ in range(60):
        w = int(rs.choice([8, 16, 32, 64])); h = int(rs.choice([8, 16, 32, 64]))
        if w > W or h > H:
            continue
        px, py = int(rs.integers(0, (W - w) // 8 + 1)) * 8, int(rs.integers(0, (H - h) // 8 + 1)) * 8
        rows.append((px, py, int(rs.integers(-70, W + 10)), int(rs.integers(-70, H + 10)), w, h, int(rs.integers(0, 2))))
    ij = np.array(rows, blocks.IMAGE_JOB_DT)
    got = blocks.image_costs(torch.from_numpy(pic.view(np.int16)).cuda(), torch.from_numpy(ref.view(np.int16)).cuda(), ij,
                             bitdepth=10).cpu().numpy().astype(np.uint32)
    for g, r in zip(got, ij):
        fn = O.kvzo_image_sad_10 if r["kind"] == 0 else O.kvzo_image_satd_10
        assert g == fn(ptr(pic), ptr(ref), W, H, int(r["pic_x"]), int(r["pic_y"]), int(r["ref_x"]), int(r["ref_y"]),
                       int(r["w"]), int(r["h"])), r
