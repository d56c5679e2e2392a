"""Shared helpers of the deblocking tests: the SCU record / config mirrors and a random but well-formed CU
quadtree (CU sizes 64..8, PU partitions incl. AMP, TU depth, per-PU motion, per-CU QP) for a picture."""
import ctypes as C

import numpy as np

SCU_DT = np.dtype([("type", "u1"), ("depth", "u1"), ("tr_depth", "u1"), ("part_size", "u1"), ("cbf_y", "u1"),
                   ("mv_dir", "u1"), ("mv_ref", "u1", (2,)), ("mv", "<i2", (2, 2)), ("qp", "i1"), ("reserved", "u1", (3,))])
assert SCU_DT.itemsize == 20

# cu.c:51-71: PU offsets and sizes in quarters of the CU width, [part mode][PU] -> (x, y, w, h)
PARTS = {0: [(0, 0, 4, 4)], 1: [(0, 0, 4, 2), (0, 2, 4, 2)], 2: [(0, 0, 2, 4), (2, 0, 2, 4)],
         3: [(0, 0, 2, 2), (2, 0, 2, 2), (0, 2, 2, 2), (2, 2, 2, 2)],
         4: [(0, 0, 4, 1), (0, 1, 4, 3)], 5: [(0, 0, 4, 3), (0, 3, 4, 1)],
         6: [(0, 0, 1, 4), (1, 0, 3, 4)], 7: [(0, 0, 3, 4), (3, 0, 1, 4)]}


class DeblockCfg(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("qp", C.c_int32), ("cu_qp", C.c_int32),
                ("beta_offset_div2", C.c_int32), ("tc_offset_div2", C.c_int32), ("slice_b", C.c_int32),
                ("chroma", C.c_int32), ("ref_LX", (C.c_uint8 * 16) * 2)]


def make_cfg(width, height, qp=32, cu_qp=0, beta=0, tc=0, slice_b=0, chroma=1, ref_LX=None):
    c = DeblockCfg()
    c.width, c.height, c.qp, c.cu_qp = width, height, qp, cu_qp
    c.beta_offset_div2, c.tc_offset_div2, c.slice_b, c.chroma = beta, tc, slice_b, chroma
    for l in range(2):
        for i in range(16):
            c.ref_LX[l][i] = ref_LX[l][i] if ref_LX is not None else (i if l == 0 else (i + 1) % 4)
    return c


def random_scus(rs, width, height, slice_b=False, p_intra=0.3, qp=32, qp_spread=0, n_refs=3, still=0.5):
    """SCU_DT array of (height / 4, width / 4).  `still`: probability that a PU copies the motion of the
    previous PU (so that strength-0 edges occur)."""
    scus = np.zeros((height // 4, width // 4), SCU_DT)
    last = {"mv_dir": 1, "mv_ref": (0, 0), "mv": ((0, 0), (0, 0))}

    def motion():
        nonlocal last
        if rs.random() < still:
            m = dict(last)
            if rs.random() < 0.3:                     # small change below / at the one-sample threshold
                mv = np.array(m["mv"]) + rs.integers(-4, 5, (2, 2))
                m["mv"] = tuple(map(tuple, mv.tolist()))
        else:
            m = {"mv_dir": int(rs.integers(1, 4)) if slice_b else 1,
                 "mv_ref": tuple(int(v) for v in rs.integers(0, n_refs, 2)),
                 "mv": tuple(map(tuple, rs.integers(-40, 41, (2, 2)).tolist()))}
        last = m
        return m

    def put_cu(x, y, size, depth):
        intra = rs.random() < p_intra
        cu_qp = int(qp + rs.integers(-qp_spread, qp_spread + 1)) if qp_spread else qp
        if intra:
            part = 3 if (size == 8 and rs.random() < 0.3) else 0
        else:
            choices = [0, 0, 1, 2] + ([4, 5, 6, 7] if size >= 16 else [])
            part = int(rs.choice(choices))
        tr_depth = max(depth, 1) + (1 if (size > 8 and rs.random() < 0.4) or part == 3 else 0)
        tr_depth = min(tr_depth, 4 if part == 3 else 3)
        cbf_cu = rs.random() < 0.6
        for (px, py, pw, ph) in PARTS[part]:
            m = motion()
            x0, y0, w, h = x + px * size // 4, y + py * size // 4, pw * size // 4, ph * size // 4
            blk = scus[y0 // 4:(y0 + h + 3) // 4, x0 // 4:(x0 + w + 3) // 4]
            blk["type"] = 1 if intra else 2
            blk["depth"], blk["tr_depth"], blk["part_size"], blk["qp"] = depth, tr_depth, part, cu_qp
            blk["cbf_y"] = int(cbf_cu and rs.random() < 0.8)
            if not intra:
                blk["mv_dir"], blk["mv_ref"], blk["mv"] = m["mv_dir"], m["mv_ref"], m["mv"]

    def split(x, y, size, depth):
        if x >= width or y >= height:
            return
        fits = x + size <= width and y + size <= height
        if size > 8 and (not fits or rs.random() < (0.9 if size == 64 else 0.5)):
            h = size // 2
            for dy in (0, h):
                for dx in (0, h):
                    split(x + dx, y + dy, h, depth + 1)
        else:
            put_cu(x, y, size, depth)

    for ly in range(0, height, 64):
        for lx in range(0, width, 64):
            split(lx, ly, 64, 0)
    return scus


def picture(rs, width, height, bitdepth=8, blockiness=12):
    """4:2:0 planes with block-wise offsets on an 8x8 grid over a smooth gradient: plenty of filtered edges,
    of both the strong and the weak kind, and plenty of edges left alone"""
    top = (1 << bitdepth) - 1
    dt = np.uint8 if bitdepth == 8 else np.uint16
    out = []
    for w, h, g in ((width, height, 8), (width // 2, height // 2, 8), (width // 2, height // 2, 8)):
        yy, xx = np.mgrid[0:h, 0:w]
        base = (xx * 0.7 + yy * 0.4) % 200 + 20
        blocks = rs.integers(-blockiness, blockiness + 1, ((h + g - 1) // g, (w + g - 1) // g))
        base = base + np.kron(blocks, np.ones((g, g)))[:h, :w] + rs.integers(-2, 3, (h, w))
        out.append(np.clip(base * (top + 1) / 256, 0, top).astype(dt))
    return out
