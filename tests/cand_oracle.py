"""Shared helpers of the inter candidate tests (A18): the frame description mirror, random frames of SCU records
(current picture + references) and PU lists cut from random CU quadtrees."""
import ctypes as C

import numpy as np

import deblock_oracle as do

MERGE_DT = np.dtype([("dir", "u1"), ("ref", "u1", (2,)), ("reserved", "u1"), ("mv", "<i2", (2, 2))])
assert MERGE_DT.itemsize == 12


class CandFrame(C.Structure):
    """kvzo_cand_frame == ref_cand_frame"""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("in_width", C.c_int32), ("in_height", C.c_int32),
                ("tile_offset_x", C.c_int32), ("tile_offset_y", C.c_int32), ("poc", C.c_int32), ("slice_b", C.c_int32),
                ("tmvp_enable", C.c_int32), ("n_refs", C.c_int32), ("ref_pocs", C.c_int32 * 16),
                ("col_ref_pocs", (C.c_int32 * 16) * 2), ("ref_LX", (C.c_uint8 * 16) * 2), ("ref_LX_size", C.c_uint8 * 2),
                ("reserved", C.c_uint8 * 2), ("ref_scu_pitch", C.c_int32), ("scus", C.c_void_p), ("ref_scus", C.c_void_p * 16)]


class Scene:
    """a current picture and its references as SCU arrays + the POC / list bookkeeping"""

    def __init__(self, rs, width, height, slice_b=False, n_refs=3, tmvp=True, p_intra=0.25, poc=8):
        self.width, self.height, self.slice_b, self.n_refs, self.tmvp, self.poc = width, height, slice_b, n_refs, tmvp, poc
        self.scus = np.ascontiguousarray(do.random_scus(rs, width, height, slice_b, p_intra, n_refs=min(n_refs, 2), still=0.4))
        self.ref_scus = [np.ascontiguousarray(do.random_scus(rs, width, height, True, 0.2, n_refs=3, still=0.3)) for _ in range(n_refs)]
        if slice_b:
            # pictures 0.. before and after the current one; L0 = past first, L1 = future first
            pocs = [poc - 1 - 2 * (i // 2) if i % 2 == 0 else poc + 1 + 2 * (i // 2) for i in range(n_refs)]
            past = [i for i in range(n_refs) if pocs[i] < poc]
            future = [i for i in range(n_refs) if pocs[i] > poc]
            self.lists = [past + future, (future + past) if future else past]
        else:
            pocs = [poc - 1 - i for i in range(n_refs)]
            self.lists = [list(range(n_refs)), []]
        self.ref_pocs = pocs
        # POCs the colocated picture's own references had, per list and index
        self.col_ref_pocs = [[int(v) for v in rs.integers(poc - 12, poc + 6, 16)] for _ in range(2)]
        # a picture never references itself (the POC distance is a divisor, inter.c:975)
        col_poc = pocs[self.lists[0][0]]
        self.col_ref_pocs = [[v if v != col_poc else col_poc - 3 for v in l] for l in self.col_ref_pocs]
        # mv_ref of the current picture's SCUs must index the lists
        for l in range(2):
            n = max(1, len(self.lists[l]))
            self.scus["mv_ref"][..., l] %= n
        if not slice_b:
            self.scus["mv_dir"][self.scus["type"] == 2] = 1
        for r in self.ref_scus:
            r["mv_ref"] %= 4

    def frame(self):
        f = CandFrame()
        f.width = f.in_width = self.width
        f.height = f.in_height = self.height
        f.poc, f.slice_b, f.tmvp_enable, f.n_refs = self.poc, int(self.slice_b), int(self.tmvp), self.n_refs
        for i, p in enumerate(self.ref_pocs):
            f.ref_pocs[i] = p
        for l in range(2):
            f.ref_LX_size[l] = len(self.lists[l])
            for i, r in enumerate(self.lists[l]):
                f.ref_LX[l][i] = r
            for i in range(16):
                f.col_ref_pocs[l][i] = self.col_ref_pocs[l][i]
        f.ref_scu_pitch = self.width // 4
        f.scus = self.scus.ctypes.data
        for i, r in enumerate(self.ref_scus):
            f.ref_scus[i] = r.ctypes.data
        return f

    def pus(self, rs, count):
        """(x, y, w, h, use_a1, use_b1) of random PUs: CU sizes 8..64 inside the frame, every partition mode"""
        out = []
        while len(out) < count:
            size = int(rs.choice([8, 16, 32, 64]))
            if size > self.width or size > self.height:
                continue
            x = int(rs.integers(0, self.width // size)) * size
            y = int(rs.integers(0, self.height // size)) * size
            part = int(rs.choice([0, 0, 1, 2] + ([4, 5, 6, 7] if size >= 16 else [])))
            for i, (px, py, pw, ph) in enumerate(do.PARTS[part]):
                w, h = pw * size // 4, ph * size // 4
                out.append((x + px * size // 4, y + py * size // 4, w, h, int(i == 0 or w >= h), int(i == 0 or w <= h)))
        return out[:count]
