"""Frame-level host helpers over the batch face of the C ABI.

torch is used for device memory and streams only; all arithmetic happens inside
libkvz_cuda.so.  Names follow the encoder's domain: planes, CTUs, PUs, SAD maps.
"""
import ctypes as C

import torch

from . import capi


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class PaddedPlane:
    """A luma/chroma plane in HBM with replicated borders.

    Reading the border is the coordinate clamp the reference applies to
    out-of-frame motion vectors (image.c:320-452, ipol-generic.c:761-781).
    """

    def __init__(self, width, height, pad, bitdepth=8, device="cuda"):
        self.width, self.height, self.pad, self.bitdepth = width, height, pad, bitdepth
        self.pitch = (width + 2 * pad + 127) & ~127          # samples per row
        dt = torch.uint8 if bitdepth == 8 else torch.int16
        self.data = torch.empty((height + 2 * pad, self.pitch), dtype=dt, device=device)

    def fill_from(self, plane):
        """plane: (H, W) device tensor, contiguous rows (pitch = plane.stride(0))."""
        assert plane.is_cuda and plane.shape == (self.height, self.width)
        capi.check(capi.lib().kvz_cuda_pad_plane(
            C.c_void_p(plane.data_ptr()), self.width, self.height, plane.stride(0),
            C.c_void_p(self.data.data_ptr()), self.pad, self.pitch, self.bitdepth, _stream_ptr()),
            "kvz_cuda_pad_plane")
        return self

    @property
    def pitch_bytes(self):
        return self.pitch * self.data.element_size()


class ImeSearch:
    """Dense integer motion estimation of one picture against one reference.

    Owns the four SAD maps (8x8, 16x16, 32x32, 64x64 PUs) for a picture size and
    search range; `run` fills them (kvz_cuda_ime_sad_maps).  A PU's map is K + 1 rows of
    `map_pitch` entries: rows 0..K-1 hold dx = -R..R-1 of each dy, row K holds the column
    dx = +R (entry dy + R) -- see include/kvz_cuda.h; `window()` undoes that.
    """

    def __init__(self, width, height, rng=32, bitdepth=8, tile=None, device="cuda", max_center=0):
        capi.init(torch.cuda.current_device())
        self.width, self.height, self.range, self.bitdepth = width, height, rng, bitdepth
        self.tile = tile or (0, 0, width, height)
        # border the reference planes need: the search range plus the farthest window centre
        self.pad = capi.ime_required_pad(rng) + ((max_center + 31) & ~31)
        self.map_pitch = capi.ime_map_pitch(rng)
        self.k = 2 * rng + 1
        tw, th = self.tile[2], self.tile[3]
        self.ctus_x, self.ctus_y = (tw + 63) // 64, (th + 63) // 64
        n_ctu = self.ctus_x * self.ctus_y
        self.maps = []
        for level in range(4):
            n = 8 >> level
            dt = torch.int16 if level < 2 else torch.int32
            self.maps.append(torch.empty((n_ctu * n * n, self.k + 1, self.map_pitch), dtype=dt, device=device))
        self.desc = capi.ImeDesc()

    def run(self, cur, ref_padded, center=(0, 0), maps=None):
        """cur: (H, W) uint8 device tensor; ref_padded: PaddedPlane with pad >= self.pad;
        center: window centre (whole samples); maps: the four tensors to fill (default self.maps)."""
        if maps is not None:
            self.maps = maps
        d = self.desc
        d.center_x, d.center_y = center
        d.width, d.height, d.bitdepth, d.range = self.width, self.height, self.bitdepth, self.range
        d.tile_x, d.tile_y, d.tile_w, d.tile_h = self.tile
        d.cur_pitch = cur.stride(0) * cur.element_size()
        d.ref_pitch = ref_padded.pitch_bytes
        d.ref_pad = ref_padded.pad
        capi.check(capi.lib().kvz_cuda_ime_sad_maps(
            C.byref(d), C.c_void_p(cur.data_ptr()), C.c_void_p(ref_padded.data.data_ptr()),
            C.c_void_p(self.maps[0].data_ptr()), C.c_void_p(self.maps[1].data_ptr()),
            C.c_void_p(self.maps[2].data_ptr()), C.c_void_p(self.maps[3].data_ptr()), _stream_ptr()),
            "kvz_cuda_ime_sad_maps")
        return self.maps

    @staticmethod
    def window(m):
        """(..., K + 1, pitch) map storage (torch or numpy) -> (..., K, K) window, [dy + R, dx + R]."""
        k = m.shape[-2] - 1
        out = m[..., :k, :k].clone() if isinstance(m, torch.Tensor) else m[..., :k, :k].copy()
        out[..., :, k - 1] = m[..., k, :k]
        return out

    def pu_index(self, level, x, y):
        """Map row of the PU of size 8<<level whose top-left luma sample is (x, y)
        (tile-relative)."""
        n = 8 >> level
        size = 8 << level
        ctu = (y // 64) * self.ctus_x + (x // 64)
        return ctu * n * n + ((y % 64) // size) * n + (x % 64) // size

    def algorithmic_bytes(self):
        """Bytes one launch must move: the four SAD maps written once (the entries that
        exist, K x K per PU) plus the two luma planes read once."""
        tw, th = self.tile[2], self.tile[3]
        kk = self.k * self.k
        pus = [(tw // (8 << l)) * (th // (8 << l)) for l in range(4)]
        return kk * (2 * pus[0] + 2 * pus[1] + 4 * pus[2] + 4 * pus[3]) + 2 * tw * th

    def algorithmic_abs_diffs(self):
        """N8 * 64 * (2R+1)^2 pixel differences per launch."""
        tw, th = self.tile[2], self.tile[3]
        return (tw // 8) * (th // 8) * 64 * self.k * self.k


class ImeBoxSearch:
    """Dense SAD maps over a rectangular box of displacements (kvz_cuda_ime_sad_box_maps), n_dx rounded up to a
    multiple of 8.  The device layout is CHUNK-MAJOR: map[level][chunk][pu][dy - y0][16] (a half last chunk is 8
    wide); `maps[level]` is the flat buffer, `row_major(level)` reassembles [pu][dy - y0][dx - x0] for tests."""

    def __init__(self, width, height, box, tile=None, device="cuda", bitdepth=8):
        capi.init(torch.cuda.current_device())
        self.width, self.height, self.bitdepth = width, height, bitdepth
        self.tile = tile or (0, 0, width, height)
        x0, y0, n_dx, n_dy = box
        n_dx = (n_dx + 7) & ~7
        self.box = (x0, y0, n_dx, n_dy)
        far = max(abs(x0), abs(x0 + n_dx), abs(y0), abs(y0 + n_dy))
        self.pad = (104 + far + 31) & ~31
        tw, th = self.tile[2], self.tile[3]
        self.ctus_x, self.ctus_y = (tw + 63) // 64, (th + 63) // 64
        n_ctu = self.ctus_x * self.ctus_y
        self.maps = []
        for level in range(4):
            n = 8 >> level
            dt = torch.int16 if level < 2 else torch.int32
            self.maps.append(torch.empty(n_ctu * n * n * n_dy * n_dx, dtype=dt, device=device))
        self.desc = capi.ImeBoxDesc()

    def row_major(self, level):
        """host copy of one level as [pu][n_dy][n_dx] (numpy)"""
        x0, y0, n_dx, n_dy = self.box
        n = 8 >> level
        pus = self.ctus_x * self.ctus_y * n * n
        flat = self.maps[level].cpu().numpy()
        parts, at = [], 0
        for c0 in range(0, n_dx, 16):
            w = min(16, n_dx - c0)
            parts.append(flat[at:at + pus * n_dy * w].reshape(pus, n_dy, w))
            at += pus * n_dy * w
        return np.concatenate(parts, axis=2)

    pu_index = ImeSearch.pu_index

    def run(self, cur, ref_padded, maps=None):
        if maps is not None:
            self.maps = maps
        d = self.desc
        self._fill_desc(d, cur, ref_padded)
        capi.check(capi.lib().kvz_cuda_ime_sad_box_maps(
            C.byref(d), C.c_void_p(cur.data_ptr()), C.c_void_p(ref_padded.data.data_ptr()),
            C.c_void_p(self.maps[0].data_ptr()), C.c_void_p(self.maps[1].data_ptr()),
            C.c_void_p(self.maps[2].data_ptr()), C.c_void_p(self.maps[3].data_ptr()), _stream_ptr()),
            "kvz_cuda_ime_sad_box_maps")
        return self.maps

    def map_bytes(self):
        return sum(m.numel() * m.element_size() for m in self.maps)

    def _fill_desc(self, d, cur, ref_padded):
        d.width, d.height, d.bitdepth = self.width, self.height, self.bitdepth
        d.tile_x, d.tile_y, d.tile_w, d.tile_h = self.tile
        d.cur_pitch = cur.stride(0) * cur.element_size()
        d.ref_pitch, d.ref_pad = ref_padded.pitch_bytes, ref_padded.pad
        d.x0, d.y0, d.n_dx, d.n_dy = self.box

    @staticmethod
    def run_many(searches, cur, refs_padded):
        """kvz_cuda_ime_sad_box_maps_multi: the boxes of several references of one current picture in a
        single launch (up to 8 per call)."""
        for i0 in range(0, len(searches), capi.IME_MAX_MULTI):
            grp = searches[i0:i0 + capi.IME_MAX_MULTI]
            arr = (capi.ImeBoxRef * len(grp))()
            for a, s, rp in zip(arr, grp, refs_padded[i0:i0 + capi.IME_MAX_MULTI]):
                s._fill_desc(a.desc, cur, rp)
                a.d_ref_padded = rp.data.data_ptr()
                a.d_sad8, a.d_sad16, a.d_sad32, a.d_sad64 = (m.data_ptr() for m in s.maps)
            capi.check(capi.lib().kvz_cuda_ime_sad_box_maps_multi(arr, len(grp), C.c_void_p(cur.data_ptr()), _stream_ptr()),
                       "kvz_cuda_ime_sad_box_maps_multi")


# numpy mirrors of the job structures of include/kvz_cuda.h
import numpy as np  # noqa: E402

MERGE_CAND_DT = np.dtype([("dir", "u1"), ("ref", "u1", (2,)), ("reserved", "u1"), ("mv", "<i2", (2, 2))])
ME_PU_DT = np.dtype([("x", "<i2"), ("y", "<i2"), ("w", "u1"), ("h", "u1"), ("num_merge_cand", "u1"),
                     ("num_jobs", "u1"), ("first_job", "<u4"), ("merge_cand", MERGE_CAND_DT, (5,))])
ME_JOB_DT = np.dtype([("mv_cand", "<i2", (2, 2)), ("extra_mv", "<i2", (2,)), ("ref_idx", "u1"),
                      ("reserved", "u1", (3,))])
ME_RESULT_DT = np.dtype([("mv", "<i2", (2,)), ("cost", "<u4"), ("bitcost", "<u4"), ("merged", "u1"),
                         ("merge_idx", "u1"), ("mv_cand_idx", "u1"), ("took", "u1")])
assert MERGE_CAND_DT.itemsize == 12 and ME_PU_DT.itemsize == 72
assert ME_JOB_DT.itemsize == 16 and ME_RESULT_DT.itemsize == 16


class MeSearch:
    """Per-PU motion decision (kvz_cuda_me_search) for one current picture against up to 16
    references; mirrors the reference loop of search_pu_inter (search_inter.c:1451-1524)."""

    def __init__(self, width, height, rng=32, fme_level=4, lambda_sqrt=27.0, tile=None, owf=0, wpp=0,
                 sao_type=3, deblock_enable=1, mv_constraint=0, max_ref_lcu=(1, 1), ref_LX=None,
                 ime_search=0, me_early_termination=0, me_max_steps=-1):
        """ime_search: 0 full search (dense maps), 1 hexbs, 2 diamond, 3 tz (probes evaluated on demand)"""
        capi.init(torch.cuda.current_device())
        fr = capi.MeFrame()
        fr.ime_search, fr.me_early_termination, fr.me_max_steps = ime_search, me_early_termination, me_max_steps
        fr.width, fr.height, fr.bitdepth, fr.range = width, height, 8, rng
        fr.tile_x, fr.tile_y, fr.tile_w, fr.tile_h = tile or (0, 0, width, height)
        fr.lambda_sqrt_i = int(lambda_sqrt + 0.5)
        fr.fme_level, fr.owf, fr.wpp, fr.sao_type = fme_level, owf, wpp, sao_type
        fr.deblock_enable, fr.mv_constraint = deblock_enable, mv_constraint
        fr.max_ref_lcu_right, fr.max_ref_lcu_down = max_ref_lcu
        for l in range(2):
            for i in range(16):
                fr.ref_LX[l][i] = ref_LX[l][i] if ref_LX is not None else i
        self.frame = fr
        self._keep = []

    def set_current(self, cur):
        self.frame.d_cur = cur.data_ptr()
        self.frame.cur_pitch = cur.stride(0) * cur.element_size()
        if not any(k is cur for k in self._keep):
            self._keep.append(cur)

    def set_refs(self, padded_planes, maps=None):
        """padded_planes: list of PaddedPlane; maps: per reference either the 4 zero-window map
        tensors, or a list of map sets -- ((center_x, center_y), [4 tensors]) for a square window,
        ((x0, y0, n_dx, n_dy), [4 tensors]) for a box -- or None."""
        self.frame.n_refs = len(padded_planes)
        for i, pp in enumerate(padded_planes):
            r = self.frame.refs[i]
            r.d_ref_padded, r.ref_pitch, r.ref_pad = pp.data.data_ptr(), pp.pitch_bytes, pp.pad
            m = maps[i] if maps is not None else None
            if m is None:
                sets = []
            elif isinstance(m[0], tuple):
                sets = m
            else:
                sets = [((0, 0), m)]
            r.n_mapsets = len(sets)
            for k, (c, t) in enumerate(sets):
                ms = r.mapsets[k]
                ms.center_x, ms.center_y = c[0], c[1]
                ms.n_dx, ms.n_dy = (c[2], c[3]) if len(c) == 4 else (0, 0)
                ms.d_sad8, ms.d_sad16, ms.d_sad32, ms.d_sad64 = [x.data_ptr() for x in t]
        self._keep.append((padded_planes, maps))

    def run(self, d_pus, n_pus, d_jobs, d_results, fused=False):
        """device tensors (uint8 views of the ME_*_DT arrays); fused: kvz_cuda_me_search_fused (one CTA per PU, one launch)"""
        fn = capi.lib().kvz_cuda_me_search_fused if fused else capi.lib().kvz_cuda_me_search
        capi.check(fn(C.byref(self.frame), C.c_void_p(d_pus.data_ptr()), n_pus, C.c_void_p(d_jobs.data_ptr()),
                      C.c_void_p(d_results.data_ptr()), _stream_ptr()), "kvz_cuda_me_search")
        return d_results


def inter_candidates(scene, d_scus, d_ref_scus, d_pus, n_pus, d_jobs, d_pu_flags=None):
    """kvz_cuda_inter_candidates: `scene` supplies the frame bookkeeping (width, height, poc, slice_b, tmvp,
    ref_pocs, col_ref_pocs, lists = [L0 picture indices, L1 picture indices]); d_scus / d_ref_scus are device uint8
    tensors of kvz_cuda_scu records; d_pus / d_jobs are filled in place (merge lists, AMVP predictors, extra_mv)."""
    capi.init(torch.cuda.current_device())
    f = capi.CandFrame()
    f.width = f.in_width = scene.width
    f.height = f.in_height = scene.height
    f.poc, f.slice_b, f.tmvp_enable, f.n_refs = scene.poc, int(scene.slice_b), int(scene.tmvp), len(d_ref_scus)
    for i, p in enumerate(scene.ref_pocs):
        f.ref_pocs[i] = p
    for l in range(2):
        f.ref_LX_size[l] = len(scene.lists[l])
        for i, r in enumerate(scene.lists[l]):
            f.ref_LX[l][i] = r
        for i in range(16):
            f.col_ref_pocs[l][i] = scene.col_ref_pocs[l][i]
    f.ref_scu_pitch = scene.width // 4
    f.d_scus = d_scus.data_ptr()
    for i, r in enumerate(d_ref_scus):
        f.d_ref_scus[i] = r.data_ptr()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    capi.check(capi.lib().kvz_cuda_inter_candidates(C.byref(f), C.c_void_p(d_pus.data_ptr()), n_pus,
                                                    C.c_void_p(d_pu_flags.data_ptr()) if d_pu_flags is not None else None,
                                                    C.c_void_p(d_jobs.data_ptr()), stream), "kvz_cuda_inter_candidates")
