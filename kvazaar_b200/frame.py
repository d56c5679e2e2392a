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
    search range; `run` fills them with one kernel launch.
    """

    def __init__(self, width, height, rng=32, bitdepth=8, tile=None, device="cuda"):
        capi.init(torch.cuda.current_device())
        self.width, self.height, self.range, self.bitdepth = width, height, rng, bitdepth
        self.tile = tile or (0, 0, width, height)
        self.pad = capi.ime_required_pad(rng)
        self.map_pitch = capi.ime_map_pitch(rng)
        self.k = 2 * rng + 1
        tw, th = self.tile[2], self.tile[3]
        self.ctus_x, self.ctus_y = (tw + 63) // 64, (th + 63) // 64
        n_ctu = self.ctus_x * self.ctus_y
        self.maps = []
        for level in range(4):
            n = 8 >> level
            dt = torch.int16 if level < 2 else torch.int32
            self.maps.append(torch.empty((n_ctu * n * n, self.k, self.map_pitch), dtype=dt, device=device))
        self.desc = capi.ImeDesc()

    def run(self, cur, ref_padded):
        """cur: (H, W) uint8 device tensor; ref_padded: PaddedPlane with pad >= self.pad."""
        d = self.desc
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
