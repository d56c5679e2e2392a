"""Frame-level motion-estimation pipeline over the batch face of the C ABI.

This is the host-side mirror of what a patched encoderstate.c / search_inter.c would do
per frame: keep the current and the reference luma planes in HBM, pad the
references, produce the dense SAD maps per reference, then run the per-PU decision stage.
torch supplies device memory, pinned host buffers, streams and events only.
"""
import ctypes as C

import torch

from . import capi
from .frame import ImeSearch, PaddedPlane


class MePipeline:
    # dram__bytes_read.sum + dram__bytes_write.sum of one ime_sad_dense_kernel<32> launch at
    # 1080p from profiles/ (ncu --set full); None until a capture has been taken
    NCU_TRAFFIC_BYTES = None

    def __init__(self, width, height, rng, n_refs, lambda_sqrt=27.0, fme_level=4, device="cuda"):
        capi.init(torch.cuda.current_device())
        self.width, self.height, self.range, self.n_refs = width, height, rng, n_refs
        self.lambda_sqrt, self.fme_level = lambda_sqrt, fme_level
        self.ime = ImeSearch(width, height, rng, device=device)
        self.cur = torch.empty((height, width), dtype=torch.uint8, device=device)
        self.refs = [torch.empty((height, width), dtype=torch.uint8, device=device) for _ in range(n_refs)]
        self.padded = [PaddedPlane(width, height, self.ime.pad, device=device) for _ in range(n_refs)]
        self.stages = ["pad_plane", "ime_sad_maps"]
        self.h2d_bytes = (1 + n_refs) * width * height
        self.d2h_bytes = 0
        self._result_host = None

    # ---- data movement ----
    def upload(self, host_cur, host_refs):
        self.cur.copy_(host_cur, non_blocking=True)
        for d, h in zip(self.refs, host_refs):
            d.copy_(h, non_blocking=True)

    # ---- one frame ----
    def step_resident(self):
        for r in range(self.n_refs):
            self.padded[r].fill_from(self.refs[r])
            self.ime.run(self.cur, self.padded[r])
            self._decide(r)

    def step_from_host(self, host_cur, host_refs):
        self.upload(host_cur, host_refs)
        self.step_resident()
        self._download()

    def _decide(self, r):
        pass

    def _download(self):
        torch.cuda.current_stream().synchronize()

    # ---- dominant kernel alone ----
    def time_sad_kernel(self, iters=10):
        self.padded[0].fill_from(self.refs[0])
        for _ in range(2):
            self.ime.run(self.cur, self.padded[0])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            self.ime.run(self.cur, self.padded[0])
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    def kernel_time_ms(self, name):
        return 0.0
