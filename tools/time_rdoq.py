"""Times kvz_cuda_rdoq_blocks on one 1080p frame's worth of transform blocks per size (CUDA events)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from kvazaar_b200 import blocks

rs = np.random.default_rng(3)
ctxs = np.frombuffer(rs.integers(0, 128, 64 * 148, dtype=np.uint8).tobytes(), blocks.RDOQ_CTX_DT).copy()
res = {}
for width in (4, 8, 16, 32):
    n = (1920 // width) * (1080 // width)
    yy, xx = np.mgrid[0:width, 0:width]
    coef = np.rint(rs.normal(0, 1, (n, width, width)) * (600.0 / (1 + 0.9 * (xx + yy)))).astype(np.int16)
    jobs = np.zeros(n, blocks.RDOQ_JOB_DT)
    jobs["coef_offset"] = np.arange(n) * width * width
    jobs["ctx_index"] = np.arange(n) % 64
    jobs["lambda"] = 40.0
    jobs["width"], jobs["qp"], jobs["cu_is_intra"] = width, 32, 1
    d = torch.from_numpy(coef.reshape(-1)).cuda()
    for sh in (False, True):
        for _ in range(3):
            out = blocks.rdoq_blocks(d, ctxs, jobs, signhide=sh)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            out = blocks.rdoq_blocks(d, ctxs, jobs, signhide=sh)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        res[f"{width}x{width}{'_sh' if sh else ''}"] = dict(blocks=n, ms=round(ms, 4), Mcoef_per_s=round(n * width * width / ms / 1e3, 1),
                                                            nonzero_frac=round(float((out != 0).float().mean()), 3))
print(json.dumps(res))
