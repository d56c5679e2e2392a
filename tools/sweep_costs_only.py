"""quick timing of the batched SAD / SATD kernels per block size (subset of tools/kernel_sweep.py, GPU only)"""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from kvazaar_b200 import blocks

def timed(fn, iters=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

rs = np.random.default_rng(5)
total = 1 << 26
a = torch.from_numpy(rs.integers(0, 256, total, dtype=np.uint8)).cuda()
b = torch.from_numpy(rs.integers(0, 256, total, dtype=np.uint8)).cuda()
out = {}
for name, kind in (("sad", 1), ("satd", 2)):
    for n in (4, 8, 16, 32, 64):
        cnt = total // (n * n)
        jobs = np.zeros(cnt, blocks.BLOCK_JOB_DT)
        jobs["a_offset"] = jobs["b_offset"] = np.arange(cnt, dtype=np.uint32) * n * n
        jobs["a_stride"] = jobs["b_stride"] = n
        jobs["w"] = jobs["h"] = n
        jobs["kind"] = kind
        blocks.keep_on_device(jobs)
        ms = timed(lambda: blocks.block_costs(a, b, jobs))
        out[f"{name}_{n}"] = {"gpu_ms": round(ms, 3), "GB_per_s": round(2 * total / ms / 1e6, 1)}
print(json.dumps(out))
