"""Development aid: time kvz_cuda_me_search variants at 1080p (which part of the decision stage costs what)."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
import kvazaar_b200 as kb, kvzlibs
from kvazaar_b200.pipeline import MePipeline

W, H = 1920, 1080
f = kvzlibs.synth_luma(2, W, H, 5)
for name, fme, zero_cands in (("full", 4, False), ("fme0", 0, False), ("zero-cands fme4", 4, True), ("zero-cands fme0", 0, True)):
    p = MePipeline(W, H, 32, 4, fme_level=fme)
    p.upload(torch.from_numpy(f[4]).pin_memory(), [torch.from_numpy(f[3 - i]).pin_memory() for i in range(4)])
    if zero_cands:
        p.d_pus.zero_(); p.d_pus.copy_(p.host_pus)
        import kvazaar_b200.frame as fr
        pus = p.host_pus.numpy().view(fr.ME_PU_DT).copy(); pus["num_merge_cand"] = 0
        jobs = p.host_jobs.numpy().view(fr.ME_JOB_DT).copy(); jobs["extra_mv"] = 0
        p.d_pus.copy_(torch.from_numpy(pus.view(np.uint8).reshape(-1)))
        p.d_jobs.copy_(torch.from_numpy(jobs.view(np.uint8).reshape(-1)))
    p.step_resident(); torch.cuda.synchronize()
    print(f"{name:20s} me_search {p.time_me_search(3):8.3f} ms")
    del p
