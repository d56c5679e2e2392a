"""Development aid: device time of the box SAD stage of the bench pipeline (all references)."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch
import kvzlibs
from kvazaar_b200.pipeline import MePipeline
W, H = 1920, 1080
f = kvzlibs.synth_luma(2, W, H, 5)
p = MePipeline(W, H, 32, 4)
p.upload(torch.from_numpy(f[4]).pin_memory(), [torch.from_numpy(f[3 - i]).pin_memory() for i in range(4)])
p.sad_maps(); torch.cuda.synchronize()
for multi in (True, False):
    p.multi_launch = multi
    print("multi" if multi else "per-ref", "SAD stage %.4f ms" % p.time_sad_kernel(20))
