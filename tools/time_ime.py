"""Quick device timing of the dense SAD kernel at 1080p (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import kvazaar_b200 as kb
import kvzlibs

w, h, r = 1920, 1080, int(sys.argv[1]) if len(sys.argv) > 1 else 32
f = kvzlibs.synth_luma(2, w, h, 2)
srch = kb.ImeSearch(w, h, r)
ref_d = torch.from_numpy(f[0]).cuda(); cur_d = torch.from_numpy(f[1]).cuda()
padded = kb.PaddedPlane(w, h, srch.pad).fill_from(ref_d)
for _ in range(3):
    srch.run(cur_d, padded)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 10
e0.record()
for _ in range(n):
    srch.run(cur_d, padded)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
ad = srch.algorithmic_abs_diffs()
cands = (32400 + 8040 + 1980 + 480) * srch.k ** 2
mapbytes = sum(m.numel() * m.element_size() for m in srch.maps)
print(f"range {r}: {ms:.3f} ms/launch  {ad/ms/1e9:.2f} Tabsdiff/s  {cands/ms/1e6:.1f} Gcand/s  "
      f"map bytes {mapbytes/1e6:.0f} MB -> {mapbytes/ms/1e6:.0f} GB/s written")
