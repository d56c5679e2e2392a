"""Development aid: BASELINE configs[3] per-GPU share -- one tile (960x1088 of 3840x2160, --tiles 4x2) of a 10-bit
picture, --me full64 zero window as one box (129 x 136 displacements), dense SAD maps only; 8-bit for comparison."""
import json, os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
import kvazaar_b200 as kb

W, H = 3840, 2160
tile = (0, 0, 960, 1088)
rs = np.random.default_rng(4)
out = {}
for bd in (10, 8):
    dt = np.uint16 if bd == 10 else np.uint8
    cur = rs.integers(0, 1 << bd, (H, W)).astype(dt); ref = rs.integers(0, 1 << bd, (H, W)).astype(dt)
    tv = (lambda a: torch.from_numpy(a.view(np.int16)).cuda()) if bd == 10 else (lambda a: torch.from_numpy(a).cuda())
    srch = kb.ImeBoxSearch(W, H, (-64, -64, 136, 129), tile=tile, bitdepth=bd)
    padded = kb.PaddedPlane(W, H, srch.pad, bitdepth=bd).fill_from(tv(ref))
    d_cur = tv(cur)
    for _ in range(2):
        srch.run(d_cur, padded)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        srch.run(d_cur, padded)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    blocks8 = (tile[2] // 8) * (1080 // 8)          # picture rows of the tile (the last CTU row is partial)
    ad = blocks8 * 64 * 136 * 129
    out[f"{bd}bit"] = {"tile": "960x1088 of 3840x2160", "box": "136x129 (full64 zero window)", "ms": round(ms, 3),
                       "Tabsdiff_per_s": round(ad / ms / 1e9, 2), "tiles_per_s": round(1e3 / ms, 1)}
print(json.dumps(out))
