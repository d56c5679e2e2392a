"""The dense SAD launch with its reference tile rows staged by TMA (the default) and by cp.async (KVZ_CUDA_SAD_TMA=0):
one subprocess per box set and staging (the switch is read once per process), digest of all maps on stdout.
python tools/tma_probe.py -> every set both ways; python tools/tma_probe.py NAME [W H] -> one run (tests/test_gpu_ime_sad.py)"""
import os
import subprocess
import sys

SETS = {
    "x-16_32x17": [(-16, -8, 32, 17)],
    "x-15_72x19": [(-15, -5, 72, 19)],
    "x-12_32x17": [(-12, -8, 32, 17)],
    "x-1_16x41": [(-1, -30, 16, 41)],
    "x+5_24x12": [(5, -7, 24, 12)],
    "five": [(-8, -8, 32, 17), (-20, 3, 40, 25), (5, -30, 16, 41), (-9, -7, 24, 12), (-16, -5, 72, 19)],
    "one_177": [(-101, -80, 64, 177)],
}

if len(sys.argv) > 1:
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    import torch
    import kvzlibs
    import kvazaar_b200 as kb
    boxes = SETS[sys.argv[1]]
    w, h = (328, 200) if len(sys.argv) < 3 else (int(sys.argv[2]), int(sys.argv[3]))
    f = kvzlibs.synth_luma(23, w, h, len(boxes) + 1)
    cur = torch.from_numpy(f[len(boxes)]).cuda()
    multi = [kb.ImeBoxSearch(w, h, b) for b in boxes]
    pad = max(s.pad for s in multi)
    padded = [kb.PaddedPlane(w, h, pad).fill_from(torch.from_numpy(f[r]).cuda()) for r in range(len(boxes))]
    kb.ImeBoxSearch.run_many(multi, cur, padded)
    torch.cuda.synchronize()
    import hashlib
    dig = hashlib.sha1()
    for s_ in multi:
        for m_ in s_.maps:
            dig.update(m_.cpu().numpy().tobytes())
    print("ok", [tuple(s.box) for s in multi], "pad", pad, "pitch", padded[0].pitch_bytes, "sha1", dig.hexdigest())
else:
    for name in SETS:
        for env in ("0", "1"):
            e = dict(os.environ, KVZ_CUDA_SAD_TMA=env)
            p = subprocess.run([sys.executable, __file__, name], env=e, capture_output=True, text=True, timeout=120)
            last = (p.stdout.strip().splitlines() or p.stderr.strip().splitlines() or ["?"])[-1]
            print(name, "TMA", env, "rc", p.returncode, last[:230], flush=True)
