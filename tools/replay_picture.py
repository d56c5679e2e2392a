"""Single-threaded replay of the device work of ONE picture of the config-2 encode, for ncu (the encoder process itself --
16 launching threads, results through mapped host memory -- could not be profiled: ncu answered 'Failed to profile
"me_pu_kernel"'): pad 4 references, ONE dense SAD launch with boxes of the size the encoder session picks at 1080p
(264 x 177 displacements per reference), then the PU searches of the picture through kvz_cuda_me_search_fused (the kernel
kvz_cuda_enc_search_pu launches per PU) as one launch over all 42 900 2Nx2N PUs with candidates that scatter round the
true motion the way the encoder's do.  Prints the CUDA-event time of both stages.
usage: python tools/replay_picture.py [n_pus]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import kvzlibs  # noqa: E402
import kvazaar_b200 as kb  # noqa: E402
from kvazaar_b200.frame import ME_RESULT_DT, ImeBoxSearch, MeSearch, PaddedPlane  # noqa: E402
from kvazaar_b200.pipeline import frame_pus  # noqa: E402

W, H, R, NREFS = 1920, 1080, 32, 4
BOX = (-128, -88, 264, 177)


def main():
    n_limit = int(sys.argv[1]) if len(sys.argv) > 1 else None
    torch.cuda.set_device(0)
    kb.init(0)
    f = kvzlibs.synth_luma(2, W, H, NREFS + 1)
    cur = torch.from_numpy(f[NREFS]).cuda()
    boxes = [ImeBoxSearch(W, H, BOX) for _ in range(NREFS)]
    pad = max(b.pad for b in boxes)
    planes = [PaddedPlane(W, H, pad).fill_from(torch.from_numpy(f[NREFS - 1 - i]).cuda()) for i in range(NREFS)]
    pus, jobs = frame_pus(W, H, NREFS)
    rs = np.random.default_rng(1)
    # the encoder's candidates scatter: neighbours of noisy blocks carry vectors a few samples off the true motion
    pus["merge_cand"]["mv"][:, 0, 0, :] += rs.integers(-12, 13, (len(pus), 2)).astype(np.int16)
    jobs["extra_mv"] += rs.integers(-8, 9, jobs["extra_mv"].shape).astype(np.int16)
    if n_limit:
        pus = pus[:n_limit].copy()
        jobs = jobs[:n_limit * NREFS].copy()
    ms = MeSearch(W, H, R, fme_level=4, lambda_sqrt=27.0)
    ms.set_current(cur)
    ms.set_refs(planes, [[(b.box, b.maps)] for b in boxes])
    d_pus = torch.from_numpy(pus.view(np.uint8).reshape(-1)).cuda()
    d_jobs = torch.from_numpy(jobs.view(np.uint8).reshape(-1)).cuda()
    d_res = torch.zeros(len(jobs) * ME_RESULT_DT.itemsize, dtype=torch.uint8, device="cuda")
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for it in range(2):                       # the second pass is the one ncu is pointed at (-s skips the first)
        ev[0].record()
        ImeBoxSearch.run_many(boxes, cur, planes)
        ev[1].record()
        ms.run(d_pus, len(pus), d_jobs, d_res, fused=True)
        ev[2].record()
        torch.cuda.synchronize()
    n8 = (W // 8) * (H // 8)
    absd = n8 * 64 * NREFS * BOX[2] * BOX[3]
    t_sad, t_pu = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])
    print(f"dense SAD (ime_sad_multi[_tma]_kernel, {NREFS} refs x {BOX[2]}x{BOX[3]}): {t_sad:.3f} ms = {absd / t_sad / 1e9:.1f} T abs-diff/s; "
          f"me_pu_kernel over {len(pus)} PUs x {NREFS} refs in one launch: {t_pu:.3f} ms = {t_pu * 1e3 / len(pus):.2f} us per PU of GPU throughput")


if __name__ == "__main__":
    main()
