"""Frame-scale timing of kvz_cuda_me_search with the heuristic integer searches (hexbs / dia / tz) on the
default 1080p / 4-reference workload of bench.py (same PUs, jobs and planes); GPU box only.
Prints ms per frame for the whole call (integer search + fractional refinement) per algorithm."""
import sys
import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from kvazaar_b200.pipeline import MePipeline          # noqa: E402
import bench                                            # noqa: E402


def main():
    torch.cuda.set_device(0)
    w, h, n_refs = 1920, 1080, 4
    pipe = MePipeline(w, h, 32, n_refs)
    frames = bench.synth_frames(n_refs + 1)          # [refs..., cur] as bench.py orders them
    pipe.upload(torch.from_numpy(frames[n_refs]), [torch.from_numpy(frames[n_refs - 1 - i]) for i in range(n_refs)])
    for r in range(n_refs):
        pipe.padded[r].fill_from(pipe.refs[r])
    pipe.sad_maps()
    names = {0: "full32 (maps)", 1: "hexbs", 2: "dia", 3: "tz"}
    for early in (0, 1):
        for algo in (0, 1, 2, 3):
            fr = pipe.search.frame
            fr.ime_search, fr.me_early_termination, fr.me_max_steps = algo, early, -1
            for _ in range(2):
                pipe.search.run(pipe.d_pus, pipe.n_pus, pipe.d_jobs, pipe.d_results)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                pipe.search.run(pipe.d_pus, pipe.n_pus, pipe.d_jobs, pipe.d_results)
            e1.record()
            torch.cuda.synchronize()
            print("me_search %-14s early_termination %d: %.3f ms per frame (%d jobs)" %
                  (names[algo], early, e0.elapsed_time(e1) / 5, pipe.n_jobs), flush=True)


if __name__ == "__main__":
    main()
