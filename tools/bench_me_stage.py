#!/usr/bin/env python
"""tools/bench_me_stage.py -- round 1's bench.py (the headline bench is now the encode, /bench.py): the flat ME stage
of BASELINE.json config 2 on B200 over PRE-MADE candidates (every PU the same merge / AMVP candidates: the best case of
the kernels, not an encode): full-search motion estimation
of 1920x1080 8-bit frames (--me full32, -r 4: every 8x8/16x16/32x32/64x64 PU of the frame
against 4 reference pictures, 65x65 integer candidates each).

One "step" = one frame: for each of the 4 references pad the reference plane, produce the
dense SAD maps (kvz_cuda_ime_sad_maps) and run the per-PU motion decision
(kvz_cuda_me_search: exact search_mv_full + search_frac semantics, search_inter.c:886-1128).

  value : ME Gcand/s with the planes already resident in HBM
  e2e   : the same through the C ABI from HOST buffers (H2D of the current + reference luma
          planes and D2H of the per-PU results inside the timed region)
  --impl reference : the reference's own CPU code (oracle/_ref, search_mv_full +
          search_frac through ref_shim) on all host threads, on a bounded sample of PUs.

Prints ONE JSON line (contract in the task statement / DESIGN.md section Measurement).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

W, H, RANGE, NREFS = 1920, 1080, 32, 4
PU_COUNTS = {8: 32400, 16: 8040, 32: 1980, 64: 480}       # 42900 PUs
METRIC = "ME Gcand/s (1080p --me full32 -r 4: integer + fractional motion search, all 2Nx2N PUs 8..64)"
UNIT = "Gcand/s"


def cands_per_frame():
    k = 2 * RANGE + 1
    return sum(PU_COUNTS.values()) * k * k * NREFS


def synth_frames(n, seed=2):
    import kvzlibs
    return kvzlibs.synth_luma(seed, W, H, n)


# ----------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ----------------------------------------------------------------------------------------
class Clocks:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.samples, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        hi = sorted(sm)[len(sm) // 2:]           # upper half = samples under load
        return {"sm_mhz": float(np.median(hi)), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------------------
# reference arm: the reference's own CPU code on host threads
# ----------------------------------------------------------------------------------------
def reference_step_fn(frames, pus_per_level, threads, fme_level=4):
    """Returns (fn, cands_per_call, description): fn() runs search_mv_full + search_frac of the
    compiled reference (oracle/_ref via ref_shim, search_inter.c:886-1128) for a sample of the
    frame's PUs -- same candidates as the GPU arm (kvazaar_b200.pipeline.frame_pus) -- against
    NREFS references on `threads` host threads, carrying *inter_cost over each PU's reference
    loop like search_pu_inter (search_inter.c:1505-1511)."""
    import kvzlibs
    from concurrent.futures import ThreadPoolExecutor
    from kvazaar_b200.pipeline import frame_pus
    L = kvzlibs.ref(1)
    cur, refs = frames[NREFS], [frames[NREFS - 1 - i] for i in range(NREFS)]
    pus, jobs = frame_pus(W, H, NREFS)
    rs = np.random.default_rng(7)
    chosen = []
    for size, n in pus_per_level.items():
        idx = np.nonzero(pus["w"] == size)[0]
        chosen.extend(rs.choice(idx, n, replace=False).tolist())
    work = []
    for pi in chosen:
        pu = pus[pi]
        per_ref = []
        for q in range(NREFS):
            jb = jobs[int(pu["first_job"]) + q]
            merge = [(int(m["dir"]), int(m["ref"][0]), int(m["ref"][1]), int(m["mv"][0][0]), int(m["mv"][0][1]),
                      int(m["mv"][1][0]), int(m["mv"][1][1])) for m in pu["merge_cand"][:int(pu["num_merge_cand"])]]
            _, rj = kvzlibs.job_pair(
                width=W, height=H, origin_x=int(pu["x"]), origin_y=int(pu["y"]), pu_w=int(pu["w"]), pu_h=int(pu["h"]),
                tile_off_x=0, tile_off_y=0, ref_idx=int(jb["ref_idx"]), lambda_sqrt=LAMBDA_SQRT,
                mv_cand=jb["mv_cand"].tolist(), merge=merge, extra_mv=tuple(int(v) for v in jb["extra_mv"]),
                owf=0, wpp=0, sao_type=3, deblock_enable=1, mv_constraint=0, max_ref_lcu_right=1,
                max_ref_lcu_down=1, fme_level=fme_level, me_early_termination=0, me_max_steps=-1,
                ime_algorithm=5)
            rj.pic_y, rj.ref_y = cur.ctypes.data, refs[q].ctypes.data
            per_ref.append(rj)
        work.append(per_ref)
    k = 2 * RANGE + 1

    def one(per_ref):
        inter_cost = 0x7FFFFFFF
        for j in per_ref:
            L.ref_shim_ime(C.byref(j))
            if fme_level > 0 and j.best_cost < inter_cost:
                L.ref_shim_fme(C.byref(j))
            elif j.best_cost < 0xFFFFFFFF:
                L.ref_shim_int_satd(C.byref(j))
            if j.best_cost < inter_cost:
                inter_cost = j.best_cost

    pool = ThreadPoolExecutor(threads)

    def fn():
        list(pool.map(one, work))
    desc = f"{sum(pus_per_level.values())} PUs ({pus_per_level}) x {NREFS} refs of one 1080p frame"
    return fn, len(work) * NREFS * k * k, desc


# bounded CPU sample: 1/6 of the frame's PUs at every size (same mix as the full frame)
REF_SAMPLE = {8: 5400, 16: 1340, 32: 330, 64: 80}
LAMBDA_SQRT = 27.0     # sqrt(lambda) of a QP-32-class frame; cost = bits * (int)(lambda_sqrt + .5)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    frames = synth_frames(NREFS + 1)
    pus = REF_SAMPLE
    fn, cands, sample = reference_step_fn(frames, pus, threads)
    for _ in range(max(1, min(args.warmup, 1))):
        fn()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    dt = (time.perf_counter() - t0) / args.steps
    v = cands / dt / 1e9
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "1920x1080 8-bit, --me full32 -r 4 (BASELINE configs[1]), ME stage",
                   "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ----------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    import kvazaar_b200 as kb

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback on the hot path)")
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION") == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"      # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    kb.init(local)

    # every rank searches its own frames (same-temporal-layer frames of the GOP-8 pyramid are
    # independent): weak scaling, no data-path collective inside the ME stage
    frames = synth_frames(NREFS + 1, seed=2 + rank)
    pipe = kb.MePipeline(W, H, RANGE, NREFS, lambda_sqrt=LAMBDA_SQRT, fme_level=4, use_boxes=not args.separate_windows)
    host_cur = torch.from_numpy(frames[NREFS]).pin_memory()
    host_refs = [torch.from_numpy(frames[NREFS - 1 - i]).pin_memory() for i in range(NREFS)]
    pipe.upload(host_cur, host_refs)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    from kvazaar_b200 import shard
    counter = [0]

    published = torch.empty_like(pipe.refs[0]) if world > 1 else None

    def exchange():
        # the one exchange step of the sharded pipeline: the rank that reconstructed the newest
        # reference publishes its luma plane to all ranks over NCCL/NVLink.  It arrives in a buffer of its own:
        # the host-buffer steps upload into pipe.refs[] on another stream, which a broadcast into the same
        # tensor would race with (the published plane is what a frame-sharded encoder would search next)
        if world > 1:
            if counter[0] % world == rank:
                published.copy_(pipe.refs[0])
            shard.publish_reference(published, src=counter[0] % world)
            counter[0] += 1

    def step_resident():
        exchange()
        pipe.step_resident()

    def step_e2e():
        exchange()
        pipe.step_from_host(host_cur, host_refs)

    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    for _ in range(max(args.warmup, 3)):
        step_resident()
    l0 = kb.launch_count()
    ms = timed(step_resident, args.steps)
    launches = kb.launch_count() - l0

    # e2e through the C ABI from host buffers: every step uploads its inputs (H2D stream) and downloads its
    # results (D2H stream); consecutive steps overlap their copies with each other's kernels
    def step_e2e_pipelined():
        exchange()
        pipe.step_from_host_pipelined(host_cur, host_refs)

    for _ in range(2):
        step_e2e()
    ms_e2e_serial = timed(step_e2e, args.steps)
    for _ in range(2):
        step_e2e_pipelined()
    pipe.flush_pipelined()

    def run_pipelined():
        step_e2e_pipelined()

    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        run_pipelined()
    pipe.flush_pipelined()                   # the last step's results are on the host
    torch.cuda.current_stream().wait_stream(pipe._d2h)
    e1.record()
    barrier()
    ms_t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms_e2e = float(ms_t.item())
    if world > 1:
        pipe.upload(host_cur, host_refs)      # restore this rank's own reference 0 for the kernel timings

    # per-kernel times, each kernel alone with events on the launching stream
    sad_ms = pipe.time_sad_kernel(iters=max(5, args.steps))
    search_ms = pipe.time_me_search(iters=max(3, args.steps // 2))
    # keep the same load up for >= 1.5 s so that nvidia-smi (100 ms period) sees it
    t_end = time.perf_counter() + 1.5
    while time.perf_counter() < t_end:
        pipe.step_resident()
        torch.cuda.synchronize()
    clk = clocks.stop() if rank == 0 else None
    if clk is not None:
        clk["note"] = "sampled every 100 ms from warm-up to the end of 1.5 s of identical untimed steps"

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    cands = cands_per_frame() * world
    value = cands / (ms / args.steps * 1e-3) / 1e9
    e2e = cands / (ms_e2e / args.steps * 1e-3) / 1e9
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    int_peak = kb.measure_int_pipe_peak()
    abs_diffs = pipe.algorithmic_abs_diffs            # of all SAD launches of one step
    map_bytes = pipe.algorithmic_map_bytes + NREFS * 2 * W * H
    n_sad = pipe.n_sad_launches
    achieved_ops = abs_diffs / (sad_ms * 1e-3)
    step_ms = ms / args.steps
    kernel = "ime_sad_multi_kernel (boxes of all references, one launch)" if pipe.use_boxes else "ime_sad_dense_kernel<32>"
    traffic = None                                    # not measured in this run; see roofline.traffic_source
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "1920x1080 8-bit, --me full32 -r 4 (BASELINE configs[1]), ME stage: "
                               "pad + dense SAD maps + per-PU decision (search_mv_full) + fractional "
                               "refinement (search_frac), 42900 PUs x 4 refs per step",
                   "candidates": "value counts the nominal window only, %d = PUs x 4225 x refs per step, for both arms; "
                                 "search_mv_full also visits the extra_mv / merge-candidate windows (%d windows of 4225 "
                                 "per PU here), which the reference arm re-evaluates where they overlap while the device "
                                 "evaluates each distinct displacement of a PU once (%d SADs per step), with "
                                 "bit-identical decisions"
                                 % (cands_per_frame(), sum(len(c) for c in pipe.centres),
                                    sum(pipe.union_area) * (32400 + 8040 + 1980 + 480)),
                   "frames_per_s": 1e3 / step_ms * world,
                   "l2": "per step %.0f MB of SAD maps are written and read back: larger than the 126 MB L2, "
                         "no flush needed" % (pipe.map_bytes / 1e6),
                   "stages": pipe.stages + (["publish_reference (NCCL broadcast of one luma plane) x1"] if world > 1 else []),
                   "parallelism": "frames of one temporal layer per GPU (weak scaling), no data-path collective in ME",
                   "kernel_ms": {"ime_sad x%d" % n_sad: sad_ms, "me_search (scan + refine x2)": search_ms,
                                 "share_sad": sad_ms / step_ms, "share_me_search": search_ms / step_ms}},
        "clocks": clk,
        "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": pipe.h2d_bytes, "d2h_bytes_per_step": pipe.d2h_bytes,
                "mode": "host buffers through the C ABI; each step copies all its inputs H2D and all its results D2H, "
                        "the copies of neighbouring steps overlap the kernels (two buffer sets, H2D / D2H streams)",
                "serial_ms_per_step": ms_e2e_serial / args.steps},
        "gpu_launches": int(launches),
        # the dense SAD kernel is bound by the integer (ALU) pipe; the denominator is
        # measured in this run by a register-only VABSDIFF4 loop (MEASURED_PEAKS.json has no integer figure).
        # achieved = the absolute differences the step's %d launches must evaluate / their summed duration.
        "roofline": {"kernel": kernel, "bound": "int_pipe",
                     "achieved": achieved_ops / 1e12, "peak": int_peak / 1e12, "unit": "Tabsdiff/s",
                     "frac": achieved_ops / int_peak,
                     "peak_source": "kvz_cuda_measure_int_pipe_peak: VABSDIFF4.U8.ACC register loop, measured in this run",
                     "algorithmic_abs_diffs": abs_diffs, "launches": n_sad, "ms_per_launch": sad_ms / n_sad,
                     "traffic": traffic, "traffic_source": pipe.NCU_TRAFFIC_SOURCE},
        "roofline_hbm": {"kernel": kernel, "bound": "hbm",
                         "achieved": map_bytes / (sad_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                         "frac": map_bytes / (sad_ms * 1e-3) / 1e9 / hbm_peak,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                         "traffic": traffic, "algorithmic_bytes": map_bytes},
    }
    # bounded CPU baseline on the same host, rank 0, N=1 only
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        fn, c, sample = reference_step_fn(frames, REF_SAMPLE, threads)
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": c / dt / 1e9, "unit": UNIT, "cores": threads, "kind": "reference",
                               "sample": sample + f", {dt:.1f} s"}
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--separate-windows", action="store_true",
                    help="one square-window SAD map set per window centre instead of one box per reference")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: whatever libraries write to file descriptor 1 while the bench runs (NCCL's
    # version banner, for one) goes to stderr instead; the result line is written to the real stdout at the end
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    lines = []
    global print
    builtin_print = print

    def print(*a, **k):                      # noqa: A001 -- the two result prints of this file
        lines.append(" ".join(str(x) for x in a))

    try:
        if args.impl == "reference":
            run_reference(args)
        else:
            run_ours(args)
    finally:
        print = builtin_print
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    for ln in lines:
        builtin_print(ln, flush=True)


if __name__ == "__main__":
    main()
