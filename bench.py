#!/usr/bin/env python
"""bench.py -- BASELINE.json's headline metric: 1080p encode frames/s of config 2
(`--preset medium --me full32 --gop 8 -r 4`, 8-bit, synthetic yuv420p of SURVEY.md 8d) through the encoder's own public
API (kvz_api: encoder_open / encoder_encode, src/kvazaar.h:545-707), with Kvazaar's inner motion loop on the B200.

GPU arm (default): kvazaar_b200/host/_build/libkvazaar_cuda.so -- the reference encoder with the CUDA path bound in
  (kvazaar_b200/host/, KVZ_CUDA_MODE=batch): per frame the source and the new reference picture go to the device, ONE
  launch produces the dense SAD maps of every PU against every reference, and every search_pu_inter of the host's CTU
  wavefront is one call of kvz_cuda_enc_search_pu (integer full search + fractional refinement + merge / AMVP choice).
  The bitstream equals the reference's (tests/test_gpu_encode_parity.py); its md5 is in the JSON line of both arms.
reference arm (--impl reference): oracle/_ref/libkvazaar_ref.so -- the UNMODIFIED reference, AVX2 strategies, same
  options, same clip, same frames, all host threads.

A step = one coded picture.  W warm-up pictures, then exactly K pictures timed (barrier + synchronize on both sides,
CUDA events, max over ranks).  `e2e` is the plain run: pictures fed from HOST memory, every step copies its source and
its new reference luma to the device and reads its results back.  `value` is the same run with the source luma of the
timed pictures preloaded into HBM before the timed region (kvz_cuda_enc_preload_source).  With N > 1 every rank encodes
its own clip on its own GPU with its share of the host threads (segment-parallel encoding: weak scaling, no data-path
collective); value = all ranks' pictures / the slowest rank's time.

The flat ME-stage benchmark of round 1 (Gcand/s over pre-made candidates) is tools/bench_me_stage.py.
Prints ONE JSON line.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

W, H = 1920, 1080
# BASELINE configs[1]; --owf is pinned (the bitstream depends on owf = 0 vs > 0 when WPP is on) to
# 3, which is what the reference's own `--owf auto` resolves to for 1080p on 16 or more threads (encoder.c:240-279): the
# reference arm runs as its CLI would by default (0.28 frames/s at --owf 0, 0.66 at 2, 0.72 at 4 on the 16-thread GPU
# box).  KVZ_BENCH_OWF overrides it for both arms.
OPTIONS = "--preset medium --me full32 --gop 8 -r 4 --owf " + os.environ.get("KVZ_BENCH_OWF", "3")
METRIC = "1080p encode fps (config 2: --preset medium --me full32 --gop 8 -r 4, 8-bit, bit-identical bitstream)"
UNIT = "frames/s"
N8 = (W // 8) * (H // 8)                                          # 8x8 blocks per picture
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libkvazaar_ref.so")


class Clocks:
    """nvidia-smi sampler (SM clock, throttle reasons) over the timed region"""
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
            threading.Thread(target=self._read, daemon=True).start()
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
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "note": "plain median of the samples taken every 100 ms over the timed regions; the GPU idles while the "
                        "host walks its CTUs, so low samples are idle clocks, not throttling"}


def make_frames(n, seed):
    import make_clip
    return list(make_clip.clip_planes(seed, W, H, n))


class Run:
    """Feeds a clip to one encoder and times the pictures [warmup, warmup + steps) as they come out."""

    def __init__(self, lib, frames, threads, warmup, steps):
        from kvazaar_b200 import encoder
        opts = encoder.options_from_cli(OPTIONS)
        opts["threads"] = str(threads)
        self.enc = encoder.Encoder(lib, W, H, opts)
        self.frames, self.warmup, self.steps = frames, warmup, steps
        self.fed, self.out = 0, 0
        self.md5 = hashlib.md5()
        self.bits = []

    def one_call(self):
        pic = None
        if self.fed < len(self.frames):
            pic = self.enc.picture(*self.frames[self.fed])
            self.fed += 1
        data, info = self.enc.encode(pic)
        if data is not None:
            self.md5.update(data)
            self.bits.append(len(data) * 8)
            self.out += 1
        elif pic is None:
            raise RuntimeError("encoder ran dry before the timed pictures were out")
        return data is not None

    def until(self, n_out):
        while self.out < n_out:
            self.one_call()

    def close(self):
        self.enc.close()


def run_reference(args):
    """the unmodified reference on the host cores: same options, clip and pictures; rank 0 only"""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    threads = os.cpu_count() or 1
    steps, warmup = args.steps, args.warmup
    frames = make_frames(warmup + steps + 9, seed=2)
    r = Run(REF_LIB, frames, threads, warmup, steps)
    r.until(warmup)
    t0 = time.perf_counter()
    r.until(warmup + steps)
    dt = time.perf_counter() - t0
    md5 = r.md5.hexdigest()
    r.close()
    v = steps / dt
    sample = f"pictures {warmup}..{warmup + steps - 1} in coding order of the 1080p clip (seed 2), {dt:.1f} s"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": dt * 1e3 / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": f"1920x1080 8-bit yuv420p, {OPTIONS} --threads {threads} (BASELINE configs[1])",
                   "api": "kvz_api of oracle/_ref/libkvazaar_ref.so: the unmodified reference, AVX2 strategies",
                   "bitstream_md5": md5, "md5_covers": f"the first {warmup + steps} pictures in coding order"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def cpu_baseline_sample(pictures=4, warmup=1):
    """the reference arm of this file in a child process on a bounded sample (rank 0, N = 1)"""
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", str(pictures),
                        "--warmup", str(warmup)], capture_output=True, text=True, timeout=1200, cwd=ROOT)
    for ln in r.stdout.splitlines():
        if ln.startswith("{"):
            d = json.loads(ln)
            cb = d["cpu_baseline"]
            cb["sample"] += " (the first pictures of a GOP have 1-3 references: cheaper than steady state)"
            cb["bitstream_md5"] = d["config"]["bitstream_md5"]
            return cb
    return {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference", "sample": "failed: " + r.stderr[-300:]}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from kvazaar_b200 import capi, encoder

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback on the hot path)")
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION") == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    os.environ["KVZ_CUDA_MODE"] = "batch"
    os.environ["KVZ_CUDA_DEVICE"] = str(local)
    threads = max(2, (os.cpu_count() or 2) // world)          # the ranks of one box share its host cores
    if os.environ.get("KVZ_BENCH_THREADS"):                  # experiments: worker threads per rank
        threads = int(os.environ["KVZ_BENCH_THREADS"])
    steps, warmup = args.steps, args.warmup
    if warmup < 2:
        raise SystemExit("bench.py: --warmup must be at least 2 (the first picture is intra: the session opens on the second)")
    frames = make_frames(warmup + steps + 9, seed=2 + rank)
    L = capi.lib()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_run(preload):
        """one full encode of the clip; returns (ms of the timed pictures, per-run record)"""
        r = Run(encoder.CUDA_LIB, frames, threads, warmup, steps)
        glue = r.enc.lib
        glue.kvz_glue_preload_source.argtypes = [C.c_int64, C.c_void_p, C.c_int]
        r.until(warmup)
        if preload:                                    # the source luma of every picture still to be fed, resident in HBM
            for i in range(r.fed, len(frames)):
                y = frames[i][0]
                if not glue.kvz_glue_preload_source(i, y.ctypes.data_as(C.c_void_p), y.strides[0]):
                    raise SystemExit("bench.py: kvz_glue_preload_source failed")
        m0, s0, c0 = (C.c_ulonglong * 8)(), (C.c_ulonglong * 8)(), (C.c_ulonglong * 4)()
        glue.kvz_glue_enc_measurements(m0)
        glue.kvz_glue_enc_stats(s0)
        glue.kvz_glue_spec_counts(c0)
        l0 = int(L.kvz_cuda_launch_count())
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        r.until(warmup + steps)
        e1.record()
        host_ms = (time.perf_counter() - t0) * 1e3
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        m1 = (C.c_ulonglong * 8)()
        glue.kvz_glue_enc_measurements(m1)
        s1, c1 = (C.c_ulonglong * 8)(), (C.c_ulonglong * 4)()
        glue.kvz_glue_enc_stats(s1)
        glue.kvz_glue_spec_counts(c1)
        rec = {"ms": float(ms.item()), "host_ms": host_ms, "launches": int(L.kvz_cuda_launch_count()) - l0,
               "m": [int(b) - int(a) for a, b in zip(m0, m1)], "pus_launched": int(s1[1]) - int(s0[1]),
               "answered_ahead": int(c1[0]) - int(c0[0]), "searched_on_the_spot": int(c1[1]) - int(c0[1]),
               "ref_jobs": int(s1[5]) - int(s0[5]), "jobs_outside_box": int(s1[6]) - int(s0[6]),
               "map_mib_last": int(s1[3]) / 2 ** 20, "md5": r.md5.hexdigest()}
        r.close()
        return rec

    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    e2e = timed_run(preload=False)                     # the plain run: everything from host memory
    res = timed_run(preload=True)                      # source luma of the timed pictures resident before the timed region
    clk = clocks.stop() if rank == 0 else None
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    fps = world * steps / (res["ms"] / 1e3)
    fps_e2e = world * steps / (e2e["ms"] / 1e3)
    m = e2e["m"]
    peak = capi.measure_int_pipe_peak()
    dense_ms = m[4] / 1e6 / max(1, m[3])
    achieved = (m[5] / max(1, m[3])) / (dense_ms / 1e3) / 1e12 if m[3] else 0.0
    hbm_peak = None
    try:
        hbm_peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    out = {
        "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": res["ms"] / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic", "gpu_launches": e2e["launches"],
        "config": {
            "workload": f"1920x1080 8-bit yuv420p, {OPTIONS} --threads {threads} (BASELINE configs[1]); a step = one coded picture",
            "api": "kvz_api (encoder_open / encoder_encode) of kvazaar_b200/host/_build/libkvazaar_cuda.so, KVZ_CUDA_MODE=batch",
            "bitstream_md5": e2e["md5"], "md5_covers": f"the first {warmup + steps} pictures in coding order of rank 0's clip (seed 2); "
                                                       "equal to the reference arm's = bit-identical",
            "bitstream_md5_preloaded_run": res["md5"],
            "host_threads_per_rank": threads,
            "value_vs_e2e": "value: source luma of the timed pictures preloaded into HBM (kvz_cuda_enc_preload_source); "
                            "e2e: every picture from host memory. Reconstructed reference pictures come from the host's own "
                            "in-loop filters in both runs.",
            "pu_searches_per_step": (e2e["answered_ahead"] + e2e["searched_on_the_spot"]) / steps,
            "share_of_searches_answered_by_a_launch_made_ahead": e2e["answered_ahead"] / max(1, e2e["answered_ahead"] + e2e["searched_on_the_spot"]),
            "pus_launched_per_step": e2e["pus_launched"] / steps,
            "reference_jobs_per_step": e2e["ref_jobs"] / steps,
            "share_of_jobs_with_a_window_outside_the_box": e2e["jobs_outside_box"] / max(1, e2e["ref_jobs"]),
            "sad_maps_mib_last_picture": e2e["map_mib_last"],
            "l2": "the SAD maps of one picture (several GB) exceed the 126 MB L2; no flush needed",
            "parallelism": "one encoder and one clip per rank (segment-parallel encoding), ranks share the host's cores; "
                           "no data-path collective",
            "device_stages_per_step": ["upload source + new reference luma, pad_plane", "ime_sad_multi_tma_kernel (dense SAD maps of all complete references in one launch; one launch per CTU for references still in flight)",
                                       "me_pu_kernel x (PU searches of the picture): box scan | CTA window scan, refinement"],
        },
        "roofline": {
            "kernel": "ime_sad_multi_tma_kernel (dense SAD maps of one picture against its references; reference tile rows staged by TMA)",
            "bound": "int_pipe", "achieved": achieved, "peak": peak / 1e12, "unit": "Tabsdiff/s",
            "frac": achieved / (peak / 1e12) if peak else None,
            "peak_source": "kvz_cuda_measure_int_pipe_peak: VABSDIFF4.U8.ACC register loop, measured in this run",
            "launches": m[3], "ms_per_launch": dense_ms,
            "algorithmic_abs_diffs_per_launch": m[5] / max(1, m[3]),
            "map_bytes_per_launch": m[6] / max(1, m[3]),
            "hbm_frac_of_map_writes": (m[6] / max(1, m[3])) / (dense_ms / 1e3) / 1e9 / hbm_peak if (hbm_peak and m[3]) else None,
            "traffic": None,
            "traffic_note": "no ncu capture of this launch inside the bench; see profiles/ for the round's captures",
            "share_of_step_ms": dense_ms / (e2e["ms"] / steps),
            # BASELINE's second metric, ME Gcand/s = (PU, reference, integer vector) costs produced per second: every map
            # entry is one; the 8x8 level has abs_diffs / 64 of them and the four levels together 85 / 64 of that
            "me_gcand_per_s_dense_kernel": (m[5] / 64 * 85 / 64) / (m[4] / 1e9) / 1e9 if m[4] else None,
            "me_gcand_per_s_in_encode": (m[5] / 64 * 85 / 64) / (e2e["ms"] / 1e3) / 1e9,
        },
        "clocks": clk,
        "e2e": {"value": fps_e2e, "unit": UNIT, "ms_per_step": e2e["ms"] / steps, "host_ms_per_step": e2e["host_ms"] / steps,
                "h2d_bytes_per_step": (m[0] + m[1]) / steps, "d2h_bytes_per_step": m[2] / steps,
                "mode": "pictures fed from host memory through kvz_api; per step: source luma + the new reference luma copied to "
                        "the device, PU / job records sent with each launch, results read back through mapped host memory"},
    }
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_sample()
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: whatever libraries write to file descriptor 1 while the bench runs goes to
    # stderr instead; the result line is written to the real stdout at the end
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    lines = []
    global print
    builtin_print = print

    def print(*a, **k):                      # noqa: A001 -- the result print of this file
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
