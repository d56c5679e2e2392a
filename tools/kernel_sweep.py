"""BASELINE config 5: kernel sweep -- SAD / SATD / quarter-pel interpolation / DCT / DCT+quant at block sizes 4..64, the
batch kernels of libkvz_cuda.so on one B200 next to the reference's own generic and AVX2 strategies on ONE host core of the
same box (tests/speed_tests.c restated as oracle/ref_shim.c:ref_shim_speed; multi-core figures would be estimates and are
not printed).  GPU inputs are larger than the 126 MB L2 (64 Mi samples per operand); CUDA events on the launching stream.
Prints one JSON object; `frac_of_hbm` uses the algorithmic bytes of SURVEY.md 8(d) and MEASURED_PEAKS.json.
usage: python tools/kernel_sweep.py [--cpu-only]"""
import ctypes as C
import json
import os
import sys

import numpy as np

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
sys.path.insert(0, os.path.join(R, "tests"))
import kvzlibs  # noqa: E402

try:
    HBM = float(json.load(open(os.path.join(R, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0))
except Exception:
    HBM = 6650.0
KINDS = (("sad", 0, (4, 8, 16, 32, 64)), ("satd", 1, (4, 8, 16, 32, 64)), ("dct", 2, (4, 8, 16, 32)),
         ("idct", 3, (4, 8, 16, 32)), ("dct_quant", 4, (4, 8, 16, 32)), ("ipol_qpel_luma", 5, (8, 16, 32, 64)))
# algorithmic bytes per sample (8-bit pixels, int16 coefficients): two blocks read | in + out | in + out + levels | src + dst
BYTES = {"sad": 2, "satd": 2, "dct": 4, "idct": 4, "dct_quant": 6, "ipol_qpel_luma": 2}


def cpu_side(out):
    for cpuid, label in ((0, "generic"), (1, "avx2")):
        L = kvzlibs.ref(cpuid)
        L.ref_shim_speed.restype = C.c_double
        for name, what, sizes in KINDS:
            for n in sizes:
                blocks = (1 << 22) // (n * n)
                t = L.ref_shim_speed(what, n, blocks, 1)
                reps = max(1, int(0.25 / max(t, 1e-6)))
                t = L.ref_shim_speed(what, n, blocks, reps)
                out.setdefault(f"{name}_{n}", {})[f"cpu_{label}_Msamples_per_s_1core"] = round(blocks * n * n * reps / t / 1e6, 1)


def gpu_side(out):
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
    total = 1 << 26                                    # samples per operand: 64 MiB of pixels, 128 MiB of coefficients
    a = torch.from_numpy(rs.integers(0, 256, total, dtype=np.uint8)).cuda()
    b = torch.from_numpy(rs.integers(0, 256, total, dtype=np.uint8)).cuda()
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
            out.setdefault(f"{name}_{n}", {}).update(gpu_ms=round(ms, 3), gpu_blocks=cnt)
    coef_total = 1 << 25
    resid = torch.from_numpy(rs.integers(-60, 61, coef_total).astype(np.int16)).cuda()
    for n in (4, 8, 16, 32):
        cnt = coef_total // (n * n)
        ms = timed(lambda: blocks.transform_blocks(resid, n))
        out.setdefault(f"dct_{n}", {}).update(gpu_ms=round(ms, 3), gpu_blocks=cnt)
        coef = blocks.transform_blocks(resid, n)
        ms_i = timed(lambda: blocks.transform_blocks(coef, n, inverse=True))
        out.setdefault(f"idct_{n}", {}).update(gpu_ms=round(ms_i, 3), gpu_blocks=cnt)
        ms_q = timed(lambda: blocks.quant_blocks(coef, n, 32, signhide=True))
        out.setdefault(f"dct_quant_{n}", {}).update(gpu_ms=round(ms + ms_q, 3), gpu_blocks=cnt, note="two launches: transform, quant")
    # quarter-pel luma interpolation of every aligned n x n block of a 8192 x 8192 plane (64 Mi samples)
    side = 8192
    plane = a.view(side, side)
    for n in (8, 16, 32, 64):
        ys, xs = np.mgrid[0:side // n, 0:side // n]
        cnt = xs.size
        jobs = np.zeros(cnt, blocks.INTERP_JOB_DT)
        jobs["x"], jobs["y"] = xs.reshape(-1) * n, ys.reshape(-1) * n
        jobs["mv_x"], jobs["mv_y"] = 1, 3
        jobs["w"] = jobs["h"] = n
        jobs["dst_offset"] = np.arange(cnt, dtype=np.uint32) * n * n
        jobs["dst_stride"] = n
        blocks.keep_on_device(jobs)
        dst = torch.empty(side * side, dtype=torch.uint8, device="cuda")
        ms = timed(lambda: blocks.interp_blocks(plane, jobs, d_dst8=dst), 3)
        out.setdefault(f"ipol_qpel_luma_{n}", {}).update(gpu_ms=round(ms, 3), gpu_blocks=cnt)
    for key, rec in out.items():
        if "gpu_ms" not in rec:
            continue
        name, n = key.rsplit("_", 1)
        n = int(n)
        samples = rec["gpu_blocks"] * n * n
        rec["gpu_Msamples_per_s"] = round(samples / rec["gpu_ms"] / 1e3, 1)
        rec["gpu_GB_per_s"] = round(samples * BYTES[name] / rec["gpu_ms"] / 1e6, 1)
        rec["frac_of_hbm"] = round(samples * BYTES[name] / rec["gpu_ms"] / 1e6 / HBM, 3)
        for label in ("generic", "avx2"):
            c = rec.get(f"cpu_{label}_Msamples_per_s_1core")
            if c:
                rec[f"gpu_over_1core_{label}"] = round(rec["gpu_Msamples_per_s"] / c, 1)


def main():
    out = {}
    cpu_side(out)
    if "--cpu-only" not in sys.argv:
        gpu_side(out)
    print(json.dumps({"hbm_peak_gbs": HBM, "host_cores": os.cpu_count(),
                      "note": "CPU columns: ONE core, the reference's own strategies (generic = --no-cpuid); "
                              "GPU: one B200, batch kernels, inputs > L2", "kernels": out}))


if __name__ == "__main__":
    main()
