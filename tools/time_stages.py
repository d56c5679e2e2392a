"""Frame-scale device timings of the batch kernels beside the ME stage (rows A13-A17 of SURVEY.md 8a): one 1080p
frame's worth of blocks per launch, CUDA events on the launching stream, inputs resident.  Prints one JSON object;
bytes are the algorithmic bytes of DESIGN.md section 4.  Job / parameter arrays are uploaded once (blocks.keep_on_device), not per timed call."""
import json, os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np
import torch
import kvzlibs
from kvazaar_b200 import blocks

W, H = 1920, 1080
f = kvzlibs.synth_luma(2, W, H, 2)
d_src, d_pred = torch.from_numpy(f[1]).cuda(), torch.from_numpy(f[0]).cuda()
rs = np.random.default_rng(1)
try:
    HBM = float(json.load(open(os.path.join(R, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0))
except Exception:
    HBM = 6650.0


ONCE = os.environ.get("KVZ_STAGES_ONCE") is not None      # one launch per stage: the run ncu captures (tools/gpu_ncu_r02e.sh)


def timed(fn, iters=10):
    if ONCE:
        fn()
        torch.cuda.synchronize()
        return 1.0
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


out = {"frame": "1920x1080 8-bit, every aligned block of the size", "hbm_peak_gbs": HBM, "stages": {}}


def put(name, ms, n_blocks, bytes_, note=""):
    out["stages"][name] = {"blocks": n_blocks, "ms": round(ms, 4), "Mblocks_per_s": round(n_blocks / ms / 1e3, 1),
                           "algorithmic_GB_per_s": round(bytes_ / ms / 1e6, 1), "frac_of_hbm": round(bytes_ / ms / 1e6 / HBM, 3), "note": note}


for n in (4, 8, 16, 32):
    cnt = (W // n) * (H // n)
    resid = torch.from_numpy(rs.integers(-60, 61, cnt * n * n).astype(np.int16)).cuda()
    put(f"dct{n}", timed(lambda: blocks.transform_blocks(resid, n)), cnt, cnt * n * n * 4)
    coef = blocks.transform_blocks(resid, n)
    put(f"idct{n}", timed(lambda: blocks.transform_blocks(coef, n, inverse=True)), cnt, cnt * n * n * 4)
    put(f"quant{n}_signhide", timed(lambda: blocks.quant_blocks(coef, n, 32, signhide=True)), cnt, cnt * n * n * 4)
    # fused residual coding: scalar quantiser and RDOQ
    ys, xs = np.mgrid[0:H // n, 0:W // n]
    off = (ys.reshape(-1) * n * W + xs.reshape(-1) * n).astype(np.uint32)
    jobs = np.zeros(cnt, blocks.RESIDUAL_JOB_DT)
    jobs["ref_offset"] = jobs["pred_offset"] = jobs["rec_offset"] = off
    jobs["coeff_offset"] = np.arange(cnt, dtype=np.uint32) * n * n
    jobs["in_stride"] = jobs["out_stride"] = W
    jobs["width"], jobs["qp"], jobs["cu_is_intra"] = n, 32, 1
    blocks.keep_on_device(jobs)
    d_rec = torch.zeros((H, W), dtype=torch.uint8, device="cuda")
    d_co = torch.zeros(cnt * n * n, dtype=torch.int16, device="cuda")
    put(f"quantize_residual{n}", timed(lambda: blocks.quantize_residuals(d_src, d_pred, d_rec, d_co, jobs, True, True)), cnt,
        cnt * n * n * 5, "residual -> T -> Q -> Q^-1 -> T^-1 -> recon, one launch")
    rj = np.zeros(cnt, blocks.RDOQ_JOB_DT)
    rj["coef_offset"] = jobs["coeff_offset"]; rj["ctx_index"] = np.arange(cnt) % 64; rj["lambda"] = 40.0
    rj["width"], rj["qp"], rj["cu_is_intra"] = n, 32, 1
    ctxs = np.frombuffer(rs.integers(0, 128, 64 * 148, dtype=np.uint8).tobytes(), blocks.RDOQ_CTX_DT).copy()
    blocks.keep_on_device(rj); blocks.keep_on_device(ctxs)
    put(f"quantize_residual{n}_rdoq", timed(lambda: blocks.quantize_residuals_rdoq(d_src, d_pred, d_rec, d_co, jobs, ctxs, rj, True, True)),
        cnt, cnt * n * n * 9, "transform | kvz_rdoq | reconstruction, three launches (latency-bound serial chain per block)")

# intra: 35 predictions + SATD/SAD per PU
for lg in (2, 3, 4, 5):
    n = 1 << lg
    cnt = (W // n) * (H // n)
    refs = rs.integers(0, 256, cnt * 2 * (2 * n + 1), dtype=np.uint8)
    ys, xs = np.mgrid[0:H // n, 0:W // n]
    jobs = np.zeros(cnt, blocks.INTRA_JOB_DT)
    jobs["ref_offset"] = np.arange(cnt, dtype=np.uint32) * 2 * (2 * n + 1)
    jobs["orig_offset"] = (ys.reshape(-1) * n * W + xs.reshape(-1) * n).astype(np.uint32)
    jobs["orig_stride"] = W
    jobs["log2_width"], jobs["filter_boundary"] = lg, 1
    jobs[jobs.dtype.names[-1]] = 1
    d_refs = torch.from_numpy(refs).cuda()
    blocks.keep_on_device(jobs)
    ms = timed(lambda: blocks.intra_mode_costs(d_refs, d_src, jobs), 5)
    out["stages"][f"intra_costs{n}"] = {"blocks": cnt, "ms": round(ms, 4), "Mblocks_per_s": round(cnt / ms / 1e3, 1),
                                        "predicted_Gsamples_per_s": round(cnt * 35 * n * n / ms / 1e6, 1),
                                        "note": "35 predictions + SATD + SAD per PU (ALU / latency bound)"}
# CABAC-exact coefficient bit cost: every block of the frame, quantised on device first
for n in (4, 8, 16, 32):
    cnt = (W // n) * (H // n)
    resid = torch.from_numpy(rs.integers(-40, 41, cnt * n * n).astype(np.int16)).cuda()
    lv = blocks.quant_blocks(blocks.transform_blocks(resid, n), n, 32, signhide=True)
    cj = np.zeros(cnt, blocks.COEFF_JOB_DT)
    cj["coef_offset"] = np.arange(cnt, dtype=np.uint32) * n * n
    cj["ctx_index"] = np.arange(cnt) % 64
    cj["range"], cj["width"] = 400, n
    cctx = blocks._dev(np.frombuffer(rs.integers(0, 126, 64 * 148, dtype=np.uint8).tobytes(), blocks.RDOQ_CTX_DT).copy())
    d_cj = blocks._dev(cj)
    put(f"coeff_bits{n}", timed(lambda: blocks.coeff_bits(lv, cctx, cj, True, False, d_jobs=d_cj)), cnt, cnt * n * n * 2,
        "kvz_get_coeff_cost CABAC path, one warp per block (serial context chain per block)")

# picture-level reductions: 4:2:0 pictures, all three planes in one launch
for name, (pw, ph), bd in (("1080p_8bit", (1920, 1080), 8), ("2160p_10bit", (3840, 2160), 10)):
    dt, top = (np.uint8, 256) if bd == 8 else (np.int16, 1024)
    mk = lambda: [torch.from_numpy(rs.integers(0, top, s).astype(dt)).cuda() for s in ((ph, pw), (ph // 2, pw // 2), (ph // 2, pw // 2))]
    a, b = mk(), mk()
    nbytes = sum(t.numel() * t.element_size() for t in a)
    for label, fn, rd in (("picture_checksum", lambda: blocks.picture_checksum(a, bd), 1), ("picture_sse", lambda: blocks.picture_sse(a, b, bd), 2)):
        ms = timed(fn, 20)
        out["stages"][f"{label}_{name}"] = {"ms": round(ms, 4), "algorithmic_GB_per_s": round(rd * nbytes / ms / 1e6, 1),
                                            "frac_of_hbm": round(rd * nbytes / ms / 1e6 / HBM, 3),
                                            "note": "one launch + one memset per picture; the planes fit in L2, so this is launch latency, not HBM"}
    # SAO: statistics pass and reconstruction of the whole picture
    st_fn = lambda: blocks.sao_ctu_stats(a, b, bd)
    ms = timed(st_fn, 20)
    out["stages"][f"sao_ctu_stats_{name}"] = {"ms": round(ms, 4), "algorithmic_GB_per_s": round(2 * nbytes / ms / 1e6, 1),
                                              "frac_of_hbm": round(2 * nbytes / ms / 1e6 / HBM, 3),
                                              "note": "all 4 edge classes + 32 bands per CTU and plane, one launch"}
    counts = blocks.sao_ctu_counts(a)
    prm = np.zeros(counts[0], blocks.SAO_PARAMS_DT); prm["type"] = 2; prm["eo_class"] = np.arange(counts[0]) % 4
    prm["offsets"] = (0, 2, 1, -1, -2, 0, 2, 1, -1, -2)
    prc = prm[:counts[1]].copy()
    d_l, d_c = blocks._dev(prm), blocks._dev(prc)
    outp = [torch.empty_like(t) for t in a]
    ms = timed(lambda: blocks.sao_reconstruct(b, outp, d_l, d_c, bd), 20)
    out["stages"][f"sao_reconstruct_{name}"] = {"ms": round(ms, 4), "algorithmic_GB_per_s": round(2 * nbytes / ms / 1e6, 1),
                                                "frac_of_hbm": round(2 * nbytes / ms / 1e6 / HBM, 3),
                                                "note": "edge offsets in every CTU, one launch"}
# inter candidate derivation (row A18): every 2Nx2N PU of the frame x 4 references
import cand_oracle as cdo
from kvazaar_b200.frame import ME_PU_DT, ME_JOB_DT, inter_candidates
sc = cdo.Scene(rs, W, H - H % 8, slice_b=False, n_refs=4)
pl_ = [(x, y, sz) for sz in (8, 16, 32, 64) for y in range(0, sc.height - sz + 1, sz) for x in range(0, sc.width - sz + 1, sz)]
P = np.zeros(len(pl_), ME_PU_DT); J = np.zeros(len(pl_) * 4, ME_JOB_DT)
P["x"], P["y"] = [p[0] for p in pl_], [p[1] for p in pl_]
P["w"] = P["h"] = [p[2] for p in pl_]
P["first_job"], P["num_jobs"] = np.arange(len(pl_)) * 4, 4
J["ref_idx"] = np.tile(np.arange(4), len(pl_))
d_P, d_J, d_S = blocks._dev(P), blocks._dev(J), blocks._dev(sc.scus)
d_R = [blocks._dev(r) for r in sc.ref_scus]
ms = timed(lambda: inter_candidates(sc, d_S, d_R, d_P, len(P), d_J), 20)
out["stages"]["inter_candidates_1080p"] = {"pus": len(P), "jobs": len(J), "ms": round(ms, 4), "Mpus_per_s": round(len(P) / ms / 1e3, 1),
                                           "note": "merge list + AMVP x4 references + colocated start vector per PU, one thread per PU"}

# deblocking: whole 4:2:0 pictures, random CU quadtrees, two launches per picture
import deblock_oracle as dbo
for name, (pw, ph), bd in (("1080p_8bit", (1920, 1080), 8), ("2160p_10bit", (3840, 2160), 10)):
    scus = blocks._dev(dbo.random_scus(rs, pw, ph, p_intra=0.3).view(blocks.SCU_DT))
    pl = [torch.from_numpy(p.view(np.int16) if bd > 8 else p).cuda() for p in dbo.picture(rs, pw, ph, bd)]
    nbytes = sum(t.numel() * t.element_size() for t in pl)
    ms = timed(lambda: blocks.deblock_picture(pl, scus, 32, bitdepth=bd), 20)
    out["stages"][f"deblock_picture_{name}"] = {"ms": round(ms, 4), "algorithmic_GB_per_s": round(2 * nbytes / ms / 1e6, 1),
                                                "frac_of_hbm": round(2 * nbytes / ms / 1e6 / HBM, 3),
                                                "note": "vertical-edge pass + horizontal-edge pass, in place; one thread per 4-sample edge segment"}
print(json.dumps(out))
