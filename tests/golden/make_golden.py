"""Generates tests/golden/ref_vectors.npz from the UNMODIFIED reference compiled into
oracle/_ref/libkvazaar_ref.so (oracle/Makefile; needs /root/reference, so this runs in the
build container only).  The vectors travel with the repo: tests/test_golden.py checks the
oracle (CPU) and the CUDA path (GPU box) against them without the reference being present.

    python tests/golden/make_golden.py

Inputs are seeded; every array the tests need (inputs and reference outputs) is stored.
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import kvzlibs  # noqa: E402
import me_oracle  # noqa: E402
from kvzlibs import ptr  # noqa: E402

W, H, RANGE, NREFS, LAMBDA, FME = 128, 72, 8, 2, 23, 4


def me_reference(L, pus, jobs, cur, refs, **cfg):
    """search_mv_full + search_frac / integer SATD of the reference per job, *inter_cost carried over
    the PU's reference loop (search_inter.c:1451-1524); merge match and select_mv_cand as
    search_inter.c:1258-1297."""
    from kvazaar_b200.frame import ME_RESULT_DT
    res = np.zeros(len(jobs), ME_RESULT_DT)
    for pu in pus:
        inter_cost = 0x7FFFFFFF
        merge = [(int(m["dir"]), int(m["ref"][0]), int(m["ref"][1]), int(m["mv"][0][0]), int(m["mv"][0][1]),
                  int(m["mv"][1][0]), int(m["mv"][1][1])) for m in pu["merge_cand"][:int(pu["num_merge_cand"])]]
        for q in range(int(pu["num_jobs"])):
            ji = int(pu["first_job"]) + q
            jb = jobs[ji]
            _, j = kvzlibs.job_pair(
                width=W, height=H, origin_x=int(pu["x"]), origin_y=int(pu["y"]), pu_w=int(pu["w"]), pu_h=int(pu["h"]),
                tile_off_x=0, tile_off_y=0, ref_idx=int(jb["ref_idx"]), lambda_sqrt=float(LAMBDA),
                mv_cand=jb["mv_cand"].tolist(), merge=merge, extra_mv=tuple(int(v) for v in jb["extra_mv"]),
                owf=cfg.get("owf", 0), wpp=cfg.get("wpp", 0), sao_type=3, deblock_enable=1,
                mv_constraint=cfg.get("mv_constraint", 0), max_ref_lcu_right=1, max_ref_lcu_down=1, fme_level=FME,
                me_early_termination=0, me_max_steps=-1, ime_algorithm=me_oracle.ALGO_OF_RANGE[RANGE])
            j.pic_y, j.ref_y = cur.ctypes.data, refs[int(jb["ref_idx"])].ctypes.data
            L.ref_shim_ime(C.byref(j))
            if FME > 0 and j.best_cost < inter_cost:
                L.ref_shim_fme(C.byref(j))
            elif j.best_cost < 0xFFFFFFFF:
                L.ref_shim_int_satd(C.byref(j))
            mvx, mvy = int(j.best_mv[0]), int(j.best_mv[1])
            merged, merge_idx = 0, 0
            for merge_idx, (d, r0, r1, a, b, c, e) in enumerate(merge):
                if d == 3:
                    continue
                mv = (a, b) if d == 1 else (c, e)
                rr = r0 if d == 1 else r1
                if mv == (mvx, mvy) and rr == int(jb["ref_idx"]):     # ref_LX is the identity here
                    merged = 1
                    break
            else:
                merge_idx = len(merge)
            cand = 0
            c0, c1 = jb["mv_cand"][0], jb["mv_cand"][1]
            if not merged and (int(c0[0]), int(c0[1])) != (int(c1[0]), int(c1[1])):
                def bits(v):
                    return 2 * (int(abs(v) + 2).bit_length() - 1)
                b0 = bits(mvx - int(c0[0])) + bits(mvy - int(c0[1]))
                b1 = bits(mvx - int(c1[0])) + bits(mvy - int(c1[1]))
                cand = 1 if b1 < b0 else 0
            took = j.best_cost < inter_cost
            if took:
                inter_cost = j.best_cost
            r = res[ji]
            r["mv"] = (mvx, mvy)
            r["cost"], r["bitcost"] = j.best_cost, j.best_bitcost
            r["merged"], r["merge_idx"], r["mv_cand_idx"], r["took"] = merged, merge_idx, cand, int(took)
    return res


def main():
    assert kvzlibs.have_ref(), "oracle/_ref/libkvazaar_ref.so missing: make -C oracle ref"
    L = kvzlibs.ref(1)
    out = {}
    rs = np.random.default_rng(2024)

    # --- motion search end to end (rows A3-A10), three restriction settings ---
    f = kvzlibs.synth_luma(31, W, H, NREFS + 1)
    cur, refs = f[NREFS], [f[NREFS - 1 - i] for i in range(NREFS)]
    pus, jobs = me_oracle.random_pus(rs, W, H, NREFS, 40, mv_span=RANGE)
    out["me_cur"], out["me_refs"] = cur, np.stack(refs)
    out["me_pus"], out["me_jobs"] = pus.view(np.uint8).reshape(-1), jobs.view(np.uint8).reshape(-1)
    for name, cfg in (("free", {}), ("owfwpp", {"owf": 1, "wpp": 1}), ("mvc4", {"mv_constraint": 4})):
        out["me_res_" + name] = me_reference(L, pus, jobs, cur, refs, **cfg).view(np.uint8).reshape(-1)

    # --- block costs (rows A1, A6, A7, A16) ---
    a = rs.integers(0, 256, (64, 64), dtype=np.uint8)
    b = np.clip(a.astype(np.int32) + rs.integers(-40, 41, (64, 64)), 0, 255).astype(np.uint8)
    out["cost_a"], out["cost_b"] = a, b
    sad, satd = [], []
    for n in (4, 8, 16, 32, 64):
        x = kvzlibs.aligned_copy(np.ascontiguousarray(a[:n, :n]))
        y = kvzlibs.aligned_copy(np.ascontiguousarray(b[:n, :n]))
        fs = kvzlibs.ref_strategy(f"kvz_sad_{n}x{n}", C.c_uint32, [C.c_void_p, C.c_void_p])
        ft = kvzlibs.ref_strategy(f"kvz_satd_{n}x{n}", C.c_uint32, [C.c_void_p, C.c_void_p])
        sad.append(fs(ptr(x), ptr(y)))
        satd.append(ft(ptr(x), ptr(y)))
    out["cost_sad_nxn"], out["cost_satd_nxn"] = np.array(sad, np.uint32), np.array(satd, np.uint32)

    # --- transforms and quantisation (rows A13, A14) ---
    for n, dst in ((4, 0), (4, 1), (8, 0), (16, 0), (32, 0)):
        blk = kvzlibs.aligned_copy(rs.integers(-255, 256, (n, n)).astype(np.int16))
        coef = kvzlibs.aligned((n, n), np.int16)
        back = kvzlibs.aligned((n, n), np.int16)
        L.ref_shim_transform2d(ptr(blk), ptr(coef), n, 0, dst, 0)
        L.ref_shim_transform2d(ptr(back), ptr(coef), n, 0, dst, 1)
        out[f"tr_{n}_{dst}_in"], out[f"tr_{n}_{dst}_coef"], out[f"tr_{n}_{dst}_back"] = blk.copy(), coef.copy(), back.copy()
        if True:
            rc = kvzlibs.RefQCfg(27, 0, 1, 0, 0, 0, 0.0)
            q = kvzlibs.aligned((n, n), np.int16)
            dq = kvzlibs.aligned((n, n), np.int16)
            L.ref_shim_quant(C.byref(rc), ptr(coef), ptr(q), n, 0, 0)
            L.ref_shim_dequant(C.byref(rc), ptr(q), ptr(dq), n, 0)
            out[f"tr_{n}_{dst}_q"], out[f"tr_{n}_{dst}_dq"] = q.copy(), dq.copy()

    np.savez_compressed(os.path.join(HERE, "ref_vectors.npz"), **out)
    print("wrote", os.path.join(HERE, "ref_vectors.npz"), {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
