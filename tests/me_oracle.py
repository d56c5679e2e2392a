"""Oracle-side restatement of the reference loop of search_pu_inter (search_inter.c:1451-1524)
on top of oracle/libkvz_oracle.so, plus random job generation shared by the GPU tests,
smoke() and the bench self-check.  TEST INFRASTRUCTURE ONLY."""
import ctypes as C

import numpy as np

import kvzlibs
from kvzlibs import ptr

ALGO_OF_RANGE = {8: 3, 16: 4, 32: 5, 64: 6}


def random_pus(rs, width, height, n_refs, count, sizes=(8, 16, 32, 64), mv_span=24, merge_prob=0.7,
               far_prob=0.05):
    """Random PU / job arrays in the layout of include/kvz_cuda.h (numpy structured arrays)."""
    from kvazaar_b200.frame import ME_JOB_DT, ME_PU_DT
    pus = np.zeros(count, ME_PU_DT)
    jobs = np.zeros(count * n_refs, ME_JOB_DT)
    for i in range(count):
        s = int(sizes[int(rs.integers(0, len(sizes)))])
        while s > width or s > height:
            s //= 2
        pus[i]["x"] = int(rs.integers(0, width // s)) * s
        pus[i]["y"] = int(rs.integers(0, height // s)) * s
        pus[i]["w"] = pus[i]["h"] = s
        nm = int(rs.integers(0, 6)) if rs.random() < merge_prob else 0
        pus[i]["num_merge_cand"] = nm
        for m in range(nm):
            mc = pus[i]["merge_cand"][m]
            mc["dir"] = int(rs.integers(1, 4))
            mc["ref"] = rs.integers(0, n_refs, 2)
            span = mv_span * 4 if rs.random() > far_prob else 1200
            mv = rs.integers(-span, span + 1, (2, 2))
            if rs.random() < 0.5:
                mv = (mv // 4) * 4            # integer vectors can match search positions exactly
            mc["mv"] = mv
        pus[i]["first_job"] = i * n_refs
        pus[i]["num_jobs"] = n_refs
        for r in range(n_refs):
            jb = jobs[i * n_refs + r]
            jb["ref_idx"] = r
            jb["mv_cand"] = rs.integers(-mv_span * 4, mv_span * 4 + 1, (2, 2))
            if rs.random() < 0.3:
                jb["mv_cand"][1] = jb["mv_cand"][0]
            jb["extra_mv"] = rs.integers(-mv_span * 4, mv_span * 4 + 1, 2) if rs.random() < 0.6 else 0
    return pus, jobs


def oracle_search(pus, jobs, cur, refs, width, height, rng, fme_level, lambda_sqrt_i, owf=0, wpp=0,
                  sao_type=3, deblock_enable=1, mv_constraint=0, max_ref_lcu=(1, 1), ref_LX=None,
                  ime_search=0, me_early_termination=0, me_max_steps=-1):
    """Runs kvzo_ime / kvzo_fme / kvzo_int_satd with *inter_cost carried over each PU's jobs
    and returns a result array in the layout of kvz_cuda_me_result."""
    from kvazaar_b200.frame import ME_RESULT_DT
    O = kvzlibs.oracle()
    res = np.zeros(len(jobs), ME_RESULT_DT)
    for pu in pus:
        inter_cost = 0x7FFFFFFF
        for q in range(int(pu["num_jobs"])):
            ji = int(pu["first_job"]) + q
            jb = jobs[ji]
            j = kvzlibs.OracleMeJob()
            j.width, j.height = width, height
            j.origin_x, j.origin_y, j.pu_w, j.pu_h = int(pu["x"]), int(pu["y"]), int(pu["w"]), int(pu["h"])
            for a in range(2):
                for b in range(2):
                    j.mv_cand[a][b] = int(jb["mv_cand"][a][b])
            j.num_merge_cand = int(pu["num_merge_cand"])
            for m in range(j.num_merge_cand):
                mc = pu["merge_cand"][m]
                j.merge_cand[m].dir = int(mc["dir"])
                j.merge_cand[m].ref[0], j.merge_cand[m].ref[1] = int(mc["ref"][0]), int(mc["ref"][1])
                for a in range(2):
                    for b in range(2):
                        j.merge_cand[m].mv[a][b] = int(mc["mv"][a][b])
            for l in range(2):
                for i in range(16):
                    j.ref_LX[l][i] = ref_LX[l][i] if ref_LX is not None else i
            j.ref_idx = int(jb["ref_idx"])
            j.extra_mv[0], j.extra_mv[1] = int(jb["extra_mv"][0]), int(jb["extra_mv"][1])
            j.lambda_sqrt_i = lambda_sqrt_i
            j.owf, j.wpp, j.sao_type, j.deblock_enable, j.mv_constraint = owf, wpp, sao_type, deblock_enable, mv_constraint
            j.max_ref_lcu_right, j.max_ref_lcu_down = max_ref_lcu
            j.fme_level, j.me_early_termination, j.me_max_steps = fme_level, me_early_termination, me_max_steps
            # kvz_cuda_ime_search: 0 full (by range), 1 hexbs (KVZ_IME_HEXBS = 0), 2 dia (KVZ_IME_DIA = 7), 3 tz (KVZ_IME_TZ = 1)
            j.ime_algorithm = (ALGO_OF_RANGE[rng], 0, 7, 1)[ime_search]
            ref = refs[j.ref_idx]
            O.kvzo_ime_8(C.byref(j), ptr(cur), ptr(ref))
            if fme_level > 0 and j.best_cost < inter_cost:
                O.kvzo_fme_8(C.byref(j), ptr(cur), ptr(ref))
            elif j.best_cost < 0xFFFFFFFF:
                O.kvzo_int_satd_8(C.byref(j), ptr(cur), ptr(ref))
            mvx, mvy = j.best_mv[0], j.best_mv[1]
            merged, merge_idx = 0, j.num_merge_cand
            for m in range(j.num_merge_cand):
                mc = j.merge_cand[m]
                if mc.dir != 3 and mc.mv[mc.dir - 1][0] == mvx and mc.mv[mc.dir - 1][1] == mvy and \
                        j.ref_LX[mc.dir - 1][mc.ref[mc.dir - 1]] == j.ref_idx:
                    merged, merge_idx = 1, m
                    break
            cand = 0 if merged else O.kvzo_select_mv_cand(j.mv_cand, mvx, mvy)
            took = j.best_cost < inter_cost
            if took:
                inter_cost = j.best_cost
            r = res[ji]
            r["mv"] = (mvx, mvy)
            r["cost"], r["bitcost"] = j.best_cost, j.best_bitcost
            r["merged"], r["merge_idx"], r["mv_cand_idx"], r["took"] = merged, merge_idx, cand, int(took)
    return res


def gpu_search(pus, jobs, cur, refs, width, height, rng, fme_level, lambda_sqrt_i, use_maps=True, centres=(), box=None,
               fused=False, pad=None, **cfg):
    """The CUDA path through the C ABI for the same inputs (numpy in, numpy out).
    centres: extra square-window map sets; box: (x0, y0, n_dx, n_dy) of one box map set per reference
    (jobs whose windows all lie inside it are searched from the box); fused: kvz_cuda_me_search_fused (one CTA per PU)."""
    import torch
    import kvazaar_b200 as kb
    from kvazaar_b200.frame import ME_RESULT_DT, MeSearch
    cur_d = torch.from_numpy(cur).cuda()
    max_c = max([0] + [max(abs(c[0]), abs(c[1])) for c in centres])
    ime = kb.ImeSearch(width, height, rng, max_center=max_c)
    bx = kb.ImeBoxSearch(width, height, box) if box is not None else None
    pad = pad if pad is not None else max(ime.pad, bx.pad if bx else 0)      # pad: explicit border (no dense maps then)
    planes, maps = [], []
    for r in refs:
        pp = kb.PaddedPlane(width, height, pad).fill_from(torch.from_numpy(r).cuda())
        planes.append(pp)
        if use_maps:
            sets = [((0, 0), [m.clone() for m in ime.run(cur_d, pp)])]
            if bx is not None:
                sets.insert(0, (bx.box, [m.clone() for m in bx.run(cur_d, pp)]))
            for c in centres:
                sets.append((tuple(c), [m.clone() for m in ime.run(cur_d, pp, center=tuple(c))]))
            maps.append(sets)
    ms = MeSearch(width, height, rng, fme_level=fme_level, lambda_sqrt=float(lambda_sqrt_i), **cfg)
    ms.set_current(cur_d)
    ms.set_refs(planes, maps if use_maps else None)
    d_pus = torch.from_numpy(pus.view(np.uint8).reshape(-1)).cuda()
    d_jobs = torch.from_numpy(jobs.view(np.uint8).reshape(-1)).cuda()
    d_res = torch.zeros(len(jobs) * ME_RESULT_DT.itemsize, dtype=torch.uint8, device="cuda")
    ms.run(d_pus, len(pus), d_jobs, d_res, fused=fused)
    torch.cuda.synchronize()
    return d_res.cpu().numpy().view(ME_RESULT_DT)


def assert_results_equal(got, want, pus=None):
    for name in ("mv", "cost", "bitcost", "merged", "merge_idx", "mv_cand_idx", "took"):
        if not np.array_equal(got[name], want[name]):
            bad = np.nonzero((got[name] != want[name]).reshape(len(got), -1).any(axis=1))[0]
            i = int(bad[0])
            raise AssertionError(f"{name} differs for {len(bad)} jobs; first job {i}: got {got[i]} want {want[i]}"
                                 + (f" pu {pus[i // int(pus[0]['num_jobs'])]}" if pus is not None else ""))
