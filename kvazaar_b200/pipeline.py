"""Frame-level motion-estimation pipeline over the batch face of the C ABI.

This is the host-side mirror of what a patched encoderstate.c / search_inter.c would do
per frame: keep the current and the reference luma planes in HBM, pad the
references, produce the dense SAD maps per reference, then run the per-PU decision stage.
torch supplies device memory, pinned host buffers, streams and events only.
"""
import ctypes as C

import numpy as np
import torch

from . import capi
from .frame import ME_JOB_DT, ME_PU_DT, ME_RESULT_DT, ImeBoxSearch, ImeSearch, MeSearch, PaddedPlane


def frame_pus(width, height, n_refs, true_mv_px=(5, 3), sizes=(64, 32, 16, 8)):
    """Every 2Nx2N PU (64..8) fully inside the picture, largest first, with the candidates a
    steady-state encode of a globally translating clip produces: merge candidate 0 and AMVP
    candidate 0 are the neighbours' vector (the true motion towards that reference), merge
    candidate 1 / AMVP candidate 1 the zero vector, and extra_mv the colocated CU's vector
    (search_inter.c:1188-1206).  true_mv_px is the motion per frame of temporal distance."""
    n = sum((width // s) * (height // s) for s in sizes)
    pus = np.zeros(n, ME_PU_DT)
    jobs = np.zeros(n * n_refs, ME_JOB_DT)
    i = 0
    for s in sizes:
        ys, xs = np.meshgrid(np.arange(height // s) * s, np.arange(width // s) * s, indexing="ij")
        m = xs.size
        pus["x"][i:i + m], pus["y"][i:i + m] = xs.ravel(), ys.ravel()
        pus["w"][i:i + m] = pus["h"][i:i + m] = s
        i += m
    pus["num_merge_cand"] = 2
    pus["num_jobs"] = n_refs
    pus["first_job"] = np.arange(n, dtype=np.uint32) * n_refs
    pus["merge_cand"]["dir"][:, 0] = 1
    pus["merge_cand"]["mv"][:, 0, 0, 0] = true_mv_px[0] * 4
    pus["merge_cand"]["mv"][:, 0, 0, 1] = true_mv_px[1] * 4
    pus["merge_cand"]["dir"][:, 1] = 1
    j = jobs.reshape(n, n_refs)
    for r in range(n_refs):
        j["ref_idx"][:, r] = r
        j["mv_cand"][:, r, 0, 0] = true_mv_px[0] * 4 * (r + 1)
        j["mv_cand"][:, r, 0, 1] = true_mv_px[1] * 4 * (r + 1)
        j["extra_mv"][:, r, 0] = true_mv_px[0] * 4 * (r + 1)
        j["extra_mv"][:, r, 1] = true_mv_px[1] * 4 * (r + 1)
    return pus, jobs


def window_centres(pus, jobs, n_refs, max_sets=3):
    """The most frequent non-zero window centres search_mv_full will visit per reference
    (search_inter.c:898-955): extra_mv >> 2 and the uni-directional merge candidates >> 2.
    Dense maps are made for these; any other window is evaluated on demand by me_search."""
    out = []
    jj = jobs.reshape(len(pus), -1)
    for r in range(n_refs):
        cnt = {}
        ex = jj["extra_mv"][:, r].astype(np.int64) >> 2
        for c, n in zip(*np.unique(ex, axis=0, return_counts=True)):
            cnt[tuple(int(v) for v in c)] = cnt.get(tuple(int(v) for v in c), 0) + int(n)
        for m in range(5):
            sel = (pus["num_merge_cand"] > m) & (pus["merge_cand"]["dir"][:, m] != 3)
            if not sel.any():
                continue
            d = pus["merge_cand"]["dir"][sel, m].astype(np.int64) - 1
            mv = pus["merge_cand"]["mv"][sel, m][np.arange(int(sel.sum())), d].astype(np.int64) >> 2
            for c, n in zip(*np.unique(mv, axis=0, return_counts=True)):
                cnt[tuple(int(v) for v in c)] = cnt.get(tuple(int(v) for v in c), 0) + int(n)
        cnt.pop((0, 0), None)
        out.append([c for c, _ in sorted(cnt.items(), key=lambda kv: -kv[1])[:max_sets]])
    return out


class MePipeline:
    # where the DRAM traffic of the dense SAD launch of one step was last measured (ncu --set full); bench lines name the
    # file instead of repeating a number as if it were measured in their run
    NCU_TRAFFIC_SOURCE = "profiles/r01e_prof_sad_summary.txt: 17.9 MB read + 2082.6 MB written for the four r01 boxes"

    def __init__(self, width, height, rng, n_refs, lambda_sqrt=27.0, fme_level=4, device="cuda", use_boxes=True):
        """use_boxes: one box of displacements per reference that covers the zero window and the frequent
        extra_mv / merge windows (each displacement computed and priced once); False: one square-window
        map set per window centre, as the reference visits them."""
        capi.init(torch.cuda.current_device())
        self.width, self.height, self.range, self.n_refs = width, height, rng, n_refs
        self.lambda_sqrt, self.fme_level = lambda_sqrt, fme_level
        pus, jobs = frame_pus(width, height, n_refs)
        self.centres = [[(0, 0)] + c for c in window_centres(pus, jobs, n_refs)]
        max_c = max(max(abs(c[0]), abs(c[1])) for cs in self.centres for c in cs)
        self.ime = ImeSearch(width, height, rng, device=device, max_center=max_c)
        self.cur = torch.empty((height, width), dtype=torch.uint8, device=device)
        self.refs = [torch.empty((height, width), dtype=torch.uint8, device=device) for _ in range(n_refs)]
        self.use_boxes = use_boxes
        self.multi_launch = True        # the boxes of all references in one launch (kvz_cuda_ime_sad_box_maps_multi)
        k = 2 * rng + 1
        n_pu = [(width // (8 << l)) * (height // (8 << l)) for l in range(4)]
        self.union_area = []
        if use_boxes:
            self.boxes = []
            for cs in self.centres:
                x0, x1 = min(c[0] for c in cs) - rng, max(c[0] for c in cs) + rng
                y0, y1 = min(c[1] for c in cs) - rng, max(c[1] for c in cs) + rng
                self.boxes.append(ImeBoxSearch(width, height, (x0, y0, x1 - x0 + 1, y1 - y0 + 1), device=device))
                cover = np.zeros((y1 - y0 + 1, x1 - x0 + 1), bool)
                for c in cs:
                    cover[c[1] - rng - y0:c[1] + rng + 1 - y0, c[0] - rng - x0:c[0] + rng + 1 - x0] = True
                self.union_area.append(int(cover.sum()))
            pad = max(self.ime.pad, max(b.pad for b in self.boxes))
            self.padded = [PaddedPlane(width, height, pad, device=device) for _ in range(n_refs)]
            self.maps = [[(b.box, b.maps)] for b in self.boxes]
            self.n_sad_launches = (n_refs + 7) // 8          # all references in one launch (up to 8 per launch)
            sad_stage = ("ime_sad_box_maps_multi: %d boxes in one launch (one box per reference over the zero window and the frequent extra_mv / "
                         "merge windows: %s displacements)" % (n_refs, "+".join("%dx%d" % (b.box[2], b.box[3]) for b in self.boxes)))
        else:
            self.padded = [PaddedPlane(width, height, self.ime.pad, device=device) for _ in range(n_refs)]
            self.maps = [[(c, [torch.empty_like(m) for m in self.ime.maps]) for c in self.centres[r]]
                         for r in range(n_refs)]
            self.union_area = [k * k * len(cs) for cs in self.centres]
            self.n_sad_launches = sum(len(c) for c in self.centres)
            sad_stage = "ime_sad_maps x%d (zero window + the frequent extra_mv / merge windows per reference)" % self.n_sad_launches
        # 64 absolute differences per 8x8 block and displacement that has to be evaluated
        self.algorithmic_abs_diffs = sum(n_pu[0] * 64 * a for a in self.union_area)
        # the maps of those displacements written once (read back once by me_search) + the planes read once
        self.algorithmic_map_bytes = sum(a * (2 * n_pu[0] + 2 * n_pu[1] + 4 * n_pu[2] + 4 * n_pu[3]) for a in self.union_area)
        self.map_bytes = sum(m.numel() * m.element_size() for sets in self.maps for _, ms in sets for m in ms)
        self.search = MeSearch(width, height, rng, fme_level=fme_level, lambda_sqrt=lambda_sqrt)
        self.search.set_current(self.cur)
        self.search.set_refs(self.padded, self.maps)
        self.stages = ["pad_plane x%d" % n_refs, sad_stage,
                       "me_search x1 = me_scan (integer decision) + me_refine x2 (fractional search, PUs <= 16 and > 16)"]
        # PUs / jobs of the frame (host, pinned) and their device copies
        self.n_pus, self.n_jobs = len(pus), len(jobs)
        self.host_pus = torch.from_numpy(pus.view(np.uint8).reshape(-1)).pin_memory()
        self.host_jobs = torch.from_numpy(jobs.view(np.uint8).reshape(-1)).pin_memory()
        self.d_pus = self.host_pus.to(device)
        self.d_jobs = self.host_jobs.to(device)
        self.d_results = torch.zeros(self.n_jobs * ME_RESULT_DT.itemsize, dtype=torch.uint8, device=device)
        self.host_results = torch.zeros(self.n_jobs * ME_RESULT_DT.itemsize, dtype=torch.uint8).pin_memory()
        self.h2d_bytes = (1 + n_refs) * width * height + self.host_pus.numel() + self.host_jobs.numel()
        self.d2h_bytes = self.host_results.numel()

    # ---- data movement ----
    def upload(self, host_cur, host_refs):
        self.cur.copy_(host_cur, non_blocking=True)
        for d, h in zip(self.refs, host_refs):
            d.copy_(h, non_blocking=True)

    # ---- one frame ----
    def step_resident(self):
        for r in range(self.n_refs):
            self.padded[r].fill_from(self.refs[r])
        self.sad_maps()
        self.search.run(self.d_pus, self.n_pus, self.d_jobs, self.d_results)

    def sad_maps(self):
        """every dense SAD launch of one frame (planes already padded)"""
        if self.use_boxes and self.multi_launch:
            ImeBoxSearch.run_many(self.boxes, self.cur, self.padded)      # all references in one launch
            return
        for r in range(self.n_refs):
            if self.use_boxes:
                self.boxes[r].run(self.cur, self.padded[r])
            else:
                for c, m in self.maps[r]:
                    self.ime.run(self.cur, self.padded[r], center=c, maps=m)

    def step_from_host(self, host_cur, host_refs):
        self.upload(host_cur, host_refs)
        self.d_pus.copy_(self.host_pus, non_blocking=True)
        self.d_jobs.copy_(self.host_jobs, non_blocking=True)
        self.step_resident()
        self.host_results.copy_(self.d_results, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    # ---- the same step as a software pipeline over two input/result buffer sets ----
    def _make_sets(self):
        dev = self.cur.device
        self._sets = []
        for i in range(2):
            st = {
                "cur": self.cur if i == 0 else torch.empty_like(self.cur),
                "refs": self.refs if i == 0 else [torch.empty_like(r) for r in self.refs],
                "pus": self.d_pus if i == 0 else torch.empty_like(self.d_pus),
                "jobs": self.d_jobs if i == 0 else torch.empty_like(self.d_jobs),
                "res": self.d_results if i == 0 else torch.empty_like(self.d_results),
                "host_res": self.host_results if i == 0 else torch.zeros_like(self.host_results).pin_memory(),
                "uploaded": torch.cuda.Event(), "computed": torch.cuda.Event(), "downloaded": torch.cuda.Event(),
                "used": False,
            }
            self._sets.append(st)
        self._h2d, self._d2h = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        self._k = 0

    def step_from_host_pipelined(self, host_cur, host_refs):
        """One step from HOST buffers, overlapped with its neighbours: the inputs of step k are copied on an
        H2D stream while step k-1 computes, its results come back on a D2H stream while step k+1 computes.
        Every step still uploads all its inputs and downloads all its results; `flush_pipelined()` (or a device
        synchronize) completes the last one.  Returns the host results of the step before this one, or None."""
        if not hasattr(self, "_sets"):
            self._make_sets()
        st, prev = self._sets[self._k & 1], self._sets[(self._k & 1) ^ 1]
        self._k += 1
        comp = torch.cuda.current_stream()
        with torch.cuda.stream(self._h2d):
            if st["used"]:
                self._h2d.wait_event(st["computed"])        # the set's previous step has finished reading it
            st["cur"].copy_(host_cur, non_blocking=True)
            for d, h in zip(st["refs"], host_refs):
                d.copy_(h, non_blocking=True)
            st["pus"].copy_(self.host_pus, non_blocking=True)
            st["jobs"].copy_(self.host_jobs, non_blocking=True)
            st["uploaded"].record(self._h2d)
        comp.wait_event(st["uploaded"])
        if st["used"]:
            comp.wait_event(st["downloaded"])               # its previous results have left the device
        for r in range(self.n_refs):
            self.padded[r].fill_from(st["refs"][r])
        if self.use_boxes and self.multi_launch:
            ImeBoxSearch.run_many(self.boxes, st["cur"], self.padded)
        else:
            cur_saved, self.cur = self.cur, st["cur"]
            self.sad_maps()
            self.cur = cur_saved
        self.search.set_current(st["cur"])
        self.search.run(st["pus"], self.n_pus, st["jobs"], st["res"])
        st["computed"].record(comp)
        with torch.cuda.stream(self._d2h):
            self._d2h.wait_event(st["computed"])
            st["host_res"].copy_(st["res"], non_blocking=True)
            st["downloaded"].record(self._d2h)
        st["used"] = True
        if prev["used"]:
            prev["downloaded"].synchronize()
            return prev["host_res"]
        return None

    def flush_pipelined(self):
        """host results of the last pipelined step"""
        st = self._sets[(self._k - 1) & 1]
        st["downloaded"].synchronize()
        self.search.set_current(self.cur)
        return st["host_res"]

    def results(self):
        """kvz_cuda_me_result per job, as a numpy structured array (after step_from_host)"""
        return self.host_results.numpy().view(ME_RESULT_DT)

    def time_stage(self, fn, iters=5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    def time_me_search(self, iters=5):
        return self.time_stage(lambda: self.search.run(self.d_pus, self.n_pus, self.d_jobs, self.d_results), iters)

    # ---- the dense SAD launches of one frame alone ----
    def time_sad_kernel(self, iters=10):
        """ms for all SAD launches of one frame (n_sad_launches of them), planes resident and padded"""
        for r in range(self.n_refs):
            self.padded[r].fill_from(self.refs[r])
        return self.time_stage(self.sad_maps, iters)

