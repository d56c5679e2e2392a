"""Generates tests/golden/filter_vectors.npz: outputs of the UNMODIFIED reference compiled into
oracle/_ref/libkvazaar_ref.so for the rows built after the ME path (needs /root/reference: build container only):
  * kvz_filter_deblock_lcu over whole pictures (filter.c:770-779),
  * the SAO strategies (strategies-sao.h:37-60) and kvz_sao_reconstruct (sao.c:277-337),
  * kvz_get_coeff_cost on its CABAC path (rdo.c:158-222),
  * kvz_inter_get_merge_cand / kvz_inter_get_mv_cand (inter.c:1209-1446),
  * kvz_array_checksum (nal-generic.c:58-82).
tests/test_golden.py checks the oracle (CPU) and the CUDA path (GPU box) against these vectors without the
reference being present.

    python tests/golden/make_golden_filters.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import kvzlibs  # noqa: E402
import cand_oracle as co  # noqa: E402
import coeff_cases as cc  # noqa: E402
import deblock_oracle as do  # noqa: E402
import sao_oracle as so  # noqa: E402
from test_oracle_deblock import ref_deblock  # noqa: E402

DEBLOCK = [dict(size=(136, 72), qp=32), dict(size=(200, 136), qp=37, slice_b=1), dict(size=(136, 136), qp=27, cu_qp=1, qp_spread=5, beta=2, tc=-1)]
CAND = [dict(size=(200, 136)), dict(size=(200, 136), slice_b=True, n_refs=4)]


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def main():
    R = kvzlibs.ref(1)
    R.ref_shim_coeff_cost.restype = C.c_uint32
    out = {}
    # ---- deblocking ----
    for i, case in enumerate(DEBLOCK):
        w, h = case["size"]
        rs = np.random.default_rng(900 + i)
        cfg = do.make_cfg(w, h, qp=case["qp"], cu_qp=case.get("cu_qp", 0), beta=case.get("beta", 0), tc=case.get("tc", 0),
                          slice_b=case.get("slice_b", 0))
        scus = do.random_scus(rs, w, h, slice_b=bool(case.get("slice_b", 0)), qp=case["qp"], qp_spread=case.get("qp_spread", 0))
        planes = do.picture(rs, w, h)
        res = ref_deblock(cfg, scus, planes)
        out[f"db{i}_scus"] = scus.view(np.uint8).reshape(scus.shape + (20,))
        for k in range(3):
            out[f"db{i}_in{k}"], out[f"db{i}_out{k}"] = planes[k], res[k]
    # ---- SAO: block strategies + picture reconstruction ----
    rs = np.random.default_rng(910)
    # the generic strategy: the reference's AVX2 calc_sao_edge_dir (sao-avx2.c:157-245) assumes block widths that are
    # multiples of 8 and, for the 12-wide chroma CTUs of this picture, folds samples of the next row into its
    # 6-sample tail -- see DESIGN.md section 7
    ed = kvzlibs.ref_strategy("kvz_calc_sao_edge_dir", None, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)],
                              cpuid=0)
    R = kvzlibs.ref(1)
    W, H = 152, 104
    orig = [rs.integers(0, 256, s).astype(np.uint8) for s in ((H, W), (H // 2, W // 2), (H // 2, W // 2))]
    rec = [np.clip(o.astype(np.int32) + rs.integers(-9, 10, o.shape), 0, 255).astype(np.uint8) for o in orig]
    stats = []
    for k in range(3):
        for _, y0, x0, bh, bw in so.ctu_blocks(orig[k], 64 if k == 0 else 32):
            ob, rb = np.ascontiguousarray(orig[k][y0:y0 + bh, x0:x0 + bw]), np.ascontiguousarray(rec[k][y0:y0 + bh, x0:x0 + bw])
            e4 = []
            for e in range(4):
                acc = (C.c_int * 10)()
                ed(vp(ob), vp(rb), e, bw, bh, acc)
                e4.append(np.array(acc[:]).reshape(2, 5))
            stats.append(np.stack(e4))
    out["sao_orig0"], out["sao_orig1"], out["sao_orig2"] = orig
    out["sao_rec0"], out["sao_rec1"], out["sao_rec2"] = rec
    out["sao_edge_stats"] = np.stack(stats).astype(np.int32)
    luma, chroma = so.random_params(rs, 6), so.random_params(rs, 6)
    out["sao_luma_params"], out["sao_chroma_params"] = luma.view(np.uint8).reshape(-1, 16), chroma.view(np.uint8).reshape(-1, 16)
    for k in range(3):
        res = rec[k].copy()
        params = luma if k == 0 else chroma
        for i, y0, x0, bh, bw in so.ctu_blocks(rec[k], 64 if k == 0 else 32):
            s = so.info_of(params[i])
            R.ref_shim_sao_reconstruct(vp(rec[k]), vp(res), W, H, x0, y0, bw, bh, C.byref(s), k)
        out[f"sao_out{k}"] = res
    # ---- coefficient bit costs ----
    rs = np.random.default_rng(920)
    prm, ctxs, coefs, bits = [], [], [], []
    for (w, t, s) in cc.SHAPES:
        for i, blk in enumerate(cc.blocks(rs, w, 6)):
            ctx = kvzlibs.random_rdoq_ctx(rs)
            rng = int(rs.integers(256, 511))
            sh, ts = i % 2, (i // 2) % 2
            b = int(R.ref_shim_coeff_cost(C.byref(ctx), C.c_uint32(rng), vp(blk), w, t, s, sh, ts))
            prm.append([w, t, s, rng, sh, ts]); ctxs.append(np.frombuffer(bytes(ctx), np.uint8)); coefs.append(blk.reshape(-1)); bits.append(b)
    out["cb_params"], out["cb_ctx"], out["cb_coef"], out["cb_bits"] = np.array(prm, np.int32), np.stack(ctxs), np.concatenate(coefs), np.array(bits, np.uint32)
    # ---- merge / AMVP candidates ----
    for i, case in enumerate(CAND):
        kw = {k: v for k, v in case.items() if k != "size"}
        rs = np.random.default_rng(930 + i)
        sc = co.Scene(rs, case["size"][0], case["size"][1], **kw)
        f = sc.frame()
        pus = sc.pus(rs, 60)
        merges, amvps = [], []
        for pu in pus:
            m = np.zeros(5, co.MERGE_DT)
            R.ref_shim_merge_cand(C.byref(f), *pu, vp(m))
            merges.append(m.view(np.uint8).reshape(5, 12))
            per = []
            for l in range(2):
                for idx in range(4):
                    a = np.zeros((2, 2), np.int16)
                    if idx < len(sc.lists[l]):
                        R.ref_shim_mv_cand(C.byref(f), *pu[:4], l, idx, vp(a))
                    per.append(a)
            amvps.append(np.stack(per))
        out[f"cand{i}_scus"] = sc.scus.view(np.uint8).reshape(sc.scus.shape + (20,))
        out[f"cand{i}_ref_scus"] = np.stack([r.view(np.uint8).reshape(r.shape + (20,)) for r in sc.ref_scus])
        out[f"cand{i}_meta"] = np.array([sc.width, sc.height, sc.poc, int(sc.slice_b), int(sc.tmvp), sc.n_refs], np.int32)
        out[f"cand{i}_ref_pocs"] = np.array(sc.ref_pocs, np.int32)
        out[f"cand{i}_col_ref_pocs"] = np.array(sc.col_ref_pocs, np.int32)
        out[f"cand{i}_list0"], out[f"cand{i}_list1"] = np.array(sc.lists[0], np.int32), np.array(sc.lists[1], np.int32)
        out[f"cand{i}_pus"] = np.array(pus, np.int32)
        out[f"cand{i}_merge"], out[f"cand{i}_amvp"] = np.stack(merges), np.stack(amvps)
    # ---- picture checksum ----
    rs = np.random.default_rng(940)
    plane = rs.integers(0, 256, (270, 520), dtype=np.uint8)
    fn = kvzlibs.ref_strategy("kvz_array_checksum", None, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_ubyte), C.c_uint8])
    o = (C.c_ubyte * 16)()
    fn(vp(plane), 270, 520, 520, o, 8)
    out["ck_plane"], out["ck_sum"] = plane, np.array([(o[0] << 24) | (o[1] << 16) | (o[2] << 8) | o[3]], np.uint32)
    np.savez_compressed(os.path.join(HERE, "filter_vectors.npz"), **out)
    print(len(out), "arrays,", os.path.getsize(os.path.join(HERE, "filter_vectors.npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
