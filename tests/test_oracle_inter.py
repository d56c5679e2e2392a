"""Pins the oracle's interpolation and motion-search restatement (SURVEY.md 8a rows A3, A4,
A5, A9, A10, A11) on the compiled reference: the FME block filters, the final MC samplers,
calc_mvd_cost, fracmv_within_tile, every IME algorithm and search_frac."""
import ctypes as C

import numpy as np
import pytest

import kvzlibs
from kvzlibs import ptr

W, H = 160, 96


@pytest.fixture(scope="module")
def clip():
    return kvzlibs.synth_luma(21, W, H, 3)


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_mc_luma_chroma_vs_reference(oracle_lib, ref_lib, clip, cpuid):
    L = kvzlibs.ref(cpuid)
    ref = clip[0]
    rs = np.random.default_rng(31)
    for (w, h) in [(8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (8, 16), (4, 4), (4, 8)]:
        for _ in range(10):
            x = int(rs.integers(0, (W - w) // 4 + 1)) * 4
            y = int(rs.integers(0, (H - h) // 4 + 1)) * 4
            mvx, mvy = int(rs.integers(-120, 121)), int(rs.integers(-120, 121))
            if w >= 8:
                want = np.zeros((h, w), np.uint8); got = np.zeros((h, w), np.uint8)
                L.ref_shim_mc_luma(ptr(ref), W, H, x, y, mvx, mvy, w, h, ptr(want), w, None)
                oracle_lib.kvzo_interp_luma_8(ptr(ref), W, H, x, y, mvx, mvy, w, h, ptr(got), w)
                assert np.array_equal(got, want), ("luma", w, h, x, y, mvx, mvy)
                want14 = np.zeros((h, w), np.int16); got14 = np.zeros((h, w), np.int16)
                L.ref_shim_mc_luma(ptr(ref), W, H, x, y, mvx, mvy, w, h, None, w, ptr(want14))
                oracle_lib.kvzo_interp_luma14_8(ptr(ref), W, H, x, y, mvx, mvy, w, h, ptr(got14), w)
                assert np.array_equal(got14, want14), ("luma14", w, h, x, y, mvx, mvy)
            if w > 32 or h > 32:
                continue   # chroma blocks are at most 32x32 (LCU_WIDTH_C)
            # chroma: the same plane used as a chroma plane, 1/8-pel vectors
            want = np.zeros((h, w), np.uint8); got = np.zeros((h, w), np.uint8)
            L.ref_shim_mc_chroma(ptr(ref), W, H, x, y, mvx, mvy, w, h, ptr(want), w, None)
            oracle_lib.kvzo_interp_chroma_8(ptr(ref), W, H, x, y, mvx, mvy, w, h, ptr(got), w)
            assert np.array_equal(got, want), ("chroma", w, h, x, y, mvx, mvy)
            want14 = np.zeros((h, w), np.int16); got14 = np.zeros((h, w), np.int16)
            L.ref_shim_mc_chroma(ptr(ref), W, H, x, y, mvx, mvy, w, h, None, w, ptr(want14))
            oracle_lib.kvzo_interp_chroma14_8(ptr(ref), W, H, x, y, mvx, mvy, w, h, ptr(got14), w)
            assert np.array_equal(got14, want14), ("chroma14", w, h, x, y, mvx, mvy)


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_fme_filter_blocks_vs_reference(oracle_lib, ref_lib, clip, cpuid):
    L = kvzlibs.ref(cpuid)
    ref = clip[0]
    rs = np.random.default_rng(32)
    for (w, h) in [(8, 8), (16, 16), (32, 32), (64, 64), (16, 8), (32, 16)]:
        for _ in range(4):
            x = int(rs.integers(0, (W - w) // 8 + 1)) * 8
            y = int(rs.integers(0, (H - h) // 8 + 1)) * 8
            mvx, mvy = int(rs.integers(-40, 41)), int(rs.integers(-40, 41))
            for step in range(4):
                hx, hy = (int(rs.integers(-1, 2)), int(rs.integers(-1, 2))) if step >= 2 else (0, 0)
                want = kvzlibs.aligned((4, 64 * 64), np.uint8)
                got = np.zeros((4, 64 * 64), np.uint8)
                L.ref_shim_fme_filter(ptr(ref), W, H, x, y, mvx, mvy, w, h, step, 4, hx, hy, ptr(want))
                oracle_lib.kvzo_fme_blocks_8(ptr(ref), W, H, x + mvx, y + mvy, w, h, step, hx, hy, ptr(got))
                wv = want.reshape(4, 64, 64)[:, :h, :w]
                gv = got.reshape(4, 64, 64)[:, :h, :w]
                assert np.array_equal(gv, wv), (w, h, x, y, mvx, mvy, step, hx, hy)


def _rand_job(rs, algo, size, fme=4, owf=0, wpp=0, extra=True):
    x = int(rs.integers(0, (W - size) // 8 + 1)) * 8
    y = int(rs.integers(0, (H - size) // 8 + 1)) * 8
    nm = int(rs.integers(0, 4))
    merge = []
    for _ in range(nm):
        d = int(rs.integers(1, 4))
        merge.append((d, int(rs.integers(0, 2)), int(rs.integers(0, 2)),
                      int(rs.integers(-60, 61)), int(rs.integers(-60, 61)),
                      int(rs.integers(-60, 61)), int(rs.integers(-60, 61))))
    return kvzlibs.job_pair(
        width=W, height=H, origin_x=x, origin_y=y, pu_w=size, pu_h=size, tile_off_x=0, tile_off_y=0,
        mv_cand=[[int(rs.integers(-40, 41)), int(rs.integers(-40, 41))],
                 [int(rs.integers(-40, 41)), int(rs.integers(-40, 41))]],
        merge=merge, ref_idx=int(rs.integers(0, 2)),
        extra_mv=(int(rs.integers(-30, 31)), int(rs.integers(-30, 31))) if extra else (0, 0),
        lambda_sqrt=float(rs.uniform(3.0, 40.0)),
        owf=owf, wpp=wpp, sao_type=3, deblock_enable=1, mv_constraint=0,
        max_ref_lcu_right=1, max_ref_lcu_down=1,
        fme_level=fme, me_early_termination=int(rs.integers(0, 3)), me_max_steps=int(rs.integers(-1, 8)),
        ime_algorithm=algo)


def test_mvd_cost_and_within_tile_vs_reference(oracle_lib, ref_lib):
    L = kvzlibs.ref(1)
    rs = np.random.default_rng(33)
    for it in range(60):
        oj, rj = _rand_job(rs, 0, 16, owf=it % 2, wpp=it % 2)
        for _ in range(30):
            mx, my = int(rs.integers(-80, 81)), int(rs.integers(-80, 81))
            for shift in (0, 1, 2):
                bw, bg = C.c_uint32(0), C.c_uint32(0)
                want = L.ref_shim_mvd_cost(C.byref(rj), mx, my, shift, C.byref(bw))
                got = oracle_lib.kvzo_mvd_cost(C.byref(oj), mx, my, shift, C.byref(bg))
                assert (got, bg.value) == (want, bw.value)
            assert oracle_lib.kvzo_fracmv_within_tile(C.byref(oj), mx * 3, my * 3) == \
                L.ref_shim_fracmv_within_tile(C.byref(rj), mx * 3, my * 3)


@pytest.mark.parametrize("algo", [kvzlibs_algo for kvzlibs_algo in (0, 1, 3, 4, 7)],
                         ids=["hexbs", "tz", "full8", "full16", "dia"])
@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_ime_vs_reference(oracle_lib, ref_lib, clip, algo, cpuid):
    L = kvzlibs.ref(cpuid)
    pic, ref = clip[2], clip[0]
    rs = np.random.default_rng(40 + algo)
    n = 6 if algo in (3, 4) else 40
    for it in range(n):
        size = (8, 16, 32, 64)[it % 4]
        oj, rj = _rand_job(rs, algo, size, owf=(it // 4) % 2, wpp=(it // 4) % 2)
        rj.pic_y, rj.ref_y = pic.ctypes.data, ref.ctypes.data
        L.ref_shim_ime(C.byref(rj))
        oracle_lib.kvzo_ime_8(C.byref(oj), ptr(pic), ptr(ref))
        assert (oj.best_mv[0], oj.best_mv[1], oj.best_cost, oj.best_bitcost) == \
               (rj.best_mv[0], rj.best_mv[1], rj.best_cost, rj.best_bitcost), (algo, it, size)


@pytest.mark.parametrize("cpuid", [0, 1], ids=["generic", "avx2"])
def test_fme_and_int_satd_vs_reference(oracle_lib, ref_lib, clip, cpuid):
    L = kvzlibs.ref(cpuid)
    pic, ref = clip[2], clip[0]
    rs = np.random.default_rng(50)
    for it in range(48):
        size = (8, 16, 32, 64)[it % 4]
        fme = (it // 4) % 4 + 1
        oj, rj = _rand_job(rs, 0, size, fme=fme, owf=(it // 16) % 2, wpp=(it // 16) % 2)
        rj.pic_y, rj.ref_y = pic.ctypes.data, ref.ctypes.data
        L.ref_shim_ime(C.byref(rj))
        for j in (oj,):
            j.best_mv[0], j.best_mv[1] = rj.best_mv[0], rj.best_mv[1]
            j.best_cost, j.best_bitcost = rj.best_cost, rj.best_bitcost
        # SATD at the integer vector
        oj2, rj2 = kvzlibs.OracleMeJob.from_buffer_copy(oj), kvzlibs.RefMeJob.from_buffer_copy(rj)
        L.ref_shim_int_satd(C.byref(rj2))
        oracle_lib.kvzo_int_satd_8(C.byref(oj2), ptr(pic), ptr(ref))
        assert oj2.best_cost == rj2.best_cost
        # fractional refinement
        L.ref_shim_fme(C.byref(rj))
        oracle_lib.kvzo_fme_8(C.byref(oj), ptr(pic), ptr(ref))
        assert (oj.best_mv[0], oj.best_mv[1], oj.best_cost, oj.best_bitcost) == \
               (rj.best_mv[0], rj.best_mv[1], rj.best_cost, rj.best_bitcost), (it, size, fme)
