"""GPU parity of the PER-CALL face: the functions with the reference's strategy signatures
(include/kvz_cuda_strategies.h), called with host pointers like the encoder would, against the
reference's known-answer vectors and the oracle; and their registration into the reference's
own strategy list (kvz_strategyselector_register)."""
import ctypes as C

import numpy as np
import pytest

import kat
import kvzlibs
from kvzlibs import ptr

pytestmark = pytest.mark.gpu
PX = C.POINTER(C.c_uint8)


def _lib():
    import kvazaar_b200 as kb
    kb.init(0)
    return kb.lib()


def _p8(a):
    return a.ctypes.data_as(PX)


def test_picture_strategies():
    L = _lib()
    O = kvzlibs.oracle()
    L.kvz_cuda_reg_sad.restype = C.c_uint
    pic, ref = kat.reg_sad_pictures()
    for (w, h) in kat.REG_SAD_SHAPES:
        want = int(np.abs(pic[:h, :w].astype(np.int64) - ref[:h, :w].astype(np.int64)).sum())
        assert L.kvz_cuda_reg_sad(_p8(pic), _p8(ref), w, h, 64, 64) == want
    for test in (0, 1, 2):
        for n in (4, 8, 16, 32, 64):
            a, b = kat.satd_buffers(test, n)
            f = getattr(L, f"kvz_cuda_satd_{n}x{n}"); f.restype = C.c_uint
            assert f(_p8(a), _p8(b)) == f(_p8(b), _p8(a)) == kat.SATD_GOLDEN[test][n]
    rs = np.random.default_rng(1)
    for n in (4, 8, 16, 32):
        preds = rs.integers(0, 256, (2, 32 * 32), dtype=np.uint8)
        orig = rs.integers(0, 256, n * n, dtype=np.uint8)
        for kind, ofn in (("sad", O.kvzo_sad_nxn_8), ("satd", O.kvzo_satd_nxn_8)):
            costs = (C.c_uint * 2)()
            getattr(L, f"kvz_cuda_{kind}_{n}x{n}_dual")(ptr(preds), _p8(orig), 2, costs)
            assert list(costs) == [ofn(n, ptr(preds[i]), ptr(orig)) for i in range(2)]
    a = rs.integers(0, 256, (64, 80), dtype=np.uint8)
    ps = [rs.integers(0, 256, (64, 64), dtype=np.uint8) for _ in range(4)]
    L.kvz_cuda_satd_any_size.restype = C.c_uint
    L.kvz_cuda_pixels_calc_ssd.restype = C.c_uint
    for (w, h) in [(8, 8), (16, 32), (64, 64), (12, 16), (24, 8), (4, 4)]:
        assert L.kvz_cuda_satd_any_size(w, h, _p8(a), 80, _p8(ps[0]), 64) == O.kvzo_satd_any_size_8(w, h, ptr(a), 80, ptr(ps[0]), 64)
        parr = (PX * 4)(*[_p8(p) for p in ps])
        got = (C.c_uint * 4)(); want = (C.c_uint * 4)()
        valid = (C.c_int8 * 4)(1, 1, 1, 1)
        L.kvz_cuda_satd_any_size_quad(w, h, parr, 64, _p8(a), 80, 4, got, valid)
        O.kvzo_satd_any_size_quad_8(w, h, (C.c_void_p * 4)(*[p.ctypes.data for p in ps]), 64, ptr(a), 80, want)
        assert list(got) == list(want)
    for n in (4, 8, 16, 32, 64):
        assert L.kvz_cuda_pixels_calc_ssd(_p8(a), _p8(ps[1]), 80, 64, n) == O.kvzo_ssd_8(ptr(a), ptr(ps[1]), 80, 64, n)


def test_dct_quant_strategies():
    L = _lib()
    O = kvzlibs.oracle()
    rs = np.random.default_rng(2)
    names = {(4, 1): "fast_forward_dst_4x4", (4, 0): "dct_4x4", (8, 0): "dct_8x8", (16, 0): "dct_16x16", (32, 0): "dct_32x32"}
    inames = {(4, 1): "fast_inverse_dst_4x4", (4, 0): "idct_4x4", (8, 0): "idct_8x8", (16, 0): "idct_16x16", (32, 0): "idct_32x32"}
    for (n, dst), nm in names.items():
        for blk in (kat.dct_gradient_block(n, 0), kat.dct_gradient_block(n, 1), rs.integers(-255, 256, (n, n)).astype(np.int16)):
            got = np.zeros((n, n), np.int16); want = np.zeros((n, n), np.int16)
            getattr(L, "kvz_cuda_" + nm)(8, ptr(blk), ptr(got))
            O.kvzo_fwd_transform(n, dst, 8, ptr(blk), ptr(want))
            assert np.array_equal(got, want), nm
            back = np.zeros((n, n), np.int16); bw = np.zeros((n, n), np.int16)
            getattr(L, "kvz_cuda_" + inames[(n, dst)])(8, ptr(want), ptr(back))
            O.kvzo_inv_transform(n, dst, 8, ptr(want), ptr(bw))
            assert np.array_equal(back, bw), inames[(n, dst)]
    c = kat.coeff_series()
    L.kvz_cuda_coeff_abs_sum.restype = C.c_uint32
    L.kvz_cuda_coeff_abs_sum.argtypes = [C.c_void_p, C.c_size_t]
    assert L.kvz_cuda_coeff_abs_sum(ptr(c), c.size) == int(np.abs(c.astype(np.int64)).sum())
    from kvazaar_b200.capi import QuantCfg
    for n in (4, 8, 16, 32):
        coef = (rs.standard_normal((n, n)) * 300).astype(np.int16)
        cfg = QuantCfg(27, 8, 0, 1, 0); oc = kvzlibs.OracleQCfg(27, 8, 0, 1, 0)
        q = np.zeros((n, n), np.int16); qw = np.zeros((n, n), np.int16)
        assert L.kvz_cuda_quant_host(C.byref(cfg), ptr(coef), ptr(q), n, n, 0, 0) == 1
        O.kvzo_quant(C.byref(oc), ptr(coef), ptr(qw), n, 0, 0)
        assert np.array_equal(q, qw)
        d = np.zeros((n, n), np.int16); dw = np.zeros((n, n), np.int16)
        assert L.kvz_cuda_dequant_host(C.byref(cfg), ptr(q), ptr(d), n, n, 0) == 1
        O.kvzo_dequant(C.byref(oc), ptr(qw), ptr(dw), n, 0)
        assert np.array_equal(d, dw)
        ref_in = rs.integers(0, 256, (n, n + 8), dtype=np.uint8)
        pred = np.clip(ref_in.astype(np.int64) + rs.integers(-30, 31, ref_in.shape), 0, 255).astype(np.uint8)
        rec = np.zeros((n, n + 8), np.uint8); recw = np.zeros((n, n + 8), np.uint8)
        co = np.zeros(n * n, np.int16); cow = np.zeros(n * n, np.int16)
        h = L.kvz_cuda_quantize_residual_host(C.byref(cfg), n, 0, 0, 0, n + 8, n + 8, _p8(ref_in), _p8(pred), _p8(rec), ptr(co))
        hw = O.kvzo_quantize_residual_8(C.byref(oc), n, 0, 0, 0, n + 8, n + 8, ptr(ref_in), ptr(pred), ptr(recw), ptr(cow))
        assert h == hw and np.array_equal(co, cow) and np.array_equal(rec[:, :n], recw[:, :n])


def test_intra_and_ipol_strategies():
    L = _lib()
    O = kvzlibs.oracle()
    rs = np.random.default_rng(3)
    for lg in (2, 3, 4, 5):
        w = 1 << lg
        top = rs.integers(0, 256, 2 * w + 1, dtype=np.uint8); left = rs.integers(0, 256, 2 * w + 1, dtype=np.uint8)
        left[0] = top[0]
        for mode in (0, 2, 9, 10, 11, 18, 25, 26, 27, 34):
            got = np.zeros(w * w, np.uint8); want = np.zeros(w * w, np.uint8)
            if mode == 0:
                L.kvz_cuda_intra_pred_planar(lg, _p8(top), _p8(left), _p8(got))
            else:
                L.kvz_cuda_angular_pred(lg, mode, _p8(top), _p8(left), _p8(got))
            O.kvzo_intra_predict_8(ptr(top), ptr(left), lg, mode, 1, 0, ptr(want))   # chroma path == raw strategy output
            assert np.array_equal(got, want), (lg, mode)
    # ipol: an extended block with margins, like kvz_get_extended_block returns
    W, H = 96, 96
    plane = kvzlibs.synth_luma(4, W, H, 1)[0]
    for (w, h) in [(8, 8), (16, 16), (32, 32), (64, 64)]:
        x0, y0 = 12, 10
        src = plane[y0:, x0:]
        mv = (C.c_int16 * 2)(int(rs.integers(-3, 4)), int(rs.integers(-3, 4)))
        got = np.zeros((h, w), np.uint8); want = np.zeros((h, w), np.uint8)
        L.kvz_cuda_sample_quarterpel_luma(None, _p8(src), W, w, h, _p8(got), w, 0, 0, mv)
        O.kvzo_interp_luma_8(ptr(plane), W, H, x0, y0, mv[0] & 3, mv[1] & 3, w, h, ptr(want), w)
        assert np.array_equal(got, want), ("luma", w)
        g14 = np.zeros((h, w), np.int16); w14 = np.zeros((h, w), np.int16)
        L.kvz_cuda_sample_14bit_quarterpel_luma(None, _p8(src), W, w, h, ptr(g14), w, 0, 0, mv)
        O.kvzo_interp_luma14_8(ptr(plane), W, H, x0, y0, mv[0] & 3, mv[1] & 3, w, h, ptr(w14), w)
        assert np.array_equal(g14, w14)
        if w <= 32:
            mvc = (C.c_int16 * 2)(int(rs.integers(-7, 8)), int(rs.integers(-7, 8)))
            L.kvz_cuda_sample_octpel_chroma(None, _p8(src), W, w, h, _p8(got), w, 0, 0, mvc)
            O.kvzo_interp_chroma_8(ptr(plane), W, H, x0, y0, mvc[0] & 7, mvc[1] & 7, w, h, ptr(want), w)
            assert np.array_equal(got, want), ("chroma", w)
            L.kvz_cuda_sample_14bit_octpel_chroma(None, _p8(src), W, w, h, ptr(g14), w, 0, 0, mvc)
            O.kvzo_interp_chroma14_8(ptr(plane), W, H, x0, y0, mvc[0] & 7, mvc[1] & 7, w, h, ptr(w14), w)
            assert np.array_equal(g14, w14)
        if w <= 64 and y0 + h + 6 < H and x0 + w + 6 < W:
            # FME filters: src = orig_topleft of the block fetched at (integer vector - 1)
            for step, nm in enumerate(["hpel_blocks_hor_ver", "hpel_blocks_diag", "qpel_blocks_hor_ver", "qpel_blocks_diag"]):
                hx, hy = (int(rs.integers(-1, 2)), int(rs.integers(-1, 2))) if step >= 2 else (0, 0)
                filt = np.zeros((4, 64 * 64), np.uint8); wantb = np.zeros((4, 64, 64), np.uint8)
                srcf = plane[y0 - 1:, x0 - 1:]
                getattr(L, f"kvz_cuda_filter_{nm}_luma")(None, _p8(srcf), W, w, h, ptr(filt), None, 4, None, hx, hy)
                O.kvzo_fme_blocks_8(ptr(plane), W, H, x0, y0, w, h, step, hx, hy, ptr(wantb))
                assert np.array_equal(filt.reshape(4, 64, 64)[:, :h, :w], wantb[:, :h, :w]), (nm, w)


def test_nal_checksum_strategy():
    # array_checksum_func (strategies-nal.h:42-47) on host planes, larger than one staging strip
    from test_oracle_picture_stats import oracle_checksum
    L = _lib()
    L.kvz_cuda_array_checksum.restype = None
    L.kvz_cuda_array_checksum.argtypes = [PX, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_ubyte), C.c_uint8]
    rs = np.random.default_rng(3)
    for h, w, pad in ((1, 1, 0), (68, 100, 12), (1080, 1920, 0), (540, 960, 64), (2160, 3840, 0)):
        big = rs.integers(0, 256, (h, w + pad), dtype=np.uint8)
        out = (C.c_ubyte * 16)()
        L.kvz_cuda_array_checksum(_p8(big), h, w, w + pad, out, 8)
        got = (out[0] << 24) | (out[1] << 16) | (out[2] << 8) | out[3]
        assert got == oracle_checksum(big[:, :w]), (h, w)


def test_sao_strategies():
    # strategies-sao.h:37-60 on host blocks, against the oracle (which is pinned on the reference's strategies)
    import sao_oracle as so
    from sao_oracle import _pp, ints
    L = _lib()
    O = kvzlibs.oracle()
    IP = C.POINTER(C.c_int)
    L.kvz_cuda_sao_edge_ddistortion.restype = C.c_int
    L.kvz_cuda_sao_edge_ddistortion.argtypes = [PX, PX, C.c_int, C.c_int, C.c_int, IP]
    L.kvz_cuda_calc_sao_edge_dir.restype = None
    L.kvz_cuda_calc_sao_edge_dir.argtypes = [PX, PX, C.c_int, C.c_int, C.c_int, IP]
    L.kvz_cuda_sao_band_ddistortion_host.restype = C.c_int
    L.kvz_cuda_sao_band_ddistortion_host.argtypes = [C.c_int, PX, PX, C.c_int, C.c_int, C.c_int, IP]
    L.kvz_cuda_sao_reconstruct_color_host.restype = None
    L.kvz_cuda_sao_reconstruct_color_host.argtypes = [C.c_int, PX, PX, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    rs = np.random.default_rng(8)
    for bw, bh in ((64, 64), (32, 32), (40, 24), (8, 8)):
        base = rs.integers(0, 256, (bh, bw)).astype(np.int32)
        base = (base + np.roll(base, 1, 0) + np.roll(base, 1, 1)) // 3
        orig = base.astype(np.uint8)
        rec = np.clip(base + rs.integers(-8, 9, base.shape), 0, 255).astype(np.uint8)
        for e in range(4):
            offs = [0] + [int(v) for v in rs.integers(-7, 8, 4)]
            assert L.kvz_cuda_sao_edge_ddistortion(_p8(orig), _p8(rec), bw, bh, e, ints(offs)) == \
                O.kvzo_sao_edge_ddistortion_8(_pp(orig), _pp(rec), bw, bh, e, ints(offs))
            a, b = (C.c_int * 10)(*range(10)), (C.c_int * 10)(*range(10))      # accumulated into
            L.kvz_cuda_calc_sao_edge_dir(_p8(orig), _p8(rec), e, bw, bh, a)
            O.kvzo_calc_sao_edge_dir_8(_pp(orig), _pp(rec), e, bw, bh, b)
            assert a[:] == b[:]
        for pos in (0, 11, 28):
            bands = [int(v) for v in rs.integers(-7, 8, 4)]
            assert L.kvz_cuda_sao_band_ddistortion_host(8, _p8(orig), _p8(rec), bw, bh, pos, ints(bands)) == \
                O.kvzo_sao_band_ddistortion_8(_pp(orig), _pp(rec), bw, bh, pos, ints(bands))
        # reconstruct_color on a block inside a larger buffer (the caller owns a one-sample border)
        plane = np.clip(np.pad(base, 1, mode="edge") + rs.integers(-20, 21, (bh + 2, bw + 2)), 0, 255).astype(np.uint8)
        for prm in so.random_params(rs, 6, p_none=0.0):
            for color in (0, 1, 2):
                got, want = np.zeros((bh, bw), np.uint8), np.zeros((bh, bw), np.uint8)
                src = plane[1:, 1:]
                p16 = np.zeros(1, so.SAO_PARAMS_DT); p16[0] = prm
                L.kvz_cuda_sao_reconstruct_color_host(8, _p8(src), _p8(got), p16.ctypes.data_as(C.c_void_p), bw + 2, bw, bw, bh, color)
                s_ = so.info_of(prm)
                O.kvzo_sao_reconstruct_color_8(_pp(src), _pp(want), C.byref(s_), bw + 2, bw, bw, bh, color)
                assert np.array_equal(got, want), (bw, bh, prm, color)


def test_registers_into_the_reference_strategy_list(ref_lib):
    L = _lib()

    class StrategyList(C.Structure):
        _fields_ = [("count", C.c_uint), ("allocated", C.c_uint), ("strategies", C.c_void_p)]

    reg = C.cast(ref_lib.kvz_strategyselector_register, C.c_void_p)
    total = 0
    for fam, n in (("picture", 24), ("dct", 10), ("quant", 1), ("intra", 2), ("ipol", 8), ("nal", 1), ("sao", 2)):
        fn = getattr(L, f"kvz_strategy_register_{fam}_cuda_with")
        fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_uint8, C.c_void_p]
        lst = StrategyList()
        assert fn(C.byref(lst), 8, reg) == 1
        assert lst.count == n, (fam, lst.count)
        total += lst.count
    assert total == 48
