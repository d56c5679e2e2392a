"""Batched block operations of the C ABI (costs, transforms, quantisation, interpolation,
intra prediction) on device-resident arrays.  numpy structured dtypes mirror the job
structures of include/kvz_cuda.h; torch is used for device memory and streams only."""
import ctypes as C

import numpy as np
import torch

from . import capi

COST_REG_SAD, COST_SAD_NXN, COST_SATD_NXN, COST_SATD_ANY, COST_SATD_QUAD, COST_SSD = range(6)
IMAGE_SAD, IMAGE_SATD = 0, 1

BLOCK_JOB_DT = np.dtype([("a_offset", "<u4"), ("b_offset", "<u4"), ("a_stride", "<i4"), ("b_stride", "<i4"),
                         ("w", "<i2"), ("h", "<i2"), ("kind", "<i4")])
IMAGE_JOB_DT = np.dtype([("pic_x", "<i4"), ("pic_y", "<i4"), ("ref_x", "<i4"), ("ref_y", "<i4"),
                         ("w", "<i2"), ("h", "<i2"), ("kind", "<i4")])
assert BLOCK_JOB_DT.itemsize == 24 and IMAGE_JOB_DT.itemsize == 24


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


_RESIDENT = {}


def keep_on_device(a):
    """Upload a numpy job / parameter array once; later calls that are handed the SAME array object reuse the device
    copy instead of uploading it again (the arrays of a frame that is processed repeatedly, the timing tools)."""
    import weakref
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
    _RESIDENT[id(a)] = (weakref.ref(a), t)
    return a


def _dev(a):
    """numpy array -> device uint8 tensor of its bytes"""
    hit = _RESIDENT.get(id(a))
    if hit is not None and hit[0]() is a:
        return hit[1]
    return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()


def _p(t):
    return C.c_void_p(t.data_ptr())


def block_costs(d_a, d_b, jobs, bitdepth=8):
    """d_a, d_b: device tensors holding the blocks; jobs: numpy BLOCK_JOB_DT array.
    Returns a device int32 tensor with one cost per job."""
    capi.init(torch.cuda.current_device())
    d_jobs = _dev(jobs)
    out = torch.zeros(len(jobs), dtype=torch.int32, device="cuda")
    capi.check(capi.lib().kvz_cuda_block_costs(_p(d_a), _p(d_b), _p(d_jobs), len(jobs), bitdepth, _p(out), _stream()),
               "kvz_cuda_block_costs")
    return out


def image_costs(d_pic, d_ref, jobs, bitdepth=8):
    """d_pic, d_ref: (H, W) device planes (unpadded); jobs: numpy IMAGE_JOB_DT array."""
    capi.init(torch.cuda.current_device())
    h, w = d_pic.shape
    d_jobs = _dev(jobs)
    out = torch.zeros(len(jobs), dtype=torch.int32, device="cuda")
    capi.check(capi.lib().kvz_cuda_image_costs(_p(d_pic), d_pic.stride(0), _p(d_ref), d_ref.stride(0), w, h,
                                               _p(d_jobs), len(jobs), bitdepth, _p(out), _stream()),
               "kvz_cuda_image_costs")
    return out


def _planes(tensors):
    arr = (capi.Plane * len(tensors))()
    for a, t in zip(arr, tensors):
        assert t.dim() == 2 and t.stride(1) == 1
        a.d_data, a.height, a.width = t.data_ptr(), t.shape[0], t.shape[1]
        a.pitch = t.stride(0) * t.element_size()
    return arr


def picture_checksum(planes, bitdepth=8):
    """planes: up to 3 (H, W) device tensors (uint8, or int16/uint16 samples at bitdepth > 8; row-strided views
    allowed).  Returns a device int32 tensor: per plane the 32-bit sum of kvz_array_checksum
    (nal-generic.c:58-82) before the reference unpacks it big-endian."""
    capi.init(torch.cuda.current_device())
    out = torch.empty(len(planes), dtype=torch.int32, device="cuda")
    arr = _planes(planes)
    capi.check(capi.lib().kvz_cuda_picture_checksum(C.addressof(arr), len(planes), bitdepth, _p(out), _stream()),
               "kvz_cuda_picture_checksum")
    return out


def picture_sse(src_planes, rec_planes, bitdepth=8):
    """Per plane the sum of squared differences compute_psnr (encmain.c:101-131) accumulates; device int64 tensor."""
    capi.init(torch.cuda.current_device())
    out = torch.empty(len(src_planes), dtype=torch.int64, device="cuda")
    a, b = _planes(src_planes), _planes(rec_planes)
    capi.check(capi.lib().kvz_cuda_picture_sse(C.addressof(a), C.addressof(b), len(src_planes), bitdepth, _p(out), _stream()),
               "kvz_cuda_picture_sse")
    return out


SAO_PARAMS_DT = np.dtype([("type", "i1"), ("eo_class", "i1"), ("band_position", "i1", (2,)), ("offsets", "i1", (10,)),
                          ("reserved", "i1", (2,))])
SAO_STATS_DT = np.dtype([("edge", "<i4", (4, 2, 5)), ("band", "<i4", (2, 32))])
assert SAO_PARAMS_DT.itemsize == 16 and SAO_STATS_DT.itemsize == 416


def sao_ctu_counts(planes):
    """CTUs per plane: 64x64 for plane 0, 32x32 for the chroma planes of a 4:2:0 picture"""
    return [((t.shape[0] + c - 1) // c) * ((t.shape[1] + c - 1) // c) for t, c in zip(planes, (64, 32, 32))]


def sao_ctu_stats(orig_planes, rec_planes, bitdepth=8):
    """kvz_cuda_sao_ctu_stats: device uint8 tensor holding SAO_STATS_DT records, plane-major, CTUs in raster order"""
    capi.init(torch.cuda.current_device())
    n = sum(sao_ctu_counts(orig_planes))
    out = torch.empty(n * SAO_STATS_DT.itemsize, dtype=torch.uint8, device="cuda")
    a, b = _planes(orig_planes), _planes(rec_planes)
    capi.check(capi.lib().kvz_cuda_sao_ctu_stats(C.addressof(a), C.addressof(b), len(orig_planes), bitdepth, _p(out), _stream()),
               "kvz_cuda_sao_ctu_stats")
    return out


def sao_reconstruct(rec_planes, out_planes, luma_params, chroma_params, bitdepth=8):
    """kvz_cuda_sao_reconstruct: *_params are numpy SAO_PARAMS_DT arrays (one record per CTU) or device tensors"""
    capi.init(torch.cuda.current_device())
    d_l = _dev(luma_params) if isinstance(luma_params, np.ndarray) else luma_params
    d_c = (_dev(chroma_params) if isinstance(chroma_params, np.ndarray) else chroma_params) if chroma_params is not None else None
    a, b = _planes(rec_planes), _planes(out_planes)
    capi.check(capi.lib().kvz_cuda_sao_reconstruct(C.addressof(a), C.addressof(b), len(rec_planes), bitdepth, _p(d_l),
                                                   _p(d_c) if d_c is not None else None, _stream()),
               "kvz_cuda_sao_reconstruct")
    return out_planes


SCU_DT = np.dtype([("type", "u1"), ("depth", "u1"), ("tr_depth", "u1"), ("part_size", "u1"), ("cbf_y", "u1"),
                   ("mv_dir", "u1"), ("mv_ref", "u1", (2,)), ("mv", "<i2", (2, 2)), ("qp", "i1"), ("reserved", "u1", (3,))])
assert SCU_DT.itemsize == 20


def deblock_picture(planes, scus, qp, cu_qp=0, beta_offset_div2=0, tc_offset_div2=0, slice_b=0, ref_LX=None, bitdepth=8):
    """kvz_cuda_deblock_picture: planes = [luma] or [luma, u, v] device tensors, filtered in place;
    scus: numpy SCU_DT array (height / 4, width / 4) or a device tensor of its bytes"""
    capi.init(torch.cuda.current_device())
    cfg = capi.DeblockCfg()
    cfg.height, cfg.width = planes[0].shape
    cfg.qp, cfg.cu_qp, cfg.beta_offset_div2, cfg.tc_offset_div2, cfg.slice_b = qp, cu_qp, beta_offset_div2, tc_offset_div2, slice_b
    for l in range(2):
        for i in range(16):
            cfg.ref_LX[l][i] = ref_LX[l][i] if ref_LX is not None else i
    d_scus = _dev(scus) if isinstance(scus, np.ndarray) else scus
    arr = _planes(planes)
    capi.check(capi.lib().kvz_cuda_deblock_picture(C.byref(cfg), _p(d_scus), C.addressof(arr), len(planes), bitdepth, _stream()),
               "kvz_cuda_deblock_picture")
    return planes


RESIDUAL_JOB_DT = np.dtype([("ref_offset", "<u4"), ("pred_offset", "<u4"), ("rec_offset", "<u4"),
                            ("coeff_offset", "<u4"), ("in_stride", "<i4"), ("out_stride", "<i4"),
                            ("width", "u1"), ("color", "u1"), ("scan_idx", "u1"), ("use_trskip", "u1"),
                            ("qp", "i1"), ("cu_is_intra", "u1"), ("reserved", "u1", (2,))])
assert RESIDUAL_JOB_DT.itemsize == 32


def transform_blocks(d_in, n, inverse=False, dst4=False, bitdepth=8):
    """d_in: device int16 tensor of `count` contiguous n x n blocks -> same-shaped output
    (kvz_dct_NxN / kvz_idct_NxN / the 4x4 DST, dct-generic.c:206-617)."""
    capi.init(torch.cuda.current_device())
    out = torch.empty_like(d_in)
    count = d_in.numel() // (n * n)
    capi.check(capi.lib().kvz_cuda_transform_blocks(n, int(dst4), int(inverse), bitdepth, _p(d_in), _p(out), count,
                                                    _stream()), "kvz_cuda_transform_blocks")
    return out


def _qcfg(qp, bitdepth, slice_is_intra, signhide, cu_is_intra):
    return capi.QuantCfg(qp, bitdepth, int(slice_is_intra), int(signhide), int(cu_is_intra))


def quant_blocks(d_coef, n, qp, type_=0, scan_idx=0, slice_is_intra=False, signhide=False, bitdepth=8):
    """kvz_quant (quant-generic.c:37-163) for contiguous n x n int16 blocks."""
    capi.init(torch.cuda.current_device())
    out = torch.empty_like(d_coef)
    cfg = _qcfg(qp, bitdepth, slice_is_intra, signhide, slice_is_intra)
    capi.check(capi.lib().kvz_cuda_quant_blocks(C.byref(cfg), n, type_, scan_idx, _p(d_coef), _p(out),
                                                d_coef.numel() // (n * n), _stream()), "kvz_cuda_quant_blocks")
    return out


def dequant_blocks(d_q, n, qp, type_=0, bitdepth=8):
    """kvz_dequant (quant-generic.c:279-321)."""
    capi.init(torch.cuda.current_device())
    out = torch.empty_like(d_q)
    cfg = _qcfg(qp, bitdepth, False, False, False)
    capi.check(capi.lib().kvz_cuda_dequant_blocks(C.byref(cfg), n, type_, _p(d_q), _p(out), d_q.numel() // (n * n),
                                                  _stream()), "kvz_cuda_dequant_blocks")
    return out


def coeff_abs_sums(d_coeffs, per_block):
    capi.init(torch.cuda.current_device())
    count = d_coeffs.numel() // per_block
    out = torch.zeros(count, dtype=torch.int32, device="cuda")
    capi.check(capi.lib().kvz_cuda_coeff_abs_sums(_p(d_coeffs), per_block, count, _p(out), _stream()),
               "kvz_cuda_coeff_abs_sums")
    return out


def quantize_residuals(d_ref, d_pred, d_rec, d_coeff, jobs, slice_is_intra=False, signhide=False, bitdepth=8):
    """Fused kvz_quantize_residual (quant-generic.c:180-273) for a batch of transform blocks;
    jobs: numpy RESIDUAL_JOB_DT array.  Returns the has_coeffs flags (device uint8)."""
    capi.init(torch.cuda.current_device())
    d_jobs = _dev(jobs)
    has = torch.zeros(len(jobs), dtype=torch.uint8, device="cuda")
    cfg = _qcfg(0, bitdepth, slice_is_intra, signhide, False)
    capi.check(capi.lib().kvz_cuda_quantize_residuals(C.byref(cfg), _p(d_ref), _p(d_pred), _p(d_rec), _p(d_coeff),
                                                      _p(d_jobs), len(jobs), _p(has), _stream()),
               "kvz_cuda_quantize_residuals")
    return has


RDOQ_CTX_DT = np.dtype([("cbf_luma", "u1", (4,)), ("cbf_chroma", "u1", (4,)), ("sig_group", "u1", (4,)),
                        ("sig_luma", "u1", (27,)), ("sig_chroma", "u1", (15,)),
                        ("last_y_luma", "u1", (15,)), ("last_y_chroma", "u1", (15,)),
                        ("last_x_luma", "u1", (15,)), ("last_x_chroma", "u1", (15,)),
                        ("one_luma", "u1", (16,)), ("one_chroma", "u1", (8,)),
                        ("abs_luma", "u1", (4,)), ("abs_chroma", "u1", (2,)),
                        ("root_cbf", "u1"), ("trskip_luma", "u1"), ("trskip_chroma", "u1"), ("reserved", "u1")])
assert RDOQ_CTX_DT.itemsize == 148
RDOQ_JOB_DT = np.dtype([("coef_offset", "<u4"), ("ctx_index", "<u4"), ("lambda", "<f8"), ("width", "u1"),
                        ("type", "u1"), ("scan_idx", "u1"), ("tr_depth", "u1"), ("qp", "i1"),
                        ("cu_is_intra", "u1"), ("reserved", "u1", (2,))])
assert RDOQ_JOB_DT.itemsize == 24


def rdoq_blocks(d_coef, ctxs, jobs, signhide=False, bitdepth=8):
    """kvz_rdoq (rdo.c:548-885) for a batch of transform blocks.  d_coef: device int16 coefficients;
    ctxs: numpy RDOQ_CTX_DT array of CABAC context snapshots; jobs: numpy RDOQ_JOB_DT array.
    Returns the levels (device int16, same layout as d_coef)."""
    capi.init(torch.cuda.current_device())
    d_jobs, d_ctx = _dev(jobs), _dev(ctxs)
    out = torch.zeros_like(d_coef)
    cfg = _qcfg(0, bitdepth, False, signhide, False)
    max_width = int(jobs["width"].max()) if len(jobs) else 4
    capi.check(capi.lib().kvz_cuda_rdoq_blocks(C.byref(cfg), _p(d_ctx), _p(d_coef), _p(out), _p(d_jobs), len(jobs),
                                               max_width, _stream()), "kvz_cuda_rdoq_blocks")
    return out


COEFF_JOB_DT = np.dtype([("coef_offset", "<u4"), ("ctx_index", "<u4"), ("range", "<u2"), ("width", "u1"), ("type", "u1"),
                         ("scan_idx", "u1"), ("reserved", "u1", (3,))])
assert COEFF_JOB_DT.itemsize == 16


def coeff_bits(d_coeffs, ctxs, jobs, signhide=False, trskip_enable=False, d_jobs=None):
    """kvz_get_coeff_cost's CABAC path (rdo.c:158-197) for a batch of coefficient blocks.  d_coeffs: device int16;
    ctxs: numpy RDOQ_CTX_DT snapshots (or a device tensor of them); jobs: numpy COEFF_JOB_DT (d_jobs: the same array
    already on the device).  Returns device int32 bits."""
    capi.init(torch.cuda.current_device())
    d_jobs = d_jobs if d_jobs is not None else _dev(jobs)
    d_ctx = _dev(ctxs) if isinstance(ctxs, np.ndarray) else ctxs
    out = torch.zeros(len(jobs), dtype=torch.int32, device="cuda")
    max_width = int(jobs["width"].max()) if len(jobs) else 4
    capi.check(capi.lib().kvz_cuda_coeff_bits(_p(d_ctx), _p(d_coeffs), _p(d_jobs), len(jobs), max_width, int(signhide),
                                              int(trskip_enable), _p(out), _stream()), "kvz_cuda_coeff_bits")
    return out


def quantize_residuals_rdoq(d_ref, d_pred, d_rec, d_coeff, jobs, ctxs, rdoq_jobs, slice_is_intra=False, signhide=False,
                            rdoq_skip=False, bitdepth=8):
    """kvz_quantize_residual with cfg.rdoq_enable (quant-generic.c:180-273 + rdo.c:548-885) for a batch of
    transform blocks.  jobs: RESIDUAL_JOB_DT, rdoq_jobs: RDOQ_JOB_DT (paired), ctxs: RDOQ_CTX_DT."""
    capi.init(torch.cuda.current_device())
    d_jobs, d_rj, d_ctx = _dev(jobs), _dev(rdoq_jobs), _dev(ctxs)
    has = torch.zeros(len(jobs), dtype=torch.uint8, device="cuda")
    work = torch.empty_like(d_coeff)
    cfg = _qcfg(0, bitdepth, slice_is_intra, signhide, False)
    max_width = int(jobs["width"].max()) if len(jobs) else 4
    capi.check(capi.lib().kvz_cuda_quantize_residuals_rdoq(
        C.byref(cfg), _p(d_ref), _p(d_pred), _p(d_rec), _p(d_coeff), _p(d_jobs), len(jobs), _p(has), _p(d_ctx), _p(d_rj),
        max_width, _p(work), int(rdoq_skip), _stream()), "kvz_cuda_quantize_residuals_rdoq")
    return has


NOALLOC_MAX_COEFFS = 1 << 20      # KVZ_CUDA_NOALLOC_MAX_COEFFS (include/kvz_cuda.h)


def quantize_residuals_rdoq_noalloc(d_ref, d_pred, d_rec, d_coeff, d_jobs, n_jobs, d_has, d_ctx, d_rdoq_jobs, max_width,
                                    slice_is_intra=False, signhide=False, rdoq_skip=False, bitdepth=8):
    """kvz_cuda_quantize_residuals_rdoq_noalloc: the call above for callers that cannot allocate.  Every argument is a
    tensor already on the device (this wrapper allocates nothing either); the transform-coefficient scratch is the
    library's built-in device array of NOALLOC_MAX_COEFFS int16.  Calls that overlap on the GPU race on it."""
    cfg = _qcfg(0, bitdepth, slice_is_intra, signhide, False)
    capi.check(capi.lib().kvz_cuda_quantize_residuals_rdoq_noalloc(
        C.byref(cfg), _p(d_ref), _p(d_pred), _p(d_rec), _p(d_coeff), d_coeff.numel(), _p(d_jobs), int(n_jobs), _p(d_has),
        _p(d_ctx), _p(d_rdoq_jobs), int(max_width), int(rdoq_skip), _stream()), "kvz_cuda_quantize_residuals_rdoq_noalloc")
    return d_has


INTERP_PIXEL, INTERP_14BIT, INTERP_FME = 0, 1, 2
INTERP_JOB_DT = np.dtype([("x", "<i4"), ("y", "<i4"), ("mv_x", "<i4"), ("mv_y", "<i4"), ("w", "<i2"), ("h", "<i2"),
                          ("chroma", "u1"), ("mode", "u1"), ("reserved", "u1", (2,)), ("dst_offset", "<u4"),
                          ("dst_stride", "<i4")])
assert INTERP_JOB_DT.itemsize == 32


def interp_blocks(d_plane, jobs, d_dst8=None, d_dst16=None, bitdepth=8):
    """Luma (8-tap, 1/4-pel) / chroma (4-tap, 1/8-pel) interpolation of a batch of blocks of one
    unpadded device plane; jobs: numpy INTERP_JOB_DT array."""
    capi.init(torch.cuda.current_device())
    h, w = d_plane.shape
    d_jobs = _dev(jobs)
    capi.check(capi.lib().kvz_cuda_interp_blocks(
        _p(d_plane), d_plane.stride(0) * d_plane.element_size(), w, h, bitdepth, _p(d_jobs), len(jobs),
        _p(d_dst8) if d_dst8 is not None else None, _p(d_dst16) if d_dst16 is not None else None, _stream()),
        "kvz_cuda_interp_blocks")


def bipred_blend(d_s0, d_s1, bitdepth=8):
    capi.init(torch.cuda.current_device())
    out = torch.empty(d_s0.shape, dtype=torch.uint8 if bitdepth == 8 else torch.int16, device="cuda")
    capi.check(capi.lib().kvz_cuda_bipred_blend(_p(d_s0), _p(d_s1), d_s0.numel(), bitdepth, _p(out), _stream()),
               "kvz_cuda_bipred_blend")
    return out


INTRA_JOB_DT = np.dtype([("ref_offset", "<u4"), ("orig_offset", "<u4"), ("orig_stride", "<i4"), ("pred_offset", "<u4"),
                         ("log2_width", "u1"), ("is_chroma", "u1"), ("filter_boundary", "u1"), ("has_orig", "u1")])
assert INTRA_JOB_DT.itemsize == 20


def intra_mode_costs(d_refs, d_orig, jobs, want_preds=False, bitdepth=8):
    """All 35 intra predictions of each PU and their {SATD, SAD} against the source block.
    Returns (costs int32 [n, 35, 2] device tensor, preds uint8 device tensor or None)."""
    capi.init(torch.cuda.current_device())
    d_jobs = _dev(jobs)
    costs = torch.zeros((len(jobs), 35, 2), dtype=torch.int32, device="cuda")
    preds = None
    if want_preds:
        total = int(jobs["pred_offset"].max()) + 35 * (1 << (2 * int(jobs["log2_width"][np.argmax(jobs["pred_offset"])])))
        preds = torch.zeros(total, dtype=torch.uint8 if bitdepth == 8 else torch.int16, device="cuda")
    capi.check(capi.lib().kvz_cuda_intra_mode_costs(
        _p(d_refs), _p(d_orig) if d_orig is not None else None, _p(d_jobs), len(jobs), bitdepth, _p(costs),
        _p(preds) if preds is not None else None, _stream()), "kvz_cuda_intra_mode_costs")
    return costs, preds
