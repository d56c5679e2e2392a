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


def _dev(a):
    """numpy array -> device uint8 tensor of its bytes"""
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
