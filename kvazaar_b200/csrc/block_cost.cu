// kvazaar_b200/csrc/block_cost.cu
//
// Batched block costs: one warp evaluates one job.  Covers, for arrays of blocks in
// device memory, the cost strategies of strategies-picture.h:102-136:
//   reg_sad            picture-generic.c:86-99          (any w x h, two strides)
//   sad_NxN / satd_NxN picture-generic.c:189-196,343-346,460-486 (as w = h = stride = N jobs)
//   satd_any_size      strategies-picture.h:62-100      (4-wide column, 4-high row, 8x8 tiles)
//   satd_any_size_quad picture-generic.c:392-456        (as 4 jobs; only the (w&~7)x(h&~7) area)
//   pixels_calc_ssd    picture-generic.c:521-536
// and the picture-level costs with the reference block clamped into the frame:
//   kvz_image_calc_sad  image.c:455-479 (clamp image.c:320-452)
//   kvz_image_calc_satd image.c:488-545
#include "common.cuh"
#include "satd.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kWarpsPerCta = 4;

// kvz_satd_any_size (strategies-picture.h:62-100): the result is NOT shifted here
template <typename FA, typename FB>
__device__ uint32_t warp_satd_any(const FA &a, const FB &b, int w, int h, int lane)
{
  uint32_t s = 0;
  int x0 = 0, y0 = 0;
  if (w & 7) {                                   // first 4 columns as 4x4 blocks
    s += warp_satd4_strip(a, b, 0, 0, 0, 4, h >> 2, lane);
    x0 = 4; w -= 4;
  }
  if (h & 7) {                                   // first 4 rows of what is left
    s += warp_satd4_strip(a, b, x0, 0, 4, 0, w >> 2, lane);
    y0 = 4; h -= 4;
  }
  s += warp_satd8_area(a, b, x0, y0, w, h, lane);
  return s;
}

template <typename PIX>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
block_cost_kernel(const PIX *__restrict__ base_a, const PIX *__restrict__ base_b,
                  const kvz_cuda_block_job *__restrict__ jobs, int n_jobs, int bitdepth, uint32_t *__restrict__ out)
{
  const int ji = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (ji >= n_jobs) return;
  const kvz_cuda_block_job jb = jobs[ji];
  const StridedBlockT<PIX> a{ base_a + jb.a_offset, jb.a_stride }, b{ base_b + jb.b_offset, jb.b_stride };
  const int sh = bitdepth - 8;
  uint32_t r = 0;
  switch (jb.kind) {
    case KVZ_CUDA_COST_REG_SAD: r = warp_sad(a, b, jb.w, jb.h, lane); break;              // no shift (picture-generic.c:86-99)
    case KVZ_CUDA_COST_SAD_NXN: r = warp_sad(a, b, jb.w, jb.h, lane) >> sh; break;
    case KVZ_CUDA_COST_SATD_NXN:                                                            // 4x4: no shift (picture-generic.c:189-196)
      r = jb.w == 4 ? warp_satd4_strip(a, b, 0, 0, 0, 0, 1, lane) : warp_satd8_area(a, b, 0, 0, jb.w, jb.h, lane) >> sh;
      break;
    case KVZ_CUDA_COST_SATD_ANY: r = warp_satd_any(a, b, jb.w, jb.h, lane) >> sh; break;
    case KVZ_CUDA_COST_SATD_QUAD: r = warp_satd8_area(a, b, 0, 0, jb.w & ~7, jb.h & ~7, lane) >> sh; break;
    case KVZ_CUDA_COST_SSD: r = warp_ssd(a, b, jb.w, jb.w, lane) >> (2 * sh); break;
    default: break;
  }
  if (lane == 0) out[ji] = r;
}

template <typename PIX>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
image_cost_kernel(const PIX *__restrict__ pic, int pic_pitch, const PIX *__restrict__ ref, int ref_pitch,
                  int width, int height, const kvz_cuda_image_cost_job *__restrict__ jobs, int n_jobs,
                  int bitdepth, uint32_t *__restrict__ out)
{
  const int ji = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (ji >= n_jobs) return;
  const kvz_cuda_image_cost_job jb = jobs[ji];
  const StridedBlockT<PIX> a{ pic + (size_t)jb.pic_y * pic_pitch + jb.pic_x, pic_pitch };
  const ClampedPlaneT<PIX> b{ ref, ref_pitch, width, height, jb.ref_x, jb.ref_y };
  const int sh = bitdepth - 8;
  uint32_t r;
  if (jb.kind == KVZ_CUDA_IMAGE_SAD) r = warp_sad(a, b, jb.w, jb.h, lane) >> sh;                 // image.c:474-477
  else r = (warp_satd_any(a, b, jb.w, jb.h, lane) >> sh) >> sh;                                // image.c:507-512: shifted twice
  if (lane == 0) out[ji] = r;
}

}  // namespace

int launch_block_cost(const void *base_a, const void *base_b, const kvz_cuda_block_job *jobs, int n_jobs,
                      int bitdepth, uint32_t *out, cudaStream_t stream)
{
  if (n_jobs <= 0) return 1;
  const int grid = (n_jobs + kWarpsPerCta - 1) / kWarpsPerCta;
  if (bitdepth == 8)
    block_cost_kernel<uint8_t><<<grid, kWarpsPerCta * 32, 0, stream>>>((const uint8_t *)base_a, (const uint8_t *)base_b, jobs,
                                                                       n_jobs, bitdepth, out);
  else   // 10-bit build: uint16_t samples, offsets and strides in samples
    block_cost_kernel<uint16_t><<<grid, kWarpsPerCta * 32, 0, stream>>>((const uint16_t *)base_a, (const uint16_t *)base_b, jobs,
                                                                        n_jobs, bitdepth, out);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_image_cost(const void *pic, int pic_pitch, const void *ref, int ref_pitch, int width, int height,
                      const kvz_cuda_image_cost_job *jobs, int n_jobs, int bitdepth, uint32_t *out, cudaStream_t stream)
{
  if (n_jobs <= 0) return 1;
  const int grid = (n_jobs + kWarpsPerCta - 1) / kWarpsPerCta;
  if (bitdepth == 8)
    image_cost_kernel<uint8_t><<<grid, kWarpsPerCta * 32, 0, stream>>>((const uint8_t *)pic, pic_pitch, (const uint8_t *)ref,
                                                                       ref_pitch, width, height, jobs, n_jobs, bitdepth, out);
  else   // pitches in samples
    image_cost_kernel<uint16_t><<<grid, kWarpsPerCta * 32, 0, stream>>>((const uint16_t *)pic, pic_pitch, (const uint16_t *)ref,
                                                                        ref_pitch, width, height, jobs, n_jobs, bitdepth, out);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
