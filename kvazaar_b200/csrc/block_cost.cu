// kvazaar_b200/csrc/block_cost.cu
//
// Batched block costs: a warp per job, or -- batches of 4x4 / 8x8 blocks -- a thread / 8 lanes per job.  Covers, for arrays of blocks in
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

// N consecutive samples of a row: as 32-bit words when the address allows, else one by one
template <int N, typename PIX>
__device__ __forceinline__ void load_row(const PIX *p, int (&v)[N])
{
  constexpr int kPerWord = 4 / (int)sizeof(PIX);
  if ((reinterpret_cast<uintptr_t>(p) & 3) == 0) {
#pragma unroll
    for (int i = 0; i < N / kPerWord; i++) {
      const uint32_t w = __ldg(reinterpret_cast<const uint32_t *>(p) + i);
#pragma unroll
      for (int k = 0; k < kPerWord; k++) v[i * kPerWord + k] = (int)((w >> (k * 8 * (int)sizeof(PIX))) & ((1u << (8 * (int)sizeof(PIX))) - 1u));
    }
  } else {
#pragma unroll
    for (int i = 0; i < N; i++) v[i] = (int)p[i];
  }
}

// one job by a whole warp
template <typename PIX>
__device__ __forceinline__ uint32_t warp_job_cost(const PIX *base_a, const PIX *base_b, const kvz_cuda_block_job &jb, int bitdepth, int lane)
{
  const StridedBlockT<PIX> a{ base_a + jb.a_offset, jb.a_stride }, b{ base_b + jb.b_offset, jb.b_stride };
  const int sh = bitdepth - 8;
  switch (jb.kind) {
    case KVZ_CUDA_COST_REG_SAD: return warp_sad(a, b, jb.w, jb.h, lane);                     // no shift (picture-generic.c:86-99)
    case KVZ_CUDA_COST_SAD_NXN: return warp_sad(a, b, jb.w, jb.h, lane) >> sh;
    case KVZ_CUDA_COST_SATD_NXN:                                                            // 4x4: no shift (picture-generic.c:189-196)
      return jb.w == 4 ? warp_satd4_strip(a, b, 0, 0, 0, 0, 1, lane) : warp_satd8_area(a, b, 0, 0, jb.w, jb.h, lane) >> sh;
    case KVZ_CUDA_COST_SATD_ANY: return warp_satd_any(a, b, jb.w, jb.h, lane) >> sh;
    case KVZ_CUDA_COST_SATD_QUAD: return warp_satd8_area(a, b, 0, 0, jb.w & ~7, jb.h & ~7, lane) >> sh;
    case KVZ_CUDA_COST_SSD: return warp_ssd(a, b, jb.w, jb.w, lane) >> (2 * sh);
    default: return 0;
  }
}

// 32 consecutive jobs by one warp.  The bulk of a frame's cost calls are 4x4 and 8x8 blocks, where a warp per job
// leaves 16 or more lanes without a sample: 32 jobs that are all 4x4 SAD / SATD are computed one per THREAD
// (satd4_thread), 32 jobs that are all 8x8 by 8 lanes each, a row per lane (satd8_rows), four jobs per pass.
// Anything else: the 32 jobs in turn, a warp each.
__device__ __forceinline__ bool small_job(const kvz_cuda_block_job &jb, int n)
{
  return (jb.kind == KVZ_CUDA_COST_REG_SAD || jb.kind == KVZ_CUDA_COST_SAD_NXN || jb.kind == KVZ_CUDA_COST_SATD_NXN) &&
         jb.w == n && jb.h == n;
}

template <typename PIX>
__device__ __forceinline__ void group_cost(const PIX *base_a, const PIX *base_b, const kvz_cuda_block_job *jobs, int n_jobs,
                                           int first, int bitdepth, int lane, uint32_t *out)
{
  const int mine = first + lane;
  kvz_cuda_block_job jb;
  jb.a_offset = jb.b_offset = 0; jb.a_stride = jb.b_stride = 0; jb.w = jb.h = 0; jb.kind = -1;
  if (mine < n_jobs) jb = jobs[mine];
  const bool all4 = __all_sync(0xffffffffu, small_job(jb, 4));
  const bool all8 = __all_sync(0xffffffffu, small_job(jb, 8));
  const int sh = bitdepth - 8;
  if (all4) {
    const PIX *a = base_a + jb.a_offset, *b = base_b + jb.b_offset;
    int d[16];
#pragma unroll
    for (int y = 0; y < 4; y++) {
      int ra[4], rb[4];
      load_row<4>(a + (size_t)y * jb.a_stride, ra);
      load_row<4>(b + (size_t)y * jb.b_stride, rb);
#pragma unroll
      for (int x = 0; x < 4; x++) d[4 * y + x] = ra[x] - rb[x];
    }
    uint32_t r = 0;
    if (jb.kind == KVZ_CUDA_COST_SATD_NXN) {
      r = satd4_thread(d);                                      // 4x4: no shift (picture-generic.c:189-196)
    } else {
#pragma unroll
      for (int i = 0; i < 16; i++) r += (uint32_t)abs(d[i]);
      if (jb.kind == KVZ_CUDA_COST_SAD_NXN) r >>= sh;
    }
    out[mine] = r;
    return;
  }
  if (all8) {
    const int lane8 = lane & 7, grp = lane >> 3;
#pragma unroll 1
    for (int pass = 0; pass < 8; pass++) {
      const int src = pass * 4 + grp;                           // the job of this group of 8 lanes
      const uint32_t ao = __shfl_sync(0xffffffffu, jb.a_offset, src), bo = __shfl_sync(0xffffffffu, jb.b_offset, src);
      const int as = __shfl_sync(0xffffffffu, jb.a_stride, src), bs = __shfl_sync(0xffffffffu, jb.b_stride, src);
      const int kind = __shfl_sync(0xffffffffu, jb.kind, src);
      const PIX *a = base_a + ao + (size_t)lane8 * as, *b = base_b + bo + (size_t)lane8 * bs;
      int d[8], ra[8], rb[8];
      uint32_t sad = 0;
      load_row<8>(a, ra);
      load_row<8>(b, rb);
#pragma unroll
      for (int x = 0; x < 8; x++) { d[x] = ra[x] - rb[x]; sad += (uint32_t)abs(d[x]); }
      sad += __shfl_xor_sync(0xffffffffu, sad, 1);
      sad += __shfl_xor_sync(0xffffffffu, sad, 2);
      sad += __shfl_xor_sync(0xffffffffu, sad, 4);
      const uint32_t satd = satd8_rows(d, lane8);
      const uint32_t r = kind == KVZ_CUDA_COST_SATD_NXN ? satd >> sh : kind == KVZ_CUDA_COST_SAD_NXN ? sad >> sh : sad;
      if (lane8 == 0) out[first + src] = r;
    }
    return;
  }
  for (int k = 0; k < 32 && first + k < n_jobs; k++) {
    const kvz_cuda_block_job j = jobs[first + k];
    const uint32_t r = warp_job_cost(base_a, base_b, j, bitdepth, lane);
    if (lane == 0) out[first + k] = r;
  }
}

// Grid-stride over the batch with a fixed number of warps.  The first 32 jobs decide how the launch is dealt out:
// small blocks -> a warp takes groups of 32 jobs (group_cost checks every group itself); otherwise a warp per job,
// which keeps every SM busy when a batch holds few, large blocks.
template <typename PIX>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
block_cost_kernel(const PIX *__restrict__ base_a, const PIX *__restrict__ base_b,
                  const kvz_cuda_block_job *__restrict__ jobs, int n_jobs, int bitdepth, uint32_t *__restrict__ out)
{
  const int lane = threadIdx.x & 31;
  const int warp = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5), warps = gridDim.x * kWarpsPerCta;
  kvz_cuda_block_job j0 = jobs[lane < n_jobs ? lane : 0];
  const bool by_groups = n_jobs >= 32 && (__all_sync(0xffffffffu, small_job(j0, 4)) || __all_sync(0xffffffffu, small_job(j0, 8)));
  if (by_groups) {
    for (int g = warp; g * 32 < n_jobs; g += warps) group_cost(base_a, base_b, jobs, n_jobs, g * 32, bitdepth, lane, out);
  } else {
    for (int ji = warp; ji < n_jobs; ji += warps) {
      const kvz_cuda_block_job jb = jobs[ji];
      const uint32_t r = warp_job_cost(base_a, base_b, jb, bitdepth, lane);
      if (lane == 0) out[ji] = r;
    }
  }
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
  int grid = (n_jobs + kWarpsPerCta - 1) / kWarpsPerCta;                          // grid-stride beyond 16 CTAs per SM
  if (grid > 148 * 16) grid = 148 * 16;
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
