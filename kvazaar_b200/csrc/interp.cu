// kvazaar_b200/csrc/interp.cu
//
// Batched HEVC fractional-sample interpolation, one CTA per block:
//   luma   8-tap, quarter-pel : kvz_sample_quarterpel_luma / kvz_sample_14bit_quarterpel_luma
//                               (ipol-generic.c:122-190) and the candidate blocks of
//                               kvz_filter_{hpel,qpel}_blocks_{hor_ver,diag}_luma (ipol-generic.c:192-658)
//   chroma 4-tap, eighth-pel  : kvz_sample_octpel_chroma / kvz_sample_14bit_octpel_chroma (ipol-generic.c:660-728)
// plus kvz_inter_recon_bipred_blend on 14-bit samples (picture-generic.c:538-588).
// The source window is fetched with the coordinate clamped into the plane, which is what
// kvz_get_extended_block (ipol-generic.c:731-784) materialises for blocks at the frame edge.
#include "common.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kThreads = 128;

__device__ __constant__ int8_t c_luma[4][8] = {          // kvz_g_luma_filter (filter.c:54-60)
  { 0, 0, 0, 64, 0, 0, 0, 0 }, { -1, 4, -10, 58, 17, -5, 1, 0 },
  { -1, 4, -11, 40, 40, -11, 4, -1 }, { 0, 1, -5, 17, 58, -10, 4, -1 } };
__device__ __constant__ int8_t c_chroma[8][4] = {        // kvz_g_chroma_filter (filter.c:62-72)
  { 0, 64, 0, 0 }, { -2, 58, 10, -2 }, { -4, 54, 16, -2 }, { -6, 46, 28, -4 },
  { -4, 36, 36, -4 }, { -4, 28, 46, -6 }, { -2, 16, 54, -4 }, { -2, 10, 58, -2 } };

template <typename PIX>
struct Smem {
  PIX win[71 * 72];              // (h + taps - 1) x (w + taps - 1), pitch 72
  int16_t hor[71 * 64];          // horizontally filtered rows, pitch w
};

// PIX = kvz_pixel of the build: uint8_t, or uint16_t for KVZ_BIT_DEPTH > 8 (pitch in bytes either way)
template <int TAPS, typename PIX>
__device__ void interp_block(Smem<PIX> &sm, const kvz_cuda_interp_job &jb, const uint8_t *plane_bytes, int pitch, int W, int H,
                             int bitdepth, PIX *dst8, int16_t *dst16)
{
  constexpr int before = TAPS / 2 - 1;
  constexpr int fshift = TAPS == 8 ? 2 : 3;                 // quarter- / eighth-pel
  constexpr int fmask = (1 << fshift) - 1;
  const int tid = threadIdx.x;
  const int w = jb.w, h = jb.h;
  const int bx = jb.x + (jb.mv_x >> fshift), by = jb.y + (jb.mv_y >> fshift);
  const int fx = jb.mv_x & fmask, fy = jb.mv_y & fmask;
  const int ww = w + TAPS - 1, wh = h + TAPS - 1;
  for (int i = tid; i < ww * wh; i += kThreads) {
    const int r = i / ww, c = i - r * ww;
    sm.win[r * 72 + c] = reinterpret_cast<const PIX *>(plane_bytes + (size_t)clampi(by - before + r, 0, H - 1) * pitch)[clampi(bx - before + c, 0, W - 1)];
  }
  __syncthreads();
  const int8_t *fh = TAPS == 8 ? c_luma[fx] : c_chroma[fx];
  const int8_t *fv = TAPS == 8 ? c_luma[fy] : c_chroma[fy];
  for (int i = tid; i < wh * w; i += kThreads) {
    const int r = i / w, x = i - r * w;
    int s = 0;
#pragma unroll
    for (int t = 0; t < TAPS; t++) s += fh[t] * (int)sm.win[r * 72 + x + t];
    sm.hor[r * w + x] = (int16_t)(s >> (bitdepth - 8));       // ipol-generic.c:147
  }
  __syncthreads();
  const int sh = 14 - bitdepth, off = 1 << (sh - 1), pix_max = (1 << bitdepth) - 1;
  for (int i = tid; i < w * h; i += kThreads) {
    const int y = i / w, x = i - y * w;
    int s = 0;
#pragma unroll
    for (int t = 0; t < TAPS; t++) s += fv[t] * (int)sm.hor[(y + t) * w + x];
    s >>= 6;                                                   // ipol-generic.c:154
    if (jb.mode == KVZ_CUDA_INTERP_14BIT) {
      dst16[(size_t)jb.dst_offset + (size_t)y * jb.dst_stride + x] = (int16_t)s;
    } else {
      if (jb.mode == KVZ_CUDA_INTERP_FME) s = (int16_t)s;      // int16 store of the block filters (ipol-generic.c:285)
      dst8[(size_t)jb.dst_offset + (size_t)y * jb.dst_stride + x] = (PIX)clampi((s + off) >> sh, 0, pix_max);
    }
  }
}

template <typename PIX>
__global__ void __launch_bounds__(kThreads)
interp_kernel(const uint8_t *__restrict__ plane, int pitch, int W, int H, int bitdepth,
              const kvz_cuda_interp_job *__restrict__ jobs, PIX *dst8, int16_t *dst16)
{
  __shared__ Smem<PIX> sm;
  const kvz_cuda_interp_job jb = jobs[blockIdx.x];
  if (jb.chroma) interp_block<4, PIX>(sm, jb, plane, pitch, W, H, bitdepth, dst8, dst16);
  else interp_block<8, PIX>(sm, jb, plane, pitch, W, H, bitdepth, dst8, dst16);
}

template <typename PIX>
__global__ void bipred_blend_kernel(const int16_t *__restrict__ s0, const int16_t *__restrict__ s1, size_t n,
                                    int bitdepth, PIX *__restrict__ dst)
{
  const int sh = 15 - bitdepth, off = 1 << (sh - 1), pix_max = (1 << bitdepth) - 1;   // picture-generic.c:553-570
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = (PIX)clampi(((int)s0[i] + (int)s1[i] + off) >> sh, 0, pix_max);
}

}  // namespace

int launch_interp(const void *plane, int pitch, int W, int H, int bitdepth, const kvz_cuda_interp_job *jobs, int n_jobs,
                  void *dst8, int16_t *dst16, cudaStream_t stream)
{
  if (n_jobs <= 0) return 1;
  if (bitdepth > 8) interp_kernel<uint16_t><<<n_jobs, kThreads, 0, stream>>>((const uint8_t *)plane, pitch, W, H, bitdepth, jobs, (uint16_t *)dst8, dst16);
  else interp_kernel<uint8_t><<<n_jobs, kThreads, 0, stream>>>((const uint8_t *)plane, pitch, W, H, bitdepth, jobs, (uint8_t *)dst8, dst16);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_bipred_blend(const int16_t *s0, const int16_t *s1, size_t n, int bitdepth, void *dst, cudaStream_t stream)
{
  if (n == 0) return 1;
  const int grid = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  if (bitdepth > 8) bipred_blend_kernel<uint16_t><<<grid, 256, 0, stream>>>(s0, s1, n, bitdepth, (uint16_t *)dst);
  else bipred_blend_kernel<uint8_t><<<grid, 256, 0, stream>>>(s0, s1, n, bitdepth, (uint8_t *)dst);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
