// kvazaar_b200/csrc/picture_stats.cu
//
// Picture-level reductions over planes that are already resident on the device:
//   * the picture-hash SEI checksum: kvz_array_checksum (strategies/generic/nal-generic.c:58-82, callers
//     kvz_image_checksum nal.c:60-76 -> add_checksum encoder_state-bitstream.c:895-940): per plane
//     sum over (x, y) of (sample byte ^ mask), mask = (x ^ y ^ (x >> 8) ^ (y >> 8)) & 0xff, both bytes of a
//     sample at KVZ_BIT_DEPTH > 8, modulo 2^32;
//   * the squared-error sums of compute_psnr (encmain.c:101-131): per plane sum of (src - rec)^2 (the reference
//     accumulates the integers in a double, exact below 2^53, so an integer sum is the same number).
// Both are single passes over 1.5 bytes per sample of a picture: HBM-bound and, at 1080p, launch-latency-bound
// (3.1 MB); all planes of a picture go through ONE launch.  MD5 (kvz_array_md5) is a serial chain over the whole
// plane by construction and stays on the host.
#include "common.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kT = 256;
constexpr int kMaxPlanes = 3;

struct Plane { const uint8_t *a, *b; int width, height, pitch_a, pitch_b; int first_item; int y0; };
struct StatsParams { Plane pl[kMaxPlanes]; int n_planes, total_items; };

// item = 16 bytes of one row
template <typename PIX>
__device__ __forceinline__ int row_items(int width) { return (width * (int)sizeof(PIX) + 15) >> 4; }

__device__ __forceinline__ uint32_t bytes_sum(uint32_t v) { return __dp4a(v, 0x01010101u, 0u); }

template <typename PIX>
__global__ void __launch_bounds__(kT) checksum_kernel(const __grid_constant__ StatsParams p, uint32_t *out)
{
  __shared__ uint32_t part[kMaxPlanes][kT / 32];
  uint32_t acc[kMaxPlanes] = { 0, 0, 0 };
  for (int it = blockIdx.x * kT + threadIdx.x; it < p.total_items; it += gridDim.x * kT) {
    int k = 0;
#pragma unroll
    for (int q = 1; q < kMaxPlanes; q++) k += (q < p.n_planes && it >= p.pl[q].first_item);
    const Plane &pl = p.pl[k];
    const int per_row = row_items<PIX>(pl.width);
    const int local = it - pl.first_item, y = local / per_row, c = local - y * per_row;
    const int row_bytes = pl.width * (int)sizeof(PIX), b0 = c * 16;
    const uint8_t *src = pl.a + (size_t)y * pl.pitch_a + b0;
    const uint32_t ym = (uint32_t)((y + pl.y0) ^ ((y + pl.y0) >> 8)) & 0xffu;   // y0: first row of a strip of a taller plane
    uint32_t s = 0;
    if (b0 + 16 <= row_bytes && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
      const uint4 v = __ldg(reinterpret_cast<const uint4 *>(src));
      const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
      for (int i = 0; i < 4; i++) {
        uint32_t m;
        if (sizeof(PIX) == 1) {
          const uint32_t x = (uint32_t)(b0 + 4 * i);                 // x .. x+3: the low two bits count 0..3, no carry
          const uint32_t xm = ((x ^ (x >> 8)) & 0xffu) ^ ym;
          m = (xm * 0x01010101u) ^ 0x03020100u;
        } else {
          const uint32_t x = (uint32_t)((b0 >> 1) + 2 * i);          // two samples, both bytes of each
          const uint32_t xm = ((x ^ (x >> 8)) & 0xffu) ^ ym;
          m = (xm * 0x01010101u) ^ 0x01010000u;
        }
        s += bytes_sum(w[i] ^ m);
      }
    } else {
      for (int b = b0; b < min(b0 + 16, row_bytes); b++) {
        const uint32_t x = (uint32_t)(b / (int)sizeof(PIX));
        s += (uint32_t)__ldg(src + (b - b0)) ^ (((x ^ (x >> 8)) & 0xffu) ^ ym);
      }
    }
#pragma unroll
    for (int q = 0; q < kMaxPlanes; q++) acc[q] += q == k ? s : 0u;
  }
#pragma unroll
  for (int q = 0; q < kMaxPlanes; q++) {
    uint32_t v = acc[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) part[q][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < p.n_planes) {
    uint32_t v = 0;
    for (int w = 0; w < kT / 32; w++) v += part[threadIdx.x][w];
    if (v) atomicAdd(out + threadIdx.x, v);
  }
}

template <typename PIX>
__global__ void __launch_bounds__(kT) sse_kernel(const __grid_constant__ StatsParams p, unsigned long long *out)
{
  __shared__ unsigned long long part[kMaxPlanes][kT / 32];
  unsigned long long acc[kMaxPlanes] = { 0, 0, 0 };
  for (int it = blockIdx.x * kT + threadIdx.x; it < p.total_items; it += gridDim.x * kT) {
    int k = 0;
#pragma unroll
    for (int q = 1; q < kMaxPlanes; q++) k += (q < p.n_planes && it >= p.pl[q].first_item);
    const Plane &pl = p.pl[k];
    const int per_row = row_items<PIX>(pl.width);
    const int local = it - pl.first_item, y = local / per_row, c = local - y * per_row;
    const int row_bytes = pl.width * (int)sizeof(PIX), b0 = c * 16;
    const uint8_t *sa = pl.a + (size_t)y * pl.pitch_a + b0, *sb = pl.b + (size_t)y * pl.pitch_b + b0;
    uint32_t s = 0;                                           // 8 samples x 1023^2 < 2^32
    if (b0 + 16 <= row_bytes && (((reinterpret_cast<uintptr_t>(sa) | reinterpret_cast<uintptr_t>(sb)) & 15) == 0)) {
      const uint4 va = __ldg(reinterpret_cast<const uint4 *>(sa)), vb = __ldg(reinterpret_cast<const uint4 *>(sb));
      const uint32_t wa[4] = { va.x, va.y, va.z, va.w }, wb[4] = { vb.x, vb.y, vb.z, vb.w };
#pragma unroll
      for (int i = 0; i < 4; i++) {
        if (sizeof(PIX) == 1) {
          const uint32_t d = __vabsdiffu4(wa[i], wb[i]);
          s = __dp4a(d, d, s);                                // sum of the four squared byte differences
        } else {
          const int d0 = (int)(wa[i] & 0xffffu) - (int)(wb[i] & 0xffffu), d1 = (int)(wa[i] >> 16) - (int)(wb[i] >> 16);
          s += (uint32_t)(d0 * d0) + (uint32_t)(d1 * d1);
        }
      }
    } else {
      const int n = (min(b0 + 16, row_bytes) - b0) / (int)sizeof(PIX);
      for (int i = 0; i < n; i++) {
        const int d = (int)reinterpret_cast<const PIX *>(sa)[i] - (int)reinterpret_cast<const PIX *>(sb)[i];
        s += (uint32_t)(d * d);
      }
    }
#pragma unroll
    for (int q = 0; q < kMaxPlanes; q++) acc[q] += q == k ? s : 0u;
  }
#pragma unroll
  for (int q = 0; q < kMaxPlanes; q++) {
    unsigned long long v = acc[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) part[q][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < p.n_planes) {
    unsigned long long v = 0;
    for (int w = 0; w < kT / 32; w++) v += part[threadIdx.x][w];
    if (v) atomicAdd(out + threadIdx.x, v);
  }
}

int fill_params(StatsParams &p, const kvz_cuda_plane *a, const kvz_cuda_plane *b, int n, int bitdepth, const char *who)
{
  if (!a || n < 1 || n > kMaxPlanes) return fail_msg(who);
  const int px = bitdepth > 8 ? 2 : 1;
  p.n_planes = n; p.total_items = 0;
  for (int k = 0; k < kMaxPlanes; k++) p.pl[k] = Plane{ nullptr, nullptr, 0, 0, 0, 0, 0x7fffffff, 0 };
  for (int k = 0; k < n; k++) {
    if (!a[k].d_data || a[k].width <= 0 || a[k].height <= 0 || a[k].pitch < a[k].width * px) return fail_msg(who);
    if (b && (!b[k].d_data || b[k].width != a[k].width || b[k].height != a[k].height || b[k].pitch < b[k].width * px)) return fail_msg(who);
    if ((long long)a[k].height * ((a[k].width * px + 15) >> 4) + p.total_items > 0x7ffffff0LL) return fail_msg(who);
    p.pl[k].a = static_cast<const uint8_t *>(a[k].d_data);
    p.pl[k].b = b ? static_cast<const uint8_t *>(b[k].d_data) : nullptr;
    p.pl[k].width = a[k].width; p.pl[k].height = a[k].height;
    p.pl[k].pitch_a = a[k].pitch; p.pl[k].pitch_b = b ? b[k].pitch : 0;
    p.pl[k].first_item = p.total_items;
    p.total_items += a[k].height * ((a[k].width * px + 15) >> 4);
  }
  return 1;
}

int stats_grid(int items)
{
  const int want = (items + kT * 4 - 1) / (kT * 4);           // about four items per thread
  return want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want);
}

}  // namespace

// y0: picture row of the planes' first row (a strip of a taller plane; the per-call face uploads planes in strips)
int launch_picture_checksum(const kvz_cuda_plane *planes, int n_planes, int bitdepth, uint32_t *d_out, cudaStream_t stream, int y0)
{
  StatsParams p;
  if (!fill_params(p, planes, nullptr, n_planes, bitdepth, "kvz_cuda_picture_checksum: bad plane")) return 0;
  for (int k = 0; k < n_planes; k++) p.pl[k].y0 = y0;
  KVZ_CUDA_TRY(cudaMemsetAsync(d_out, 0, sizeof(uint32_t) * n_planes, stream));
  if (bitdepth > 8) checksum_kernel<uint16_t><<<stats_grid(p.total_items), kT, 0, stream>>>(p, d_out);
  else checksum_kernel<uint8_t><<<stats_grid(p.total_items), kT, 0, stream>>>(p, d_out);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_picture_sse(const kvz_cuda_plane *src, const kvz_cuda_plane *rec, int n_planes, int bitdepth,
                       unsigned long long *d_out, cudaStream_t stream)
{
  StatsParams p;
  if (!rec) return fail_msg("kvz_cuda_picture_sse: bad plane");
  if (!fill_params(p, src, rec, n_planes, bitdepth, "kvz_cuda_picture_sse: bad plane")) return 0;
  KVZ_CUDA_TRY(cudaMemsetAsync(d_out, 0, sizeof(unsigned long long) * n_planes, stream));
  if (bitdepth > 8) sse_kernel<uint16_t><<<stats_grid(p.total_items), kT, 0, stream>>>(p, d_out);
  else sse_kernel<uint8_t><<<stats_grid(p.total_items), kT, 0, stream>>>(p, d_out);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
