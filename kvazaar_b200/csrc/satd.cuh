// kvazaar_b200/csrc/satd.cuh -- Hadamard building blocks shared by the cost kernels
// (picture-generic.c:148-183 satd_4x4, :201-327 satd_8x8 sub-block)
#pragma once
#include <stdint.h>
#include "common.cuh"

namespace kvz {

__device__ __forceinline__ uint32_t vsad4(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t d;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// 8-point Hadamard of one row in registers
__device__ __forceinline__ void hadamard8(int (&d)[8])
{
#pragma unroll
  for (int len = 1; len < 8; len <<= 1) {
#pragma unroll
    for (int i = 0; i < 8; i += 2 * len) {
#pragma unroll
      for (int k = i; k < i + len; k++) {
        const int p = d[k], q = d[k + len];
        d[k] = p + q;
        d[k + len] = p - q;
      }
    }
  }
}

// sum |H D H^T| of an 8x8 difference block held one row per lane by 8 consecutive lanes;
// every lane returns the block's SATD = (sum + 2) >> 2 (picture-generic.c:321-327)
__device__ __forceinline__ uint32_t satd8_rows(int (&d)[8], int lane8)
{
  hadamard8(d);
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) {
    const int sgn = (lane8 & o) ? -1 : 1;            // upper lane of the pair: other - d, lower: other + d
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const int other = __shfl_xor_sync(0xffffffffu, d[i], o);
      d[i] = other + sgn * d[i];
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += (uint32_t)abs(d[i]);
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  return (s + 2) >> 2;
}

// 4x4 SATD of a difference block held entirely by one thread: (sum |H4 D H4| + 1) >> 1
// (picture-generic.c:148-183)
__device__ __forceinline__ uint32_t satd4_thread(int (&d)[16])
{
#pragma unroll
  for (int r = 0; r < 4; r++) {
    const int a = d[4 * r] + d[4 * r + 1], b = d[4 * r] - d[4 * r + 1];
    const int c = d[4 * r + 2] + d[4 * r + 3], e = d[4 * r + 2] - d[4 * r + 3];
    d[4 * r] = a + c; d[4 * r + 1] = b + e; d[4 * r + 2] = a - c; d[4 * r + 3] = b - e;
  }
  uint32_t s = 0;
#pragma unroll
  for (int c = 0; c < 4; c++) {
    const int a = d[c] + d[4 + c], b = d[c] - d[4 + c];
    const int g = d[8 + c] + d[12 + c], e = d[8 + c] - d[12 + c];
    s += (uint32_t)(abs(a + g) + abs(b + e) + abs(a - g) + abs(b - e));
  }
  return (s + 1) >> 1;
}

// ---- warp-level costs over pixel fetchers ----

__device__ __forceinline__ uint32_t warp_sum(uint32_t v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// pixel fetchers: plain strided block, or plane with the coordinate clamped into it
// (PIX = uint8_t for the 8-bit build of the reference, uint16_t for KVZ_BIT_DEPTH 10; strides in samples)
template <typename PIX>
struct StridedBlockT {
  const PIX *p;
  int stride;
  __device__ __forceinline__ int operator()(int x, int y) const { return p[(size_t)y * stride + x]; }
};
using StridedBlock = StridedBlockT<uint8_t>;

template <typename PIX>
struct ClampedPlaneT {
  const PIX *p;
  int pitch, w, h, x0, y0;
  __device__ __forceinline__ int operator()(int x, int y) const
  {
    return p[(size_t)clampi(y0 + y, 0, h - 1) * pitch + clampi(x0 + x, 0, w - 1)];
  }
};
using ClampedPlane = ClampedPlaneT<uint8_t>;

template <typename FA, typename FB>
__device__ inline uint32_t warp_sad(const FA &a, const FB &b, int w, int h, int lane)
{
  uint32_t s = 0;
  for (int i = lane; i < w * h; i += 32) {
    const int y = i / w, x = i - y * w;
    s += (uint32_t)abs(a(x, y) - b(x, y));
  }
  return warp_sum(s);
}

template <typename FA, typename FB>
__device__ inline uint32_t warp_ssd(const FA &a, const FB &b, int w, int h, int lane)
{
  uint32_t s = 0;
  for (int i = lane; i < w * h; i += 32) {
    const int y = i / w, x = i - y * w;
    const int d = a(x, y) - b(x, y);
    s += (uint32_t)(d * d);
  }
  return warp_sum(s);
}

// sum of 8x8 SATDs over the tiles of the area [x0, x0+w) x [y0, y0+h) (w, h multiples of 8)
template <typename FA, typename FB>
__device__ inline uint32_t warp_satd8_area(const FA &a, const FB &b, int x0, int y0, int w, int h, int lane)
{
  const int tx_n = w >> 3, tiles = tx_n * (h >> 3);
  const int lane8 = lane & 7, sub = lane >> 3;
  uint32_t total = 0;
  for (int t0 = 0; t0 < tiles; t0 += 4) {
    const int t = t0 + sub;
    const bool on = t < tiles;
    const int ty = on ? t / tx_n : 0, tx = on ? t - ty * tx_n : 0;
    int d[8];
#pragma unroll
    for (int x = 0; x < 8; x++)
      d[x] = a(x0 + tx * 8 + x, y0 + ty * 8 + lane8) - b(x0 + tx * 8 + x, y0 + ty * 8 + lane8);
    const uint32_t s = satd8_rows(d, lane8);
    if (on && lane8 == 0) total += s;
  }
  return warp_sum(total);
}

// sum of 4x4 SATDs over `count` blocks starting at (x0, y0) and stepping by (sx, sy)
template <typename FA, typename FB>
__device__ inline uint32_t warp_satd4_strip(const FA &a, const FB &b, int x0, int y0, int sx, int sy, int count, int lane)
{
  uint32_t total = 0;
  for (int i = lane; i < count; i += 32) {
    int d[16];
#pragma unroll
    for (int y = 0; y < 4; y++)
#pragma unroll
      for (int x = 0; x < 4; x++)
        d[4 * y + x] = a(x0 + i * sx + x, y0 + i * sy + y) - b(x0 + i * sx + x, y0 + i * sy + y);
    total += satd4_thread(d);
  }
  return warp_sum(total);
}

}  // namespace kvz
