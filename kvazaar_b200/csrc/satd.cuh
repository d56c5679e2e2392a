// kvazaar_b200/csrc/satd.cuh -- Hadamard building blocks shared by the cost kernels
// (picture-generic.c:148-183 satd_4x4, :201-327 satd_8x8 sub-block)
#pragma once
#include <stdint.h>

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
    const bool upper = (lane8 & o) != 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const int other = __shfl_xor_sync(0xffffffffu, d[i], o);
      d[i] = upper ? other - d[i] : d[i] + other;
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

}  // namespace kvz
