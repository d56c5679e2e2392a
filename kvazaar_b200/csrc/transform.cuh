// kvazaar_b200/csrc/transform.cuh -- device building blocks of the HEVC integer transforms
// and the scalar quantiser, shared by the batch kernels and the fused residual kernel.
//
// Transforms (dct-generic.c:26-108 matrices, :206-617 passes): every N-point DCT basis is a
// row subset of the 32-point one, M_N[k][n] = M_32[k * 32/N][n], so one 32x32 table serves
// N = 4..32; the 4x4 DST has its own 4x4 matrix.  A pass is a plain matrix product in exact
// int32 (|sum| <= 32 * 90 * 32767 < 2^31), so any summation order gives the reference's bits.
#pragma once
#include <stdint.h>
#include "common.cuh"

namespace kvz {

// cos table: round(64*sqrt(2)*cos(m*pi/64)) as standardised for HEVC, m = 0..32 (m = 0 and 16 -> 64)
static __device__ __constant__ int8_t c_hevc_cos_q[33] = {
  64, 90, 90, 90, 89, 88, 87, 85, 83, 82, 80, 78, 75, 73, 70, 67, 64,
  61, 57, 54, 50, 46, 43, 38, 36, 31, 25, 22, 18, 13, 9, 4, 0 };
static __device__ __constant__ int8_t c_dst4[4][4] = {           // dct-generic.c:26-32 (kvz_g_dst_4)
  { 29, 55, 74, 84 }, { 74, 74, 0, -74 }, { 84, -29, -74, 55 }, { 55, -84, 74, -29 } };

// A group of T threads working on one transform block: the whole CTA (T = blockDim.x) or one warp (T = 32)
template <int T>
__device__ __forceinline__ void group_sync()
{
  if (T == 32) __syncwarp(); else __syncthreads();
}

template <int MAXN>
struct TransformTablesT {
  int16_t m[MAXN * 33];    // M_N[k][n] at m[k * 33 + n] for the N of this group (pitch 33: conflict-free columns)
};
using TransformTables = TransformTablesT<32>;

// fills t.m with the NxN matrix: M_N[k][n] = cos((32/N) k (2n+1) pi/64) scaled, or the 4x4 DST
template <typename TT>
__device__ inline void load_transform_matrix(TT &t, int N, bool dst4, int tid, int nthreads)
{
  for (int i = tid; i < N * N; i += nthreads) {
    const int k = i / N, n = i - k * N;
    int v;
    if (dst4) v = c_dst4[k][n];
    else {
      int m = ((32 / N) * k * (2 * n + 1)) & 127;
      if (m > 64) m = 128 - m;
      v = m <= 32 ? c_hevc_cos_q[m] : -c_hevc_cos_q[64 - m];
    }
    t.m[k * 33 + n] = (int16_t)v;
  }
}

// One forward pass over a block held in shared memory with row pitch P (int16):
//   out[k * P + j] = (int16)((sum_n M[k][n] * in[j * P + n] + rnd) >> shift)      (dct-generic.c:258, 571-572)
template <int P, typename TT>
__device__ __forceinline__ void fwd_pass_smem(const TT &t, int N, const int16_t *in, int16_t *out, int shift,
                                              int tid, int nthreads)
{
  const int rnd = 1 << (shift - 1);
  for (int idx = tid; idx < N * N; idx += nthreads) {
    const int k = idx / N, j = idx - k * N;
    int s = 0;
    for (int n = 0; n < N; n++) s += (int)t.m[k * 33 + n] * (int)in[j * P + n];
    out[k * P + j] = (int16_t)((s + rnd) >> shift);
  }
}

// One inverse pass:  out[j * P + n] = clip16((sum_k M[k][n] * in[k * P + j] + rnd) >> shift)   (dct-generic.c:582-583)
template <int P, typename TT>
__device__ __forceinline__ void inv_pass_smem(const TT &t, int N, const int16_t *in, int16_t *out, int shift,
                                              int tid, int nthreads)
{
  const int rnd = 1 << (shift - 1);
  for (int idx = tid; idx < N * N; idx += nthreads) {
    const int j = idx / N, n = idx - j * N;
    int s = 0;
    for (int k = 0; k < N; k++) s += (int)t.m[k * 33 + n] * (int)in[k * P + j];
    s = (s + rnd) >> shift;
    out[j * P + n] = (int16_t)clampi(s, -32768, 32767);
  }
}

__device__ __forceinline__ int ilog2_pow2(int n) { return 31 - __clz(n); }

// kvz_get_scaled_qp (transform.c:129-143) with kvz_g_chroma_scale (HEVC table 8-10)
static __device__ __constant__ uint8_t c_chroma_scale[58] = {
  0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25,
  26, 27, 28, 29, 29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37, 38, 39, 40, 41, 42,
  43, 44, 45, 46, 47, 48, 49, 50, 51 };

__device__ inline int scaled_qp(int type, int qp, int qp_offset)
{
  if (type == 0) return qp + qp_offset;
  const int q = clampi(qp, -qp_offset, 57);
  return q < 0 ? q + qp_offset : c_chroma_scale[q] + qp_offset;
}

// Scan tables of a (1 << lg)^2 block in shared memory: pos[i] = raster position of scan index i,
// kvz_g_sig_last_scan[scan_idx][lg - 1] (tables.c): 4x4 coefficient groups, the groups and the
// samples inside a group both in `scan_idx` order (0 up-right diagonal, 1 horizontal, 2 vertical).
template <int MAXN>
struct ScanTableT {
  uint16_t pos[MAXN * MAXN];
  uint8_t gx[(MAXN / 4) * (MAXN / 4)], gy[(MAXN / 4) * (MAXN / 4)], px[16], py[16];
};
using ScanTable = ScanTableT<32>;

__device__ inline void square_scan(int scan_idx, int n, uint8_t *xs, uint8_t *ys)
{
  int c = 0;
  if (scan_idx == 0) {
    for (int d = 0; d <= 2 * (n - 1); d++)
      for (int x = max(0, d - (n - 1)); x <= min(d, n - 1); x++) { xs[c] = (uint8_t)x; ys[c] = (uint8_t)(d - x); c++; }
  } else if (scan_idx == 1) {
    for (int y = 0; y < n; y++) for (int x = 0; x < n; x++) { xs[c] = (uint8_t)x; ys[c] = (uint8_t)y; c++; }
  } else {
    for (int x = 0; x < n; x++) for (int y = 0; y < n; y++) { xs[c] = (uint8_t)x; ys[c] = (uint8_t)y; c++; }
  }
}

// all T threads of the group call this; contains two group barriers
template <int T, typename ST>
__device__ inline void build_scan_table_t(ST &t, int scan_idx, int lg, int tid)
{
  const int n = 1 << lg;
  if (tid == 0) square_scan(scan_idx, n >> 2, t.gx, t.gy);
  if (tid == 32 % T) square_scan(scan_idx, 4, t.px, t.py);
  group_sync<T>();
  for (int i = tid; i < n * n; i += T) {
    const int g = i >> 4, k = i & 15;
    t.pos[i] = (uint16_t)((t.gy[g] * 4 + t.py[k]) * n + t.gx[g] * 4 + t.px[k]);
  }
  group_sync<T>();
}

// CTA-wide form (nthreads == blockDim.x, 32 or 128)
__device__ inline void build_scan_table(ScanTable &t, int scan_idx, int lg, int tid, int nthreads)
{
  if (nthreads == 32) build_scan_table_t<32>(t, scan_idx, lg, tid);
  else build_scan_table_t<128>(t, scan_idx, lg, tid);
}

}  // namespace kvz
