// kvazaar_b200/csrc/transform.cu
//
// Batched residual coding primitives, one CTA per transform block:
//   * forward / inverse DCT 4..32 and the 4x4 DST     (dct-generic.c:206-617, selector strategies-dct.c:59-113)
//   * kvz_quant incl. sign-bit hiding                  (quant-generic.c:37-163, flat scaling list)
//   * kvz_dequant                                      (quant-generic.c:279-321)
//   * kvz_coeff_abs_sum                                (quant-generic.c:323-330)
//   * the fused kvz_quantize_residual                  (quant-generic.c:180-273): residual -> T -> Q ->
//     Q^-1 -> T^-1 -> reconstruction, the block never leaving shared memory.
// N <= 16 is HBM-bound; every block is read once and written once.
#include <mutex>

#include "common.cuh"
#include "transform.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kThreads = 128;
constexpr int P = 34;                 // int16 row pitch of a block in shared memory (conflict-free columns)

__device__ __constant__ int32_t c_quant_scale[6] = { 26214, 23302, 20560, 18396, 16384, 14564 };   // scalinglist.c:66
__device__ __constant__ int32_t c_inv_quant_scale[6] = { 40, 45, 51, 57, 64, 72 };                 // scalinglist.c:67

// what one thread group needs for a transform block of up to MAXN x MAXN: 13 KB for 32 (a CTA of kThreads
// threads), 2.1 KB for 8 (one warp: blocks of 4x4 and 8x8 are processed one per WARP, four per CTA)
template <int MAXN>
struct BlockSmemT {
  int16_t a[MAXN * P];
  int16_t b[MAXN * P];
  int32_t delta[MAXN * MAXN];
  TransformTablesT<MAXN> tt;
  ScanTableT<MAXN> scan;
  uint32_t acc;
  int last_cg;
  int any;
};
using BlockSmem = BlockSmemT<32>;
using SmallSmem = BlockSmemT<8>;
constexpr int kSmallN = 8;            // largest block of the warp-per-block path
constexpr int kWarps = kThreads / 32;

template <int T, typename SM>
__device__ void forward2d(SM &sm, int N, int bitdepth, int tid)
{
  // a -> b -> a      (dct-generic.c:571-572)
  const int lg = ilog2_pow2(N);
  fwd_pass_smem<P>(sm.tt, N, sm.a, sm.b, lg - 1 + (bitdepth - 8), tid, T);
  group_sync<T>();
  fwd_pass_smem<P>(sm.tt, N, sm.b, sm.a, lg + 6, tid, T);
  group_sync<T>();
}

template <int T, typename SM>
__device__ void inverse2d(SM &sm, int N, int bitdepth, int tid)
{
  // a -> b -> a      (dct-generic.c:582-583)
  inv_pass_smem<P>(sm.tt, N, sm.a, sm.b, 7, tid, T);
  group_sync<T>();
  inv_pass_smem<P>(sm.tt, N, sm.b, sm.a, 12 - (bitdepth - 8), tid, T);
  group_sync<T>();
}

// kvz_quant (quant-generic.c:37-163): coefficients in sm.a, levels to sm.b.
// The scan table must have been built when cfg.signhide is set.
template <int T, typename SM>
__device__ void quant_block(SM &sm, const kvz_cuda_quant_cfg &c, int N, int type, int tid)
{
  const int lg = ilog2_pow2(N);
  const int qp_s = scaled_qp(type, c.qp, (c.bitdepth - 8) * 6);
  const long long qc = c_quant_scale[qp_s % 6];
  const int q_bits = 14 + qp_s / 6 + (15 - c.bitdepth - lg);
  const int add = (c.slice_is_intra ? 171 : 85) << (q_bits - 9);
  if (tid == 0) { sm.acc = 0; sm.last_cg = -1; }
  group_sync<T>();
  uint32_t my_sum = 0;
  for (int i = tid; i < N * N; i += T) {
    const int y = i / N, x = i - y * N;
    const int cf = sm.a[y * P + x];
    const long long prod = (long long)abs(cf) * qc;
    int level = (int)((prod + add) >> q_bits);
    my_sum += (uint32_t)level;
    sm.delta[i] = (int)((prod - ((long long)level << q_bits)) >> (q_bits - 8));
    if (cf < 0) level = -level;
    sm.b[y * P + x] = (int16_t)clampi(level, -32768, 32767);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) my_sum += __shfl_xor_sync(0xffffffffu, my_sum, o);
  if ((tid & 31) == 0) atomicAdd(&sm.acc, my_sum);
  group_sync<T>();
  if (!c.signhide || sm.acc < 2) return;          // uniform

  // sign-bit hiding: one thread per 16-coefficient group (quant-generic.c:69-162)
  const int n_cg = (N * N) >> 4;
  auto q_at = [&](int pos) -> int16_t & { return sm.b[(pos / N) * P + (pos % N)]; };
  auto c_at = [&](int pos) -> int { return sm.a[(pos / N) * P + (pos % N)]; };
  if constexpr (T == 32) {
    // a warp owns the block (4x4 / 8x8: 1 or 4 groups): lanes 0..15 take the 16 scan positions of a group, the
    // reference's backward walk with its strict '<' (the first minimum met coming down from the top wins) becomes an
    // arg-min over (cost, 15 - n)
    const int n = tid & 15;
    const bool act = tid < 16;
    int last_cg = -1;
    for (int g = 0; g < n_cg; g++)
      if (__ballot_sync(0xffffffffu, act && q_at(sm.scan.pos[g * 16 + n]) != 0)) last_cg = g;
    for (int g = 0; g < n_cg; g++) {
      const int p = sm.scan.pos[g * 16 + n];
      const int q = q_at(p), dl = sm.delta[p], cf = c_at(p);
      const unsigned nz = __ballot_sync(0xffffffffu, act && q != 0);
      if (!nz) continue;
      const int first = __ffs(nz) - 1, last = 31 - __clz(nz);
      if (last - first < 4) continue;
      int sum = act && n >= first && n <= last ? q : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      const int signbit = __shfl_sync(0xffffffffu, q, first) > 0 ? 0 : 1;
      if (signbit == (sum & 1)) continue;
      const int top = g == last_cg ? last : 15;
      int cost = 0x7FFFFFFF, cur_change = 0;
      if (act && n <= top) {
        if (q != 0) {
          if (dl > 0) { cost = -dl; cur_change = 1; }
          else if (n == first && abs(q) == 1) { cost = 0x7FFFFFFF; }
          else { cost = dl; cur_change = -1; }
        } else if (n < first && ((cf >= 0) ? 0 : 1) != signbit) {
          cost = 0x7FFFFFFF;
        } else {
          cost = -dl; cur_change = 1;
        }
      }
      long long key = (long long)cost * 256 + (15 - n) + (act ? 0 : 64);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const long long t = __shfl_xor_sync(0xffffffffu, key, o);
        key = t < key ? t : key;
      }
      const int wn = 15 - (int)(key & 63);                       // (cost * 256 + r) & 63 == r for r < 64, any sign of cost
      const int best = __shfl_sync(0xffffffffu, cost, wn);
      if (best == 0x7FFFFFFF) continue;
      if (tid == wn) {
        int change = cur_change;
        int16_t &qq = q_at(p);
        if (qq == 32767 || qq == -32768) change = -1;
        qq = (int16_t)(cf >= 0 ? qq + change : qq - change);
      }
    }
    group_sync<T>();
  } else {
  for (int g = tid; g < n_cg; g += T) {
    bool any = false;
    for (int n = 0; n < 16; n++) any |= q_at(sm.scan.pos[g * 16 + n]) != 0;
    if (any) atomicMax(&sm.last_cg, g);
  }
  group_sync<T>();
  for (int g = tid; g < n_cg; g += T) {
    const uint16_t *sc = sm.scan.pos + g * 16;
    int first = 16, last = -1, sum = 0;
    for (int n = 15; n >= 0; n--) if (q_at(sc[n])) { last = n; break; }
    for (int n = 0; n < 16; n++) if (q_at(sc[n])) { first = n; break; }
    for (int n = first; n <= last; n++) sum += q_at(sc[n]);
    if (last - first < 4) continue;
    const bool is_last_group = g == sm.last_cg;
    const int signbit = q_at(sc[first]) > 0 ? 0 : 1;
    if (signbit == (sum & 1)) continue;
    int best = 0x7FFFFFFF, pos = -1, change = 0, cur_change = 0;
    for (int n = is_last_group ? last : 15; n >= 0; n--) {
      const int p = sc[n];
      const int q = q_at(p), dl = sm.delta[p];
      int cost = 0x7FFFFFFF;
      if (q != 0) {
        if (dl > 0) { cost = -dl; cur_change = 1; }
        else if (n == first && abs(q) == 1) { cost = 0x7FFFFFFF; }
        else { cost = dl; cur_change = -1; }
      } else if (n < first && ((c_at(p) >= 0) ? 0 : 1) != signbit) {
        cost = 0x7FFFFFFF;
      } else {
        cost = -dl; cur_change = 1;
      }
      if (cost < best) { best = cost; change = cur_change; pos = p; }
    }
    if (pos < 0) continue;
    int16_t &qq = q_at(pos);
    if (qq == 32767 || qq == -32768) change = -1;
    qq = (int16_t)(c_at(pos) >= 0 ? qq + change : qq - change);
  }
  group_sync<T>();
  }
}

// kvz_dequant (quant-generic.c:279-321, flat list): levels in sm.b -> coefficients in sm.a
template <int T, typename SM>
__device__ void dequant_block(SM &sm, const kvz_cuda_quant_cfg &c, int N, int type, int tid)
{
  const int lg = ilog2_pow2(N);
  const int qp_s = scaled_qp(type, c.qp, (c.bitdepth - 8) * 6);
  const int shift = 20 - 14 - (15 - c.bitdepth - lg);
  const int scale = c_inv_quant_scale[qp_s % 6] << (qp_s / 6);
  const int add = 1 << (shift - 1);
  for (int i = tid; i < N * N; i += T) {
    const int y = i / N, x = i - y * N;
    const int v = ((int)sm.b[y * P + x] * scale + add) >> shift;
    sm.a[y * P + x] = (int16_t)clampi(v, -32768, 32767);
  }
  group_sync<T>();
}

template <int T>
__device__ __forceinline__ void load_block(int16_t *dst, const int16_t *src, int N, int tid)
{
  for (int i = tid; i < N * N; i += T) dst[(i / N) * P + (i % N)] = src[i];
}

template <int T>
__device__ __forceinline__ void store_block(int16_t *dst, const int16_t *src, int N, int tid)
{
  for (int i = tid; i < N * N; i += T) dst[i] = src[(i / N) * P + (i % N)];
}

// ---- batch kernels ----
// Blocks of 16x16 and 32x32 get the whole CTA (kThreads threads); blocks of 4x4 and 8x8 get one WARP each, four
// per CTA, with a 2 KB shared-memory record per warp and warp barriers only -- a CTA per 16-sample block would
// spend its time on barriers and idle lanes (measured: 129600 4x4 DCTs 0.24 ms with a CTA per block).

__global__ void __launch_bounds__(kThreads)
transform_batch_kernel(int N, int dst4, int inverse, int bitdepth, const int16_t *__restrict__ in,
                       int16_t *__restrict__ out, int count)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  if (N <= kSmallN) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    SmallSmem &sm = reinterpret_cast<SmallSmem *>(smem_raw)[warp];
    load_transform_matrix(sm.tt, N, dst4 != 0, lane, 32);
    for (int blk = blockIdx.x * kWarps + warp; blk < count; blk += gridDim.x * kWarps) {
      __syncwarp();
      load_block<32>(sm.a, in + (size_t)blk * N * N, N, lane);
      __syncwarp();
      if (inverse) inverse2d<32>(sm, N, bitdepth, lane);
      else forward2d<32>(sm, N, bitdepth, lane);
      store_block<32>(out + (size_t)blk * N * N, sm.a, N, lane);
    }
    return;
  }
  BlockSmem &sm = *reinterpret_cast<BlockSmem *>(smem_raw);
  const int tid = threadIdx.x;
  load_transform_matrix(sm.tt, N, dst4 != 0, tid, kThreads);
  for (int blk = blockIdx.x; blk < count; blk += gridDim.x) {
    __syncthreads();
    load_block<kThreads>(sm.a, in + (size_t)blk * N * N, N, tid);
    __syncthreads();
    if (inverse) inverse2d<kThreads>(sm, N, bitdepth, tid);
    else forward2d<kThreads>(sm, N, bitdepth, tid);
    store_block<kThreads>(out + (size_t)blk * N * N, sm.a, N, tid);
  }
}

template <int T, typename SM>
__device__ __forceinline__ void quant_one(SM &sm, const kvz_cuda_quant_cfg &c, int N, int type, int dequant,
                                          const int16_t *in, int16_t *out, int tid)
{
  group_sync<T>();
  if (dequant) {
    load_block<T>(sm.b, in, N, tid);
    group_sync<T>();
    dequant_block<T>(sm, c, N, type, tid);
    store_block<T>(out, sm.a, N, tid);
  } else {
    load_block<T>(sm.a, in, N, tid);
    group_sync<T>();
    quant_block<T>(sm, c, N, type, tid);
    group_sync<T>();
    store_block<T>(out, sm.b, N, tid);
  }
}

__global__ void __launch_bounds__(kThreads)
quant_batch_kernel(kvz_cuda_quant_cfg c, int N, int type, int scan_idx, int dequant, const int16_t *__restrict__ in,
                   int16_t *__restrict__ out, int count)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  if (N <= kSmallN) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    SmallSmem &sm = reinterpret_cast<SmallSmem *>(smem_raw)[warp];
    if (!dequant && c.signhide) build_scan_table_t<32>(sm.scan, scan_idx, ilog2_pow2(N), lane);
    for (int blk = blockIdx.x * kWarps + warp; blk < count; blk += gridDim.x * kWarps)
      quant_one<32>(sm, c, N, type, dequant, in + (size_t)blk * N * N, out + (size_t)blk * N * N, lane);
    return;
  }
  BlockSmem &sm = *reinterpret_cast<BlockSmem *>(smem_raw);
  const int tid = threadIdx.x;
  if (!dequant && c.signhide) build_scan_table_t<kThreads>(sm.scan, scan_idx, ilog2_pow2(N), tid);
  for (int blk = blockIdx.x; blk < count; blk += gridDim.x)
    quant_one<kThreads>(sm, c, N, type, dequant, in + (size_t)blk * N * N, out + (size_t)blk * N * N, tid);
}

__global__ void __launch_bounds__(256)
coeff_abs_sum_kernel(const int16_t *__restrict__ coeffs, int per_block, int count, uint32_t *__restrict__ out)
{
  const int blk = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (blk >= count) return;
  uint32_t s = 0;
  for (int i = lane; i < per_block; i += 32) s += (uint32_t)abs((int)coeffs[(size_t)blk * per_block + i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) out[blk] = s;
}

// One kvz_quantize_residual call (quant-generic.c:180-273) by a group of T threads.
// STAGE 0: the whole function with the scalar quantiser (cfg.rdoq_enable == 0).
// With RDOQ the function is three launches around rdoq_kernel (rdoq.cu):
// STAGE 1: residual -> transform, coefficients to `work` (4x4 blocks under cfg.rdoq_skip are quantised here,
//          quant-generic.c:213-215); STAGE 2: levels from coeff_out -> dequant -> inverse -> reconstruction.
// PIX = kvz_pixel of the build (uint8_t, or uint16_t for KVZ_BIT_DEPTH > 8); job offsets and strides in samples
template <int STAGE, int T, typename SM, typename PIX>
__device__ void residual_job(SM &sm, kvz_cuda_quant_cfg c, const PIX *__restrict__ ref, const PIX *__restrict__ pred,
                             PIX *__restrict__ rec, int16_t *__restrict__ coeff_out, const kvz_cuda_residual_job &jb,
                             uint8_t *has_coeff, int16_t *__restrict__ work, int rdoq_skip, int tid, int &tt_key,
                             int &scan_key)
{
  const int N = jb.width, lg = ilog2_pow2(N);
  const bool luma = jb.color == 0;
  c.qp = jb.qp;
  c.cu_is_intra = jb.cu_is_intra;
  const bool dst4 = N == 4 && luma && jb.cu_is_intra;          // strategies-dct.c:59-113
  group_sync<T>();                                             // the group's previous job is done with sm
  // the transform matrix and the scan table stay in shared memory while consecutive jobs of the group use the same
  if (!jb.use_trskip && tt_key != (N | (dst4 ? 64 : 0))) {
    load_transform_matrix(sm.tt, N, dst4, tid, T);
    tt_key = N | (dst4 ? 64 : 0);
  }
  const bool scalar_quant = STAGE == 0 || (STAGE == 1 && N == 4 && rdoq_skip);
  if (c.signhide && scalar_quant && scan_key != (N | (jb.scan_idx << 8))) {
    build_scan_table_t<T>(sm.scan, jb.scan_idx, lg, tid);
    scan_key = N | (jb.scan_idx << 8);
  }
  if (tid == 0) sm.any = 0;
  const PIX *rp = ref + jb.ref_offset, *pp = pred + jb.pred_offset;
  const int ts_shift = 15 - c.bitdepth - lg;
  int16_t *co = coeff_out + jb.coeff_offset;
  if (STAGE != 2) {
    for (int i = tid; i < N * N; i += T) {
      const int y = i / N, x = i - y * N;
      sm.a[y * P + x] = (int16_t)((int)rp[(size_t)y * jb.in_stride + x] - (int)pp[(size_t)y * jb.in_stride + x]);
    }
    group_sync<T>();
    if (jb.use_trskip) {                                           // transform.c:151-161
      for (int i = tid; i < N * N; i += T) {
        int16_t &v = sm.a[(i / N) * P + (i % N)];
        v = (int16_t)((int)v << ts_shift);
      }
      group_sync<T>();
    } else {
      forward2d<T>(sm, N, c.bitdepth, tid);
    }
    if (!scalar_quant) {                                           // STAGE 1: the coefficients go to rdoq_kernel
      int16_t *wo = work + jb.coeff_offset;
      for (int i = tid; i < N * N; i += T) wo[i] = sm.a[(i / N) * P + (i % N)];
      return;
    }
    quant_block<T>(sm, c, N, luma ? 0 : 2, tid);
    group_sync<T>();
  } else {
    for (int i = tid; i < N * N; i += T) sm.b[(i / N) * P + (i % N)] = co[i];
    group_sync<T>();
  }
  int any = 0;
  for (int i = tid; i < N * N; i += T) {
    const int16_t q = sm.b[(i / N) * P + (i % N)];
    if (STAGE != 2) co[i] = q;
    any |= q != 0;
  }
  if (STAGE == 1) return;
  if (any) sm.any = 1;
  group_sync<T>();
  PIX *op = rec + jb.rec_offset;
  const int has = sm.any;
  if (has) {
    dequant_block<T>(sm, c, N, luma ? 0 : (jb.color == 1 ? 2 : 3), tid);
    if (jb.use_trskip) {                                         // transform.c:169-181
      const int off = 1 << (ts_shift - 1);
      for (int i = tid; i < N * N; i += T) {
        int16_t &v = sm.a[(i / N) * P + (i % N)];
        v = (int16_t)(((int)v + off) >> ts_shift);
      }
      group_sync<T>();
    } else {
      inverse2d<T>(sm, N, c.bitdepth, tid);
    }
    const int pix_max = (1 << c.bitdepth) - 1;
    for (int i = tid; i < N * N; i += T) {
      const int y = i / N, x = i - y * N;
      const int16_t val = (int16_t)((int)sm.a[y * P + x] + (int)pp[(size_t)y * jb.in_stride + x]);
      op[(size_t)y * jb.out_stride + x] = (PIX)clampi(val, 0, pix_max);
    }
  } else {
    for (int i = tid; i < N * N; i += T) {
      const int y = i / N, x = i - y * N;
      op[(size_t)y * jb.out_stride + x] = pp[(size_t)y * jb.in_stride + x];
    }
  }
  if (tid == 0) *has_coeff = (uint8_t)has;
}

// a CTA takes kWarps consecutive jobs: one per warp when all of them are 4x4 / 8x8, else one after the other
template <int STAGE, typename PIX>
__global__ void __launch_bounds__(kThreads)
quantize_residual_kernel(kvz_cuda_quant_cfg c, const PIX *__restrict__ ref, const PIX *__restrict__ pred,
                         PIX *__restrict__ rec, int16_t *__restrict__ coeff_out,
                         const kvz_cuda_residual_job *__restrict__ jobs, int n_jobs, uint8_t *__restrict__ has_coeffs,
                         int16_t *__restrict__ work, int rdoq_skip, int skip_all4)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int tt_small = -1, scan_small = -1, tt_big = -1, scan_big = -1;     // what the warp's / the CTA's tables hold
  for (int j0 = blockIdx.x * kWarps; j0 < n_jobs; j0 += gridDim.x * kWarps) {
    const int nj = min(kWarps, n_jobs - j0);
    if (skip_all4) {                            // groups of 32 jobs that are all 4x4 were done by quantize_residual4_kernel
      const int g0 = j0 & ~31, li = g0 + (threadIdx.x & 31);
      const int is4 = li >= n_jobs || jobs[li].width == 4;
      if (__syncthreads_and(is4)) continue;
    }
    bool small = true;
    for (int q = 0; q < nj; q++) small &= jobs[j0 + q].width <= kSmallN;
    __syncthreads();                            // the two paths overlay the same shared memory
    if (small) {
      const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
      tt_big = scan_big = -1;
      if (warp < nj) {
        SmallSmem &sm = reinterpret_cast<SmallSmem *>(smem_raw)[warp];
        residual_job<STAGE, 32>(sm, c, ref, pred, rec, coeff_out, jobs[j0 + warp], has_coeffs + j0 + warp, work, rdoq_skip,
                                lane, tt_small, scan_small);
      }
    } else {
      BlockSmem &sm = *reinterpret_cast<BlockSmem *>(smem_raw);
      tt_small = scan_small = -1;
      for (int q = 0; q < nj; q++)
        residual_job<STAGE, kThreads>(sm, c, ref, pred, rec, coeff_out, jobs[j0 + q], has_coeffs + j0 + q, work, rdoq_skip,
                                      threadIdx.x, tt_big, scan_big);
    }
  }
}

// raster index (y * 4 + x) of the k-th coefficient of a 4x4 group in the three scans (square_scan of transform.cuh)
__host__ __device__ constexpr int cg_raster(int scan, int k)     // raster index (y * 4 + x) of the k-th coefficient of a group
{
  constexpr int diag[16] = { 0, 4, 1, 8, 5, 2, 12, 9, 6, 3, 13, 10, 7, 14, 11, 15 };
  return scan == 0 ? diag[k] : scan == 1 ? k : (k & 3) * 4 + (k >> 2);
}

// ---- kvz_quantize_residual (scalar quantiser) for 4x4 blocks: one block per THREAD ----
// A frame has more 4x4 transform blocks than anything else, and a warp per 16-sample block spends its time on barriers
// and idle lanes.  A warp takes 32 consecutive jobs; when all of them are 4x4 every thread does one: residual, DCT /
// DST / transform skip, quantiser with sign hiding over the block's one coefficient group, dequantiser, inverse and
// reconstruction, all in registers -- the same arithmetic as residual_job (transform.c:151-181, quant-generic.c:180-273).
// Groups that hold another size are left to quantize_residual_kernel (skip_all4).
template <int SCAN>
__device__ __forceinline__ bool quant4_regs(const int (&a)[16], int (&q)[16], const kvz_cuda_quant_cfg &c, int type)
{
  const int qp_s = scaled_qp(type, c.qp, (c.bitdepth - 8) * 6);
  const long long qc = c_quant_scale[qp_s % 6];
  const int q_bits = 14 + qp_s / 6 + (15 - c.bitdepth - 2);
  const int add = (c.slice_is_intra ? 171 : 85) << (q_bits - 9);
  int lv[16], dl[16];
  uint32_t neg = 0, acc = 0;
  bool any = false;
#pragma unroll
  for (int n = 0; n < 16; n++) {
    const int cf = a[cg_raster(SCAN, n)];
    const long long prod = (long long)abs(cf) * qc;
    int level = (int)((prod + add) >> q_bits);
    acc += (uint32_t)level;
    dl[n] = (int)((prod - ((long long)level << q_bits)) >> (q_bits - 8));
    if (cf < 0) { level = -level; neg |= 1u << n; }
    lv[n] = clampi(level, -32768, 32767);
    any |= lv[n] != 0;
  }
  if (c.signhide && acc >= 2) {
    int first = 16, last = -1, qfirst = 0;
#pragma unroll
    for (int n = 0; n < 16; n++) if (lv[n] != 0) last = n;
#pragma unroll
    for (int n = 15; n >= 0; n--) if (lv[n] != 0) { first = n; qfirst = lv[n]; }
    if (last - first >= 4) {
      int sum = 0;
#pragma unroll
      for (int n = 0; n < 16; n++) if (n >= first && n <= last) sum += lv[n];
      const int signbit = qfirst > 0 ? 0 : 1;
      if (signbit != (sum & 1)) {
        int best = 0x7FFFFFFF, bestn = -1, change = 0;          // the block's only group is its last one: the walk starts at `last`
#pragma unroll
        for (int n = 15; n >= 0; n--) {
          if (n > last) continue;
          int cost = 0x7FFFFFFF, cur_change = 0;
          if (lv[n] != 0) {
            if (dl[n] > 0) { cost = -dl[n]; cur_change = 1; }
            else if (n == first && abs(lv[n]) == 1) { cost = 0x7FFFFFFF; }
            else { cost = dl[n]; cur_change = -1; }
          } else if (n < first && (int)((neg >> n) & 1u) != signbit) {
            cost = 0x7FFFFFFF;
          } else {
            cost = -dl[n]; cur_change = 1;
          }
          if (cost < best) { best = cost; change = cur_change; bestn = n; }
        }
#pragma unroll
        for (int n = 0; n < 16; n++)
          if (n == bestn) {
            int ch = change;
            if (lv[n] == 32767 || lv[n] == -32768) ch = -1;
            lv[n] = (int)(int16_t)(((neg >> n) & 1u) ? lv[n] - ch : lv[n] + ch);
          }
      }
    }
  }
  any = false;
#pragma unroll
  for (int n = 0; n < 16; n++) { q[cg_raster(SCAN, n)] = lv[n]; any |= lv[n] != 0; }
  return any;
}

template <typename PIX>
__global__ void __launch_bounds__(kThreads)
quantize_residual4_kernel(kvz_cuda_quant_cfg c, const PIX *__restrict__ ref, const PIX *__restrict__ pred,
                          PIX *__restrict__ rec, int16_t *__restrict__ coeff_out,
                          const kvz_cuda_residual_job *__restrict__ jobs, int n_jobs, uint8_t *__restrict__ has_coeffs)
{
  __shared__ int16_t mats[2][16];                            // [0] the 4-point DCT, [1] the 4x4 DST
  if (threadIdx.x < 32) {
    const int k = (threadIdx.x & 15) >> 2, n = threadIdx.x & 3;
    int v;
    if (threadIdx.x >= 16) v = c_dst4[k][n];
    else {
      int q = (8 * k * (2 * n + 1)) & 127;
      if (q > 64) q = 128 - q;
      v = q <= 32 ? c_hevc_cos_q[q] : -c_hevc_cos_q[64 - q];
    }
    mats[threadIdx.x >> 4][threadIdx.x & 15] = (int16_t)v;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warps = gridDim.x * (kThreads / 32);
  for (int g = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5); g * 32 < n_jobs; g += warps) {
    const int ji = g * 32 + lane;
    const bool on = ji < n_jobs;
    kvz_cuda_residual_job jb;
    if (on) jb = jobs[ji]; else jb.width = 4;
    if (!__all_sync(0xffffffffu, jb.width == 4)) continue;
    if (!on) continue;
    c.qp = jb.qp;
    c.cu_is_intra = jb.cu_is_intra;
    const bool luma = jb.color == 0;
    const int16_t *m = mats[luma && jb.cu_is_intra ? 1 : 0];          // strategies-dct.c:59-113
    const PIX *rp = ref + jb.ref_offset, *pp = pred + jb.pred_offset;
    int a[16], p[16];
#pragma unroll
    for (int y = 0; y < 4; y++)
#pragma unroll
      for (int x = 0; x < 4; x++) {
        p[4 * y + x] = (int)pp[(size_t)y * jb.in_stride + x];
        a[4 * y + x] = (int)(int16_t)((int)rp[(size_t)y * jb.in_stride + x] - p[4 * y + x]);
      }
    const int ts_shift = 15 - c.bitdepth - 2;
    int b[16];
    if (jb.use_trskip) {                                               // transform.c:151-161
#pragma unroll
      for (int i = 0; i < 16; i++) a[i] = (int)(int16_t)(a[i] << ts_shift);
    } else {                                                           // dct-generic.c:258, 571-572
      const int s1 = 1 + (c.bitdepth - 8), r1 = 1 << (s1 - 1);
#pragma unroll
      for (int k = 0; k < 4; k++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
          int s = 0;
#pragma unroll
          for (int n = 0; n < 4; n++) s += (int)m[k * 4 + n] * a[j * 4 + n];
          b[k * 4 + j] = (int)(int16_t)((s + r1) >> s1);
        }
#pragma unroll
      for (int k = 0; k < 4; k++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
          int s = 0;
#pragma unroll
          for (int n = 0; n < 4; n++) s += (int)m[k * 4 + n] * b[j * 4 + n];
          a[k * 4 + j] = (int)(int16_t)((s + 128) >> 8);
        }
    }
    int q[16];
    const int qtype = luma ? 0 : 2;
    bool any;
    if (jb.scan_idx == 0) any = quant4_regs<0>(a, q, c, qtype);
    else if (jb.scan_idx == 1) any = quant4_regs<1>(a, q, c, qtype);
    else any = quant4_regs<2>(a, q, c, qtype);
    int16_t *co = coeff_out + jb.coeff_offset;
#pragma unroll
    for (int i = 0; i < 16; i++) co[i] = (int16_t)q[i];
    PIX *op = rec + jb.rec_offset;
    if (any) {
      const int dtype = luma ? 0 : (jb.color == 1 ? 2 : 3);           // kvz_dequant (quant-generic.c:279-321, flat list)
      const int qp_s = scaled_qp(dtype, c.qp, (c.bitdepth - 8) * 6);
      const int shift = 20 - 14 - (15 - c.bitdepth - 2);
      const int scale = c_inv_quant_scale[qp_s % 6] << (qp_s / 6);
      const int dadd = 1 << (shift - 1);
#pragma unroll
      for (int i = 0; i < 16; i++) a[i] = clampi((q[i] * scale + dadd) >> shift, -32768, 32767);
      if (jb.use_trskip) {                                             // transform.c:169-181
        const int off = 1 << (ts_shift - 1);
#pragma unroll
        for (int i = 0; i < 16; i++) a[i] = (int)(int16_t)((a[i] + off) >> ts_shift);
      } else {                                                         // dct-generic.c:582-583
        const int s2 = 12 - (c.bitdepth - 8);
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
          for (int n = 0; n < 4; n++) {
            int s = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) s += (int)m[k * 4 + n] * a[k * 4 + j];
            b[j * 4 + n] = clampi((s + 64) >> 7, -32768, 32767);
          }
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
          for (int n = 0; n < 4; n++) {
            int s = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) s += (int)m[k * 4 + n] * b[k * 4 + j];
            a[j * 4 + n] = clampi((s + (1 << (s2 - 1))) >> s2, -32768, 32767);
          }
      }
      const int pix_max = (1 << c.bitdepth) - 1;
#pragma unroll
      for (int y = 0; y < 4; y++)
#pragma unroll
        for (int x = 0; x < 4; x++)
          op[(size_t)y * jb.out_stride + x] = (PIX)clampi((int)(int16_t)(a[4 * y + x] + p[4 * y + x]), 0, pix_max);
    } else {
#pragma unroll
      for (int y = 0; y < 4; y++)
#pragma unroll
        for (int x = 0; x < 4; x++) op[(size_t)y * jb.out_stride + x] = (PIX)p[4 * y + x];
    }
    has_coeffs[ji] = (uint8_t)(any ? 1 : 0);
  }
}

bool valid_n(int n) { return n == 4 || n == 8 || n == 16 || n == 32; }

}  // namespace

static int g_sms = 0;
static int sm_count()
{
  if (!g_sms) { int d = 0; cudaGetDevice(&d); cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, d); }
  return g_sms;
}

static int ilog2_host(int n) { int l = 0; while ((1 << l) < n) l++; return l; }

// persistent CTAs: 4 per SM; a CTA serves kWarps blocks at a time when they are 4x4 / 8x8
static int batch_grid(int n, int count)
{
  const int units = n <= kSmallN ? (count + kWarps - 1) / kWarps : count;
  return units < sm_count() * 4 ? units : sm_count() * 4;
}

// persistent CTAs over groups of kWarps jobs, 8 per SM
static int residual_grid(int n_jobs)
{
  const int groups = (n_jobs + kWarps - 1) / kWarps;
  return groups < sm_count() * 8 ? groups : sm_count() * 8;
}

// ---- 4x4 and 8x8 transforms of a batch: one block per THREAD ----
// A frame's worth of small blocks is where a warp (let alone a CTA) per block idles: here the block lives in the
// thread's registers, both passes are fully unrolled exact int32 products (the same sums as fwd_pass_smem /
// inv_pass_smem, dct-generic.c:258, 571-583), the NxN matrix is read from shared memory (every lane reads the same
// entry: a broadcast), and the block moves as 16-byte vectors.
template <int N>
__global__ void __launch_bounds__(kThreads)
transform_small_kernel(int dst4, int inverse, int bitdepth, const int16_t *__restrict__ in, int16_t *__restrict__ out, int count)
{
  __shared__ int16_t m[N * N];
  for (int i = threadIdx.x; i < N * N; i += kThreads) {
    const int k = i / N, n = i - k * N;
    int v;
    if (dst4) v = c_dst4[k][n];
    else {
      int q = ((32 / N) * k * (2 * n + 1)) & 127;
      if (q > 64) q = 128 - q;
      v = q <= 32 ? c_hevc_cos_q[q] : -c_hevc_cos_q[64 - q];
    }
    m[i] = (int16_t)v;
  }
  __syncthreads();
  constexpr int lg = N == 4 ? 2 : 3;
  constexpr int kVec = N * N / 8;                       // 16-byte vectors per block
  for (int blk = blockIdx.x * kThreads + threadIdx.x; blk < count; blk += gridDim.x * kThreads) {
    int a[N * N], b[N * N];
    const uint4 *src = reinterpret_cast<const uint4 *>(in + (size_t)blk * N * N);
#pragma unroll
    for (int v = 0; v < kVec; v++) {
      const uint4 q = __ldg(src + v);
      const uint32_t w[4] = { q.x, q.y, q.z, q.w };
#pragma unroll
      for (int i = 0; i < 4; i++) {
        a[v * 8 + 2 * i] = (int)(int16_t)(w[i] & 0xFFFFu);
        a[v * 8 + 2 * i + 1] = (int)(int16_t)(w[i] >> 16);
      }
    }
    if (!inverse) {
      const int s1 = lg - 1 + (bitdepth - 8), s2 = lg + 6;
      const int r1 = s1 > 0 ? 1 << (s1 - 1) : 0, r2 = 1 << (s2 - 1);
#pragma unroll
      for (int k = 0; k < N; k++)
#pragma unroll
        for (int j = 0; j < N; j++) {
          int s = 0;
#pragma unroll
          for (int n = 0; n < N; n++) s += (int)m[k * N + n] * a[j * N + n];
          b[k * N + j] = (int)(int16_t)((s + r1) >> s1);
        }
#pragma unroll
      for (int k = 0; k < N; k++)
#pragma unroll
        for (int j = 0; j < N; j++) {
          int s = 0;
#pragma unroll
          for (int n = 0; n < N; n++) s += (int)m[k * N + n] * b[j * N + n];
          a[k * N + j] = (int)(int16_t)((s + r2) >> s2);
        }
    } else {
      const int s2 = 12 - (bitdepth - 8);
#pragma unroll
      for (int j = 0; j < N; j++)
#pragma unroll
        for (int n = 0; n < N; n++) {
          int s = 0;
#pragma unroll
          for (int k = 0; k < N; k++) s += (int)m[k * N + n] * a[k * N + j];
          b[j * N + n] = clampi((s + 64) >> 7, -32768, 32767);
        }
#pragma unroll
      for (int j = 0; j < N; j++)
#pragma unroll
        for (int n = 0; n < N; n++) {
          int s = 0;
#pragma unroll
          for (int k = 0; k < N; k++) s += (int)m[k * N + n] * b[k * N + j];
          a[j * N + n] = clampi((s + (1 << (s2 - 1))) >> s2, -32768, 32767);
        }
    }
    uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)blk * N * N);
#pragma unroll
    for (int v = 0; v < kVec; v++) {
      uint32_t w[4];
#pragma unroll
      for (int i = 0; i < 4; i++) w[i] = ((uint32_t)a[v * 8 + 2 * i] & 0xFFFFu) | ((uint32_t)a[v * 8 + 2 * i + 1] << 16);
      dst[v] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
}

// ---- 16x16 and 32x32 transforms of a batch: one block row per LANE ----
// N lanes hold the N rows of a block in registers (two 16x16 blocks or one 32x32 block per warp).  A pass is N*N
// multiply-adds per lane whose matrix entries are immediates (hevc_m<N>(k, n) folds after unrolling); between the
// passes the block is transposed through a pitch N + 2 shared-memory tile (conflict-free both ways), with warp
// barriers only.  Same sums, shifts and clips as fwd_pass_smem / inv_pass_smem (dct-generic.c:258, 571-583).
__host__ __device__ constexpr int hevc_cos_at(int m)
{
  switch (m) {
    case 0: return 64; case 1: return 90; case 2: return 90; case 3: return 90; case 4: return 89; case 5: return 88;
    case 6: return 87; case 7: return 85; case 8: return 83; case 9: return 82; case 10: return 80; case 11: return 78;
    case 12: return 75; case 13: return 73; case 14: return 70; case 15: return 67; case 16: return 64; case 17: return 61;
    case 18: return 57; case 19: return 54; case 20: return 50; case 21: return 46; case 22: return 43; case 23: return 38;
    case 24: return 36; case 25: return 31; case 26: return 25; case 27: return 22; case 28: return 18; case 29: return 13;
    case 30: return 9; case 31: return 4; default: return 0;
  }
}
template <int N>
__host__ __device__ constexpr int hevc_m(int k, int n)
{
  int q = ((32 / N) * k * (2 * n + 1)) & 127;
  if (q > 64) q = 128 - q;
  return q <= 32 ? hevc_cos_at(q) : -hevc_cos_at(64 - q);
}

template <int N, bool INV>
__global__ void __launch_bounds__(kThreads)
transform_rows_kernel(int bitdepth, const int16_t *__restrict__ in, int16_t *__restrict__ out, int count)
{
  constexpr int kPerWarp = 32 / N, kPitch = N + 2, lg = N == 16 ? 4 : 5;
  __shared__ int16_t tiles[kWarps][kPerWarp][N * kPitch];
  // 32x32: 2 x 1024 multiply-adds with immediate operands would be 32 KB of straight-line code per pass pair (measured:
  // 5.5 ms per 32768 blocks, instruction fetch bound); the matrix rows come from shared memory instead (every lane
  // reads the same 16 bytes: a broadcast) in a loop over the output index.  INV reads columns: the table is transposed.
  __shared__ __align__(16) int32_t mtab[N == 32 ? 32 * 32 : 4];
  if (N == 32) {
    for (int i = threadIdx.x; i < 32 * 32; i += kThreads) {
      const int r = i >> 5, c = i & 31;
      mtab[i] = INV ? hevc_m<32>(c, r) : hevc_m<32>(r, c);      // forward: row k of M; inverse: column n of M
    }
    __syncthreads();
  }
  auto dot = [&](int o, const int (&v)[N]) {                   // forward: sum_n M[o][n] v[n]; inverse: sum_k M[k][o] v[k]
    int s = 0;
    if (N == 32) {
      const int4 *row = reinterpret_cast<const int4 *>(mtab + o * 32);
#pragma unroll
      for (int q = 0; q < N / 4; q++) {
        const int4 mm = row[q];
        s += mm.x * v[4 * q] + mm.y * v[4 * q + 1] + mm.z * v[4 * q + 2] + mm.w * v[4 * q + 3];
      }
    } else {
#pragma unroll
      for (int i = 0; i < N; i++) s += (INV ? hevc_m<N>(i, o) : hevc_m<N>(o, i)) * v[i];
    }
    return s;
  };
  constexpr int kUnroll = N == 32 ? 2 : N;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, sub = lane / N, j = lane % N;
  int16_t *t = tiles[warp][sub];
  for (int base = (blockIdx.x * kWarps + warp) * kPerWarp; base < count; base += gridDim.x * kWarps * kPerWarp) {
    const int blk = base + sub;
    const bool on = blk < count;
    int x[N], y[N];
    {
      const uint4 *src = reinterpret_cast<const uint4 *>(in + ((size_t)(on ? blk : base) * N + j) * N);
#pragma unroll
      for (int v = 0; v < N / 8; v++) {
        const uint4 q = __ldg(src + v);
        const uint32_t w[4] = { q.x, q.y, q.z, q.w };
#pragma unroll
        for (int i = 0; i < 4; i++) {
          x[v * 8 + 2 * i] = (int)(int16_t)(w[i] & 0xFFFFu);
          x[v * 8 + 2 * i + 1] = (int)(int16_t)(w[i] >> 16);
        }
      }
    }
    if (!INV) {
      const int s1 = lg - 1 + (bitdepth - 8), s2 = lg + 6;
      const int r1 = 1 << (s1 - 1), r2 = 1 << (s2 - 1);
      // pass 1: lane j makes column j of b:  b[k][j] = (sum_n M[k][n] a[j][n] + r1) >> s1
#pragma unroll kUnroll
      for (int k = 0; k < N; k++) t[k * kPitch + j] = (int16_t)((dot(k, x) + r1) >> s1);
      __syncwarp();
#pragma unroll
      for (int n = 0; n < N; n++) x[n] = t[j * kPitch + n];                 // row j of b
      __syncwarp();
      // pass 2: column j of the result:  c[k][j] = (sum_n M[k][n] b[j][n] + r2) >> s2
#pragma unroll kUnroll
      for (int k = 0; k < N; k++) t[k * kPitch + j] = (int16_t)((dot(k, x) + r2) >> s2);
      __syncwarp();
#pragma unroll
      for (int n = 0; n < N; n++) y[n] = t[j * kPitch + n];                 // row j of the result
      __syncwarp();
    } else {
      const int s2 = 12 - (bitdepth - 8);
#pragma unroll
      for (int n = 0; n < N; n++) t[j * kPitch + n] = (int16_t)x[n];
      __syncwarp();
#pragma unroll
      for (int k = 0; k < N; k++) x[k] = t[k * kPitch + j];                 // column j of a
      __syncwarp();
      // pass 1: row j of b:  b[j][n] = clip16((sum_k M[k][n] a[k][j] + 64) >> 7)
#pragma unroll kUnroll
      for (int n = 0; n < N; n++) t[j * kPitch + n] = (int16_t)clampi((dot(n, x) + 64) >> 7, -32768, 32767);
      __syncwarp();
#pragma unroll
      for (int k = 0; k < N; k++) x[k] = t[k * kPitch + j];                 // column j of b
      __syncwarp();
      // pass 2: row j of the result:  c[j][n] = clip16((sum_k M[k][n] b[k][j] + r) >> s2)
      // (the result goes through the tile as well when the loop is not unrolled: y[] must stay in registers)
#pragma unroll kUnroll
      for (int n = 0; n < N; n++) t[j * kPitch + n] = (int16_t)clampi((dot(n, x) + (1 << (s2 - 1))) >> s2, -32768, 32767);
      __syncwarp();
#pragma unroll
      for (int n = 0; n < N; n++) y[n] = t[j * kPitch + n];
      __syncwarp();
    }
    if (on) {
      uint4 *dst = reinterpret_cast<uint4 *>(out + ((size_t)blk * N + j) * N);
#pragma unroll
      for (int v = 0; v < N / 8; v++) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) w[i] = ((uint32_t)y[v * 8 + 2 * i] & 0xFFFFu) | ((uint32_t)y[v * 8 + 2 * i + 1] << 16);
        dst[v] = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  }
}

// ---- kvz_quant of a batch, one 16-coefficient group per THREAD (blocks of 4x4, 8x8, 16x16) ----
// The quantiser proper is per coefficient and the sign-bit hiding per coefficient group (quant-generic.c:69-162); what
// a block shares is the sum of its levels (hiding needs >= 2) and which group is the last one with a level.  So a
// thread takes a group -- its 4x4 corner of the block as four 8-byte loads, visited in the scan's order through a
// compile-time permutation, everything in registers -- and the (N/4)^2 lanes of a block exchange the two block-wide
// facts with shuffles.  Same arithmetic as quant_block; the backward walk with its strict '<' is kept as it is.

template <int N, int SCAN>
__global__ void __launch_bounds__(kThreads)
quant_groups_kernel(kvz_cuda_quant_cfg c, int type, const int16_t *__restrict__ in, int16_t *__restrict__ out, int count)
{
  constexpr int G = (N / 4) * (N / 4), kPerCta = kThreads / G, lg = N == 4 ? 2 : N == 8 ? 3 : 4;
  __shared__ uint8_t gxs[16], gys[16];
  if (threadIdx.x == 0) square_scan(SCAN, N / 4, gxs, gys);
  __syncthreads();
  const int qp_s = scaled_qp(type, c.qp, (c.bitdepth - 8) * 6);
  const long long qc = c_quant_scale[qp_s % 6];
  const int q_bits = 14 + qp_s / 6 + (15 - c.bitdepth - lg);
  const int add = (c.slice_is_intra ? 171 : 85) << (q_bits - 9);
  const int g = threadIdx.x % G;
  const int gx = gxs[g], gy = gys[g];
  for (int base = blockIdx.x * kPerCta; base < count; base += gridDim.x * kPerCta) {
    const int blk = base + threadIdx.x / G;
    const bool on = blk < count;
    const size_t corner = (size_t)(on ? blk : base) * N * N + (size_t)gy * 4 * N + gx * 4;
    int a[16];
#pragma unroll
    for (int y = 0; y < 4; y++) {
      const uint2 v = __ldg(reinterpret_cast<const uint2 *>(in + corner + (size_t)y * N));
      a[4 * y] = (int)(int16_t)(v.x & 0xFFFFu); a[4 * y + 1] = (int)(int16_t)(v.x >> 16);
      a[4 * y + 2] = (int)(int16_t)(v.y & 0xFFFFu); a[4 * y + 3] = (int)(int16_t)(v.y >> 16);
    }
    int q[16], dl[16];
    uint32_t neg = 0, my_sum = 0;
    bool any = false;
#pragma unroll
    for (int n = 0; n < 16; n++) {
      const int cf = a[cg_raster(SCAN, n)];
      const long long prod = (long long)abs(cf) * qc;
      int level = (int)((prod + add) >> q_bits);
      my_sum += (uint32_t)level;
      dl[n] = (int)((prod - ((long long)level << q_bits)) >> (q_bits - 8));
      if (cf < 0) { level = -level; neg |= 1u << n; }
      q[n] = clampi(level, -32768, 32767);
      any |= q[n] != 0;
    }
    uint32_t acc = my_sum;
    int last_cg = any ? g : -1;
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
      acc += __shfl_xor_sync(0xffffffffu, acc, o);
      last_cg = max(last_cg, __shfl_xor_sync(0xffffffffu, last_cg, o));
    }
    if (c.signhide && acc >= 2) {
      int first = 16, last = -1, qfirst = 0;
#pragma unroll
      for (int n = 0; n < 16; n++) if (q[n] != 0) last = n;
#pragma unroll
      for (int n = 15; n >= 0; n--) if (q[n] != 0) { first = n; qfirst = q[n]; }
      if (last - first >= 4) {
        int sum = 0;
#pragma unroll
        for (int n = 0; n < 16; n++) if (n >= first && n <= last) sum += q[n];
        const int signbit = qfirst > 0 ? 0 : 1;
        if (signbit != (sum & 1)) {
          const int top = g == last_cg ? last : 15;
          int best = 0x7FFFFFFF, bestn = -1, change = 0;
#pragma unroll
          for (int n = 15; n >= 0; n--) {
            if (n > top) continue;
            int cost = 0x7FFFFFFF, cur_change = 0;
            if (q[n] != 0) {
              if (dl[n] > 0) { cost = -dl[n]; cur_change = 1; }
              else if (n == first && abs(q[n]) == 1) { cost = 0x7FFFFFFF; }
              else { cost = dl[n]; cur_change = -1; }
            } else if (n < first && (int)((neg >> n) & 1u) != signbit) {
              cost = 0x7FFFFFFF;
            } else {
              cost = -dl[n]; cur_change = 1;
            }
            if (cost < best) { best = cost; change = cur_change; bestn = n; }
          }
          if (bestn >= 0) {
#pragma unroll
            for (int n = 0; n < 16; n++)
              if (n == bestn) {
                int ch = change;
                if (q[n] == 32767 || q[n] == -32768) ch = -1;
                q[n] = (int)(int16_t)(((neg >> n) & 1u) ? q[n] - ch : q[n] + ch);
              }
          }
        }
      }
    }
    if (on) {
#pragma unroll
      for (int n = 0; n < 16; n++) a[cg_raster(SCAN, n)] = q[n];
#pragma unroll
      for (int y = 0; y < 4; y++) {
        const uint32_t lo = ((uint32_t)a[4 * y] & 0xFFFFu) | ((uint32_t)a[4 * y + 1] << 16);
        const uint32_t hi = ((uint32_t)a[4 * y + 2] & 0xFFFFu) | ((uint32_t)a[4 * y + 3] << 16);
        *reinterpret_cast<uint2 *>(out + corner + (size_t)y * N) = make_uint2(lo, hi);
      }
    }
  }
}

// kvz_dequant with the flat scaling list (quant-generic.c:279-321) is per coefficient: 8 levels per thread
__global__ void __launch_bounds__(256)
dequant_flat_kernel(kvz_cuda_quant_cfg c, int lg, int type, const int16_t *__restrict__ in, int16_t *__restrict__ out, size_t n8)
{
  const int qp_s = scaled_qp(type, c.qp, (c.bitdepth - 8) * 6);
  const int shift = 20 - 14 - (15 - c.bitdepth - lg);
  const int scale = c_inv_quant_scale[qp_s % 6] << (qp_s / 6);
  const int add = 1 << (shift - 1);
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n8; i += (size_t)gridDim.x * 256) {
    const uint4 v = __ldg(reinterpret_cast<const uint4 *>(in) + i);
    uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int lo = clampi(((int)(int16_t)(w[k] & 0xFFFFu) * scale + add) >> shift, -32768, 32767);
      const int hi = clampi(((int)(int16_t)(w[k] >> 16) * scale + add) >> shift, -32768, 32767);
      w[k] = ((uint32_t)lo & 0xFFFFu) | ((uint32_t)hi << 16);
    }
    reinterpret_cast<uint4 *>(out)[i] = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

template <typename K>
static int opt_in_smem(K kern)
{
  KVZ_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BlockSmem)));
  return 1;
}

int launch_transform_batch(int n, int dst4, int inverse, int bitdepth, const int16_t *in, int16_t *out, int count,
                           cudaStream_t stream)
{
  if (!valid_n(n) || (dst4 && n != 4)) return fail_msg("transform: n must be 4, 8, 16 or 32 (DST only for 4)");
  if (count <= 0) return 1;
  if (n <= 8 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {     // a block per thread
    int grid = (count + kThreads - 1) / kThreads;
    if (grid > sm_count() * 16) grid = sm_count() * 16;
    if (n == 4) transform_small_kernel<4><<<grid, kThreads, 0, stream>>>(dst4, inverse, bitdepth, in, out, count);
    else transform_small_kernel<8><<<grid, kThreads, 0, stream>>>(dst4, inverse, bitdepth, in, out, count);
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
    return 1;
  }
  if (n >= 16 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {    // a block row per lane
    const int per_cta = kWarps * (32 / n);
    int grid = (count + per_cta - 1) / per_cta;
    if (grid > sm_count() * 16) grid = sm_count() * 16;
    if (n == 16 && !inverse) transform_rows_kernel<16, false><<<grid, kThreads, 0, stream>>>(bitdepth, in, out, count);
    else if (n == 16) transform_rows_kernel<16, true><<<grid, kThreads, 0, stream>>>(bitdepth, in, out, count);
    else if (!inverse) transform_rows_kernel<32, false><<<grid, kThreads, 0, stream>>>(bitdepth, in, out, count);
    else transform_rows_kernel<32, true><<<grid, kThreads, 0, stream>>>(bitdepth, in, out, count);
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
    return 1;
  }
  if (!opt_in_smem(transform_batch_kernel)) return 0;
  const int grid = batch_grid(n, count);
  transform_batch_kernel<<<grid, kThreads, sizeof(BlockSmem), stream>>>(n, dst4, inverse, bitdepth, in, out, count);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_quant_batch(const kvz_cuda_quant_cfg &c, int n, int type, int scan_idx, int dequant, const int16_t *in,
                       int16_t *out, int count, cudaStream_t stream)
{
  if (!valid_n(n)) return fail_msg("quant: n must be 4, 8, 16 or 32");
  if (count <= 0) return 1;
  const bool aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
  if (aligned && dequant) {
    const size_t n8 = (size_t)count * n * n / 8;
    size_t grid = (n8 + 255) / 256;
    if (grid > (size_t)sm_count() * 16) grid = (size_t)sm_count() * 16;
    dequant_flat_kernel<<<(int)grid, 256, 0, stream>>>(c, ilog2_host(n), type, in, out, n8);
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
    return 1;
  }
  if (aligned && n <= 16 && scan_idx >= 0 && scan_idx <= 2) {        // a coefficient group per thread
    const int per_cta = kThreads / ((n / 4) * (n / 4));
    int grid = (count + per_cta - 1) / per_cta;
    if (grid > sm_count() * 16) grid = sm_count() * 16;
#define KVZ_QUANT_GROUPS(NN, SS) quant_groups_kernel<NN, SS><<<grid, kThreads, 0, stream>>>(c, type, in, out, count)
    if (n == 4) { if (scan_idx == 0) KVZ_QUANT_GROUPS(4, 0); else if (scan_idx == 1) KVZ_QUANT_GROUPS(4, 1); else KVZ_QUANT_GROUPS(4, 2); }
    else if (n == 8) { if (scan_idx == 0) KVZ_QUANT_GROUPS(8, 0); else if (scan_idx == 1) KVZ_QUANT_GROUPS(8, 1); else KVZ_QUANT_GROUPS(8, 2); }
    else { if (scan_idx == 0) KVZ_QUANT_GROUPS(16, 0); else if (scan_idx == 1) KVZ_QUANT_GROUPS(16, 1); else KVZ_QUANT_GROUPS(16, 2); }
#undef KVZ_QUANT_GROUPS
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
    return 1;
  }
  if (!opt_in_smem(quant_batch_kernel)) return 0;
  const int grid = batch_grid(n, count);
  quant_batch_kernel<<<grid, kThreads, sizeof(BlockSmem), stream>>>(c, n, type, scan_idx, dequant, in, out, count);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_coeff_abs_sum(const int16_t *coeffs, int per_block, int count, uint32_t *out, cudaStream_t stream)
{
  if (count <= 0) return 1;
  coeff_abs_sum_kernel<<<(count + 7) / 8, 256, 0, stream>>>(coeffs, per_block, count, out);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_quantize_residual(const kvz_cuda_quant_cfg &c, const void *ref, const void *pred, void *rec, int16_t *coeff,
                             const kvz_cuda_residual_job *jobs, int n_jobs, uint8_t *has_coeffs, cudaStream_t stream)
{
  if (n_jobs <= 0) return 1;
  if (!opt_in_smem(quantize_residual_kernel<0, uint8_t>) || !opt_in_smem(quantize_residual_kernel<0, uint16_t>)) return 0;
  // groups of 32 jobs that are all 4x4: a block per thread; everything else: the group kernel, which skips those groups
  int grid4 = (n_jobs + kThreads - 1) / kThreads;
  if (grid4 > sm_count() * 16) grid4 = sm_count() * 16;
  if (c.bitdepth > 8) {
    quantize_residual4_kernel<uint16_t><<<grid4, kThreads, 0, stream>>>(c, (const uint16_t *)ref, (const uint16_t *)pred,
                                                                       (uint16_t *)rec, coeff, jobs, n_jobs, has_coeffs);
    quantize_residual_kernel<0, uint16_t><<<residual_grid(n_jobs), kThreads, sizeof(BlockSmem), stream>>>(
        c, (const uint16_t *)ref, (const uint16_t *)pred, (uint16_t *)rec, coeff, jobs, n_jobs, has_coeffs, nullptr, 0, 1);
  } else {
    quantize_residual4_kernel<uint8_t><<<grid4, kThreads, 0, stream>>>(c, (const uint8_t *)ref, (const uint8_t *)pred,
                                                                      (uint8_t *)rec, coeff, jobs, n_jobs, has_coeffs);
    quantize_residual_kernel<0, uint8_t><<<residual_grid(n_jobs), kThreads, sizeof(BlockSmem), stream>>>(
        c, (const uint8_t *)ref, (const uint8_t *)pred, (uint8_t *)rec, coeff, jobs, n_jobs, has_coeffs, nullptr, 0, 1);
  }
  count_launch();
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_rdoq_batch(const kvz_cuda_quant_cfg &cfg, const kvz_cuda_rdoq_ctx *d_ctx, const int16_t *d_coef,
                      int16_t *d_q_coef, const kvz_cuda_rdoq_job *d_jobs, int n_jobs, int max_width, int skip4,
                      cudaStream_t stream);

// kvz_quantize_residual with cfg.rdoq_enable: transform | RDOQ | reconstruction, three launches on one stream
int launch_quantize_residual_rdoq(const kvz_cuda_quant_cfg &c, const void *ref, const void *pred, void *rec, int16_t *coeff,
                                  const kvz_cuda_residual_job *jobs, int n_jobs, uint8_t *has_coeffs,
                                  const kvz_cuda_rdoq_ctx *d_ctx, const kvz_cuda_rdoq_job *rdoq_jobs, int max_width,
                                  int16_t *work, int rdoq_skip, cudaStream_t stream)
{
  if (n_jobs <= 0) return 1;
  if (!opt_in_smem(quantize_residual_kernel<1, uint8_t>) || !opt_in_smem(quantize_residual_kernel<2, uint8_t>) ||
      !opt_in_smem(quantize_residual_kernel<1, uint16_t>) || !opt_in_smem(quantize_residual_kernel<2, uint16_t>)) return 0;
  const bool wide = c.bitdepth > 8;
  if (wide)
    quantize_residual_kernel<1, uint16_t><<<residual_grid(n_jobs), kThreads, sizeof(BlockSmem), stream>>>(
        c, (const uint16_t *)ref, (const uint16_t *)pred, (uint16_t *)rec, coeff, jobs, n_jobs, has_coeffs, work, rdoq_skip, 0);
  else
    quantize_residual_kernel<1, uint8_t><<<residual_grid(n_jobs), kThreads, sizeof(BlockSmem), stream>>>(
        c, (const uint8_t *)ref, (const uint8_t *)pred, (uint8_t *)rec, coeff, jobs, n_jobs, has_coeffs, work, rdoq_skip, 0);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  if (!launch_rdoq_batch(c, d_ctx, work, coeff, rdoq_jobs, n_jobs, max_width, rdoq_skip, stream)) return 0;
  if (wide)
    quantize_residual_kernel<2, uint16_t><<<residual_grid(n_jobs), kThreads, sizeof(BlockSmem), stream>>>(
        c, (const uint16_t *)ref, (const uint16_t *)pred, (uint16_t *)rec, coeff, jobs, n_jobs, has_coeffs, work, rdoq_skip, 0);
  else
    quantize_residual_kernel<2, uint8_t><<<residual_grid(n_jobs), kThreads, sizeof(BlockSmem), stream>>>(
        c, (const uint8_t *)ref, (const uint8_t *)pred, (uint8_t *)rec, coeff, jobs, n_jobs, has_coeffs, work, rdoq_skip, 0);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}


// ---- the same without any allocation: the scratch is a file-scope device array --------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses that forbid it).
// No counterpart in the reference.
//
// g_noalloc_work is a host-side stand-in in host code, NOT a device address: its address on the current device
// is fetched once per device with cudaGetSymbolAddress and that pointer is what the kernels receive.  Every GPU
// that loads the library holds the array (kNoAllocMaxCoeffs * 2 bytes = 2 MiB) until exit.  All calls share it
// and it keeps the last call's contents; nothing in it has to start at zero (stage 1 writes every coefficient
// the RDOQ stage reads, see residual_job), so nothing is cleared.  Calls that overlap on the GPU race on it.
constexpr size_t kNoAllocMaxCoeffs = (size_t)KVZ_CUDA_NOALLOC_MAX_COEFFS;
static __device__ __align__(256) int16_t g_noalloc_work[kNoAllocMaxCoeffs];

constexpr int kNoAllocMaxDevices = 64;
static int16_t *g_noalloc_ptr[kNoAllocMaxDevices];     // device address of g_noalloc_work, per device
static std::mutex g_noalloc_lock;

int launch_quantize_residual_rdoq_noalloc(const kvz_cuda_quant_cfg &c, const void *ref, const void *pred, void *rec,
                                          int16_t *coeff, size_t coeff_elems, const kvz_cuda_residual_job *jobs, int n_jobs,
                                          uint8_t *has_coeffs, const kvz_cuda_rdoq_ctx *d_ctx,
                                          const kvz_cuda_rdoq_job *rdoq_jobs, int max_width, int rdoq_skip,
                                          cudaStream_t stream)
{
  if (coeff_elems > kNoAllocMaxCoeffs) {
    snprintf(g_last_error, sizeof(g_last_error),
             "kvz_cuda_quantize_residuals_rdoq_noalloc: %zu coefficients exceed the built-in scratch of %zu "
             "(KVZ_CUDA_NOALLOC_MAX_COEFFS); split the batch or use kvz_cuda_quantize_residuals_rdoq with your own d_work",
             coeff_elems, kNoAllocMaxCoeffs);
    return 0;
  }
  int dev = -1;
  KVZ_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail_msg("kvz_cuda_quantize_residuals_rdoq_noalloc: device index out of range");
  int16_t *work = nullptr;
  {
    std::lock_guard<std::mutex> guard(g_noalloc_lock);
    if (!g_noalloc_ptr[dev]) {                  // set-up, once per device, with that device current
      void *p = nullptr;
      KVZ_CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_work));
      g_noalloc_ptr[dev] = static_cast<int16_t *>(p);
    }
    work = g_noalloc_ptr[dev];
  }
  return launch_quantize_residual_rdoq(c, ref, pred, rec, coeff, jobs, n_jobs, has_coeffs, d_ctx, rdoq_jobs, max_width, work,
                                       rdoq_skip, stream);
}

}  // namespace kvz
