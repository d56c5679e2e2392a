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
#include "common.cuh"
#include "transform.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kThreads = 128;
constexpr int P = 34;                 // int16 row pitch of a block in shared memory (conflict-free columns)

__device__ __constant__ int32_t c_quant_scale[6] = { 26214, 23302, 20560, 18396, 16384, 14564 };   // scalinglist.c:66
__device__ __constant__ int32_t c_inv_quant_scale[6] = { 40, 45, 51, 57, 64, 72 };                 // scalinglist.c:67

struct BlockSmem {
  int16_t a[32 * P];
  int16_t b[32 * P];
  int32_t delta[32 * 32];
  TransformTables tt;
  ScanTable scan;
  uint32_t acc;
  int last_cg;
};

__device__ void forward2d(BlockSmem &sm, int N, int bitdepth, int tid)
{
  // a -> b -> a      (dct-generic.c:571-572)
  const int lg = ilog2_pow2(N);
  fwd_pass_smem<P>(sm.tt, N, sm.a, sm.b, lg - 1 + (bitdepth - 8), tid, kThreads);
  __syncthreads();
  fwd_pass_smem<P>(sm.tt, N, sm.b, sm.a, lg + 6, tid, kThreads);
  __syncthreads();
}

__device__ void inverse2d(BlockSmem &sm, int N, int bitdepth, int tid)
{
  // a -> b -> a      (dct-generic.c:582-583)
  inv_pass_smem<P>(sm.tt, N, sm.a, sm.b, 7, tid, kThreads);
  __syncthreads();
  inv_pass_smem<P>(sm.tt, N, sm.b, sm.a, 12 - (bitdepth - 8), tid, kThreads);
  __syncthreads();
}

// kvz_quant (quant-generic.c:37-163): coefficients in sm.a, levels to sm.b.
// The scan table must have been built when cfg.signhide is set.
__device__ void quant_block(BlockSmem &sm, const kvz_cuda_quant_cfg &c, int N, int type, int tid)
{
  const int lg = ilog2_pow2(N);
  const int qp_s = scaled_qp(type, c.qp, (c.bitdepth - 8) * 6);
  const long long qc = c_quant_scale[qp_s % 6];
  const int q_bits = 14 + qp_s / 6 + (15 - c.bitdepth - lg);
  const int add = (c.slice_is_intra ? 171 : 85) << (q_bits - 9);
  if (tid == 0) { sm.acc = 0; sm.last_cg = -1; }
  __syncthreads();
  uint32_t my_sum = 0;
  for (int i = tid; i < N * N; i += kThreads) {
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
  __syncthreads();
  if (!c.signhide || sm.acc < 2) return;          // uniform

  // sign-bit hiding: one thread per 16-coefficient group (quant-generic.c:69-162)
  const int n_cg = (N * N) >> 4;
  auto q_at = [&](int pos) -> int16_t & { return sm.b[(pos / N) * P + (pos % N)]; };
  auto c_at = [&](int pos) -> int { return sm.a[(pos / N) * P + (pos % N)]; };
  for (int g = tid; g < n_cg; g += kThreads) {
    bool any = false;
    for (int n = 0; n < 16; n++) any |= q_at(sm.scan.pos[g * 16 + n]) != 0;
    if (any) atomicMax(&sm.last_cg, g);
  }
  __syncthreads();
  for (int g = tid; g < n_cg; g += kThreads) {
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
  __syncthreads();
}

// kvz_dequant (quant-generic.c:279-321, flat list): levels in sm.b -> coefficients in sm.a
__device__ void dequant_block(BlockSmem &sm, const kvz_cuda_quant_cfg &c, int N, int type, int tid)
{
  const int lg = ilog2_pow2(N);
  const int qp_s = scaled_qp(type, c.qp, (c.bitdepth - 8) * 6);
  const int shift = 20 - 14 - (15 - c.bitdepth - lg);
  const int scale = c_inv_quant_scale[qp_s % 6] << (qp_s / 6);
  const int add = 1 << (shift - 1);
  for (int i = tid; i < N * N; i += kThreads) {
    const int y = i / N, x = i - y * N;
    const int v = ((int)sm.b[y * P + x] * scale + add) >> shift;
    sm.a[y * P + x] = (int16_t)clampi(v, -32768, 32767);
  }
  __syncthreads();
}

__device__ __forceinline__ void load_block(int16_t *dst, const int16_t *src, int N, int tid)
{
  for (int i = tid; i < N * N; i += kThreads) dst[(i / N) * P + (i % N)] = src[i];
}

__device__ __forceinline__ void store_block(int16_t *dst, const int16_t *src, int N, int tid)
{
  for (int i = tid; i < N * N; i += kThreads) dst[i] = src[(i / N) * P + (i % N)];
}

// ---- batch kernels ----

__global__ void __launch_bounds__(kThreads)
transform_batch_kernel(int N, int dst4, int inverse, int bitdepth, const int16_t *__restrict__ in,
                       int16_t *__restrict__ out, int count)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BlockSmem &sm = *reinterpret_cast<BlockSmem *>(smem_raw);
  const int tid = threadIdx.x;
  load_transform_matrix(sm.tt, N, dst4 != 0, tid, kThreads);
  for (int blk = blockIdx.x; blk < count; blk += gridDim.x) {
    __syncthreads();
    load_block(sm.a, in + (size_t)blk * N * N, N, tid);
    __syncthreads();
    if (inverse) inverse2d(sm, N, bitdepth, tid);
    else forward2d(sm, N, bitdepth, tid);
    store_block(out + (size_t)blk * N * N, sm.a, N, tid);
  }
}

__global__ void __launch_bounds__(kThreads)
quant_batch_kernel(kvz_cuda_quant_cfg c, int N, int type, int scan_idx, int dequant, const int16_t *__restrict__ in,
                   int16_t *__restrict__ out, int count)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BlockSmem &sm = *reinterpret_cast<BlockSmem *>(smem_raw);
  const int tid = threadIdx.x;
  if (!dequant && c.signhide) build_scan_table(sm.scan, scan_idx, ilog2_pow2(N), tid, kThreads);
  for (int blk = blockIdx.x; blk < count; blk += gridDim.x) {
    __syncthreads();
    if (dequant) {
      load_block(sm.b, in + (size_t)blk * N * N, N, tid);
      __syncthreads();
      dequant_block(sm, c, N, type, tid);
      store_block(out + (size_t)blk * N * N, sm.a, N, tid);
    } else {
      load_block(sm.a, in + (size_t)blk * N * N, N, tid);
      __syncthreads();
      quant_block(sm, c, N, type, tid);
      __syncthreads();
      store_block(out + (size_t)blk * N * N, sm.b, N, tid);
    }
  }
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

// kvz_quantize_residual (quant-generic.c:180-273), one CTA per transform block.
// STAGE 0: the whole function with the scalar quantiser (cfg.rdoq_enable == 0).
// With RDOQ the function is three launches around rdoq_kernel (rdoq.cu):
// STAGE 1: residual -> transform, coefficients to `work` (4x4 blocks under cfg.rdoq_skip are quantised here,
//          quant-generic.c:213-215); STAGE 2: levels from coeff_out -> dequant -> inverse -> reconstruction.
template <int STAGE>
__global__ void __launch_bounds__(kThreads)
quantize_residual_kernel(kvz_cuda_quant_cfg c, const uint8_t *__restrict__ ref, const uint8_t *__restrict__ pred,
                         uint8_t *__restrict__ rec, int16_t *__restrict__ coeff_out,
                         const kvz_cuda_residual_job *__restrict__ jobs, int n_jobs, uint8_t *__restrict__ has_coeffs,
                         int16_t *__restrict__ work, int rdoq_skip)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BlockSmem &sm = *reinterpret_cast<BlockSmem *>(smem_raw);
  __shared__ int s_any;
  const int tid = threadIdx.x;
  const kvz_cuda_residual_job jb = jobs[blockIdx.x];
  const int N = jb.width, lg = ilog2_pow2(N);
  const bool luma = jb.color == 0;
  c.qp = jb.qp;
  c.cu_is_intra = jb.cu_is_intra;
  const bool dst4 = N == 4 && luma && jb.cu_is_intra;          // strategies-dct.c:59-113
  if (!jb.use_trskip) load_transform_matrix(sm.tt, N, dst4, tid, kThreads);
  const bool scalar_quant = STAGE == 0 || (STAGE == 1 && N == 4 && rdoq_skip);
  if (c.signhide && scalar_quant) build_scan_table(sm.scan, jb.scan_idx, lg, tid, kThreads);
  if (tid == 0) s_any = 0;
  const uint8_t *rp = ref + jb.ref_offset, *pp = pred + jb.pred_offset;
  const int ts_shift = 15 - c.bitdepth - lg;
  int16_t *co = coeff_out + jb.coeff_offset;
  if (STAGE != 2) {
    for (int i = tid; i < N * N; i += kThreads) {
      const int y = i / N, x = i - y * N;
      sm.a[y * P + x] = (int16_t)((int)rp[(size_t)y * jb.in_stride + x] - (int)pp[(size_t)y * jb.in_stride + x]);
    }
    __syncthreads();
    if (jb.use_trskip) {                                           // transform.c:151-161
      for (int i = tid; i < N * N; i += kThreads) {
        int16_t &v = sm.a[(i / N) * P + (i % N)];
        v = (int16_t)((int)v << ts_shift);
      }
      __syncthreads();
    } else {
      forward2d(sm, N, c.bitdepth, tid);
    }
    if (!scalar_quant) {                                           // STAGE 1: the coefficients go to rdoq_kernel
      int16_t *wo = work + jb.coeff_offset;
      for (int i = tid; i < N * N; i += kThreads) wo[i] = sm.a[(i / N) * P + (i % N)];
      return;
    }
    quant_block(sm, c, N, luma ? 0 : 2, tid);
    __syncthreads();
  } else {
    for (int i = tid; i < N * N; i += kThreads) sm.b[(i / N) * P + (i % N)] = co[i];
    __syncthreads();
  }
  int any = 0;
  for (int i = tid; i < N * N; i += kThreads) {
    const int16_t q = sm.b[(i / N) * P + (i % N)];
    if (STAGE != 2) co[i] = q;
    any |= q != 0;
  }
  if (STAGE == 1) return;
  if (any) s_any = 1;
  __syncthreads();
  uint8_t *op = rec + jb.rec_offset;
  if (s_any) {
    dequant_block(sm, c, N, luma ? 0 : (jb.color == 1 ? 2 : 3), tid);
    if (jb.use_trskip) {                                         // transform.c:169-181
      const int off = 1 << (ts_shift - 1);
      for (int i = tid; i < N * N; i += kThreads) {
        int16_t &v = sm.a[(i / N) * P + (i % N)];
        v = (int16_t)(((int)v + off) >> ts_shift);
      }
      __syncthreads();
    } else {
      inverse2d(sm, N, c.bitdepth, tid);
    }
    const int pix_max = (1 << c.bitdepth) - 1;
    for (int i = tid; i < N * N; i += kThreads) {
      const int y = i / N, x = i - y * N;
      const int16_t val = (int16_t)((int)sm.a[y * P + x] + (int)pp[(size_t)y * jb.in_stride + x]);
      op[(size_t)y * jb.out_stride + x] = (uint8_t)clampi(val, 0, pix_max);
    }
  } else {
    for (int i = tid; i < N * N; i += kThreads) {
      const int y = i / N, x = i - y * N;
      op[(size_t)y * jb.out_stride + x] = pp[(size_t)y * jb.in_stride + x];
    }
  }
  if (tid == 0) has_coeffs[blockIdx.x] = (uint8_t)s_any;
}

bool valid_n(int n) { return n == 4 || n == 8 || n == 16 || n == 32; }

}  // namespace

static int g_sms = 0;
static int sm_count()
{
  if (!g_sms) { int d = 0; cudaGetDevice(&d); cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, d); }
  return g_sms;
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
  if (!opt_in_smem(transform_batch_kernel)) return 0;
  const int grid = count < sm_count() * 4 ? count : sm_count() * 4;      // persistent CTAs, 4 per SM
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
  if (!opt_in_smem(quant_batch_kernel)) return 0;
  const int grid = count < sm_count() * 4 ? count : sm_count() * 4;
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
  if (!opt_in_smem(quantize_residual_kernel<0>)) return 0;
  quantize_residual_kernel<0><<<n_jobs, kThreads, sizeof(BlockSmem), stream>>>(
      c, (const uint8_t *)ref, (const uint8_t *)pred, (uint8_t *)rec, coeff, jobs, n_jobs, has_coeffs, nullptr, 0);
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
  if (!opt_in_smem(quantize_residual_kernel<1>) || !opt_in_smem(quantize_residual_kernel<2>)) return 0;
  quantize_residual_kernel<1><<<n_jobs, kThreads, sizeof(BlockSmem), stream>>>(
      c, (const uint8_t *)ref, (const uint8_t *)pred, (uint8_t *)rec, coeff, jobs, n_jobs, has_coeffs, work, rdoq_skip);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  if (!launch_rdoq_batch(c, d_ctx, work, coeff, rdoq_jobs, n_jobs, max_width, rdoq_skip, stream)) return 0;
  quantize_residual_kernel<2><<<n_jobs, kThreads, sizeof(BlockSmem), stream>>>(
      c, (const uint8_t *)ref, (const uint8_t *)pred, (uint8_t *)rec, coeff, jobs, n_jobs, has_coeffs, work, rdoq_skip);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
