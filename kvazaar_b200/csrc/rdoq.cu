// kvazaar_b200/csrc/rdoq.cu
//
// Rate-distortion optimised quantisation for a batch of transform blocks:
//   kvz_rdoq (rdo.c:548-885), kvz_get_coded_level (rdo.c:297-339), kvz_get_ic_rate (rdo.c:233-278),
//   calc_last_bits / get_rate_last (rdo.c:350-395), kvz_rdoq_sign_hiding (rdo.c:405-540),
//   significance context selection (context.c:303-385).
//
// The algorithm is a serial chain per block: the Rice / greater-than-1 state runs from coefficient to
// coefficient, and the running cost is a sum of doubles whose order decides ties.  Parallelism therefore
// comes from the batch (one warp per transform block, thousands of blocks per launch) and from the parts
// of a block that ARE independent:
//   * all lanes: scan table, |coef| * scale, the distortion of dropping each coefficient, the start of
//     the scan (a warp max), the final sign / zero pass, and sign hiding (one lane per coefficient group:
//     a group's parity fix touches only its own 16 levels);
//   * lane 0: level choice per position, group zeroing, last-position search -- in the reference's exact
//     double-precision operation order (compiled with -fmad=false, so no contraction).
// Everything a block needs lives in shared memory (51 KB for 32x32, 3.5 KB for 8x8): one read of the
// coefficients from HBM, one write of the levels.
#include "common.cuh"
#include "transform.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kOneBit = 32768;                  // 1 << CTX_FRAC_BITS (rdo.h:65-66)
constexpr double kBigCost = 1.7e+308;           // MAX_DOUBLE (global.h:269)

// bits (x 2^15) of a bin in a context state, index = state ^ bin: the HM 12.0 estimates of rdo.c:53-63
__device__ __constant__ uint32_t c_entropy_bits[128] = {
  0x08000, 0x08000, 0x076da, 0x089a0, 0x06e92, 0x09340, 0x0670a, 0x09cdf, 0x06029, 0x0a67f, 0x059dd,
  0x0b01f, 0x05413, 0x0b9bf, 0x04ebf, 0x0c35f, 0x049d3, 0x0ccff, 0x04546, 0x0d69e, 0x0410d, 0x0e03e,
  0x03d22, 0x0e9de, 0x0397d, 0x0f37e, 0x03619, 0x0fd1e, 0x032ee, 0x106be, 0x02ffa, 0x1105d, 0x02d37,
  0x119fd, 0x02aa2, 0x1239d, 0x02836, 0x12d3d, 0x025f2, 0x136dd, 0x023d1, 0x1407c, 0x021d2, 0x14a1c,
  0x01ff2, 0x153bc, 0x01e2f, 0x15d5c, 0x01c87, 0x166fc, 0x01af7, 0x1709b, 0x0197f, 0x17a3b, 0x0181d,
  0x183db, 0x016d0, 0x18d7b, 0x01595, 0x1971b, 0x0146c, 0x1a0bb, 0x01354, 0x1aa5a, 0x0124c, 0x1b3fa,
  0x01153, 0x1bd9a, 0x01067, 0x1c73a, 0x00f89, 0x1d0da, 0x00eb7, 0x1da79, 0x00df0, 0x1e419, 0x00d34,
  0x1edb9, 0x00c82, 0x1f759, 0x00bda, 0x200f9, 0x00b3c, 0x20a99, 0x00aa5, 0x21438, 0x00a17, 0x21dd8,
  0x00990, 0x22778, 0x00911, 0x23118, 0x00898, 0x23ab8, 0x00826, 0x24458, 0x007ba, 0x24df7, 0x00753,
  0x25797, 0x006f2, 0x26137, 0x00696, 0x26ad7, 0x0063f, 0x27477, 0x005ed, 0x27e17, 0x0059f, 0x287b6,
  0x00554, 0x29156, 0x0050e, 0x29af6, 0x004cc, 0x2a497, 0x0048d, 0x2ae35, 0x00451, 0x2b7d6, 0x00418,
  0x2c176, 0x003e2, 0x2cb15, 0x003af, 0x2d4b5, 0x0037f, 0x2de55
};
__device__ __constant__ int32_t c_rq_quant_scale[6] = { 26214, 23302, 20560, 18396, 16384, 14564 };   // scalinglist.c:66
__device__ __constant__ int32_t c_rq_inv_quant_scale[6] = { 40, 45, 51, 57, 64, 72 };                 // scalinglist.c:67
__device__ __constant__ uint8_t c_last_group[32] = { 0, 1, 2, 3, 4, 4, 5, 5, 6, 6, 6, 6, 7, 7, 7, 7,   // encoderstate.h:345-349
                                                     8, 8, 8, 8, 8, 8, 8, 8, 9, 9, 9, 9, 9, 9, 9, 9 };
__device__ __constant__ uint8_t c_sig_map4x4[16] = { 0, 1, 4, 5, 2, 3, 4, 5, 6, 6, 8, 8, 7, 7, 8, 8 };   // context.c:357-363

__device__ __forceinline__ int bin_bits(uint8_t state, int bin) { return (int)c_entropy_bits[state ^ bin]; }

// fixed part of a block's shared memory; the per-coefficient arrays follow it
struct RdoqHead {
  ScanTable scan;
  double group_sig_cost[64];
  int32_t bits_x[12], bits_y[12];
  kvz_cuda_rdoq_ctx ctx;
  uint8_t group_sig[64];
  int32_t coded_end;
  uint32_t abs_sum;
};

struct Arrays {
  double *cost, *cost_sig, *cost_zero;     // by scan position
  int4 *hide;                              // by scan position: x inc, y dec, z sig_inc, w delta
  int32_t *scaled;                         // by scan position: min(|coef| * scale, cap)
  int16_t *coef, *lv;                      // by block position
};

__host__ __device__ inline size_t rdoq_smem_bytes(int nn)
{
  return ((sizeof(RdoqHead) + 15) & ~(size_t)15) + (size_t)nn * (3 * 8 + 16 + 4 + 2 + 2);
}

// the context models of the block's colour channel
struct Models {
  const uint8_t *sig, *one, *abs;
};

// greater-than-1 / greater-than-2 / Rice state inside a coefficient group
struct LevelState {
  int ctx_set, c1, c2, rice, c1_idx, c2_idx;
  __device__ __forceinline__ int base_level() const { return c1_idx < 8 ? (c2_idx < 1 ? 3 : 2) : 1; }
};

// rdo.c:233-278
__device__ int level_rate(const Models &m, int level, const LevelState &s)
{
  const int one_ctx = 4 * s.ctx_set + s.c1, abs_ctx = s.ctx_set + s.c2;
  const int base = s.base_level();
  int rate = kOneBit;
  if (level >= base) {
    int rem = level - base;
    const int k = s.rice;
    if (rem < (3 << k)) {
      rate += ((rem >> k) + 1 + k) * kOneBit;
    } else {
      rem -= 3 << k;
      // escape: the prefix grows by one bit per doubling -> position of the top bit of (rem >> k) + 1
      const int len = k + (31 - __clz((rem >> k) + 1));
      rate += (3 + len + 1 - k + len) * kOneBit;
    }
    if (s.c1_idx < 8) {
      rate += bin_bits(m.one[one_ctx], 1);
      if (s.c2_idx < 1) rate += bin_bits(m.abs[abs_ctx], 1);
    }
  } else if (level == 1) {
    rate += bin_bits(m.one[one_ctx], 0);
  } else if (level == 2) {
    rate += bin_bits(m.one[one_ctx], 1) + bin_bits(m.abs[abs_ctx], 0);
  }
  return rate;
}

// context.c:354-385
__device__ int sig_ctx_inc(int pattern, int scan_idx, int x, int y, int lg, bool chroma)
{
  if (x + y == 0) return 0;
  if (lg == 2) return c_sig_map4x4[4 * y + x];
  const int base = lg == 3 ? (scan_idx == 0 ? 9 : 15) : (chroma ? 12 : 21);
  const int xs = x & 3, ys = y & 3;
  int cnt;
  if (pattern == 0) cnt = xs + ys == 0 ? 2 : (xs + ys <= 2 ? 1 : 0);
  else if (pattern == 1) cnt = ys == 0 ? 2 : (ys == 1 ? 1 : 0);
  else if (pattern == 2) cnt = xs == 0 ? 2 : (xs == 1 ? 1 : 0);
  else cnt = 2;
  return ((!chroma && ((x >> 2) + (y >> 2)) > 0) ? 3 : 0) + base + cnt;
}

// context.c:303-339: bit 0 = the group to the right is significant, bit 1 = the group below
__device__ __forceinline__ int neighbour_groups(const uint8_t *group_sig, int gx, int gy, int gside)
{
  int r = 0, b = 0;
  if (gx < gside - 1) r = group_sig[gy * gside + gx + 1] != 0;
  if (gy < gside - 1) b = group_sig[(gy + 1) * gside + gx] != 0;
  return r | (b << 1);
}

// rdo.c:366-395
__device__ void last_prefix_bits(const uint8_t *models, int lg, bool chroma, int32_t *bits)
{
  const int cb = lg - 2;
  const int offset = chroma ? 0 : cb * 3 + ((cb + 1) >> 2);
  const int shift = chroma ? cb : (cb + 3) >> 2;
  const int groups = c_last_group[(1 << lg) - 1];
  int run = 0;
  for (int g = 0; g < groups; g++) {
    const uint8_t st = models[offset + (g >> shift)];
    bits[g] = run + bin_bits(st, 0);
    run += bin_bits(st, 1);
  }
  bits[groups] = run;
}

// rdo.c:350-364
__device__ __forceinline__ double last_pos_cost(double lambda, int a, int b, const int32_t *bits_x, const int32_t *bits_y)
{
  const int ga = c_last_group[a], gb = c_last_group[b];
  double bits = (double)(bits_x[ga] + bits_y[gb]);
  if (ga > 3) bits += (double)(kOneBit * ((ga - 2) >> 1));
  if (gb > 3) bits += (double)(kOneBit * ((gb - 2) >> 1));
  return lambda * bits;
}

// ---- lane 0: level choice, group zeroing and last-position search (rdo.c:663-900) ----
// Returns with A.lv holding unsigned levels for scan positions <= last_pos and H.coded_end set.
__device__ void serial_pass(RdoqHead &H, const Arrays &A, const kvz_cuda_rdoq_job &jb, bool signhide, int lg, int q_bits,
                            double err_scale, int last_pos)
{
  const int width = 1 << lg, gside = width >> 2;
  const bool chroma = jb.type != 0;
  const int scan_idx = jb.scan_idx;
  const double lambda = jb.lambda;
  const uint16_t *scan = H.scan.pos;
  const Models m = { chroma ? H.ctx.sig_chroma : H.ctx.sig_luma, chroma ? H.ctx.one_chroma : H.ctx.one_luma,
                     chroma ? H.ctx.abs_chroma : H.ctx.abs_luma };
  const uint8_t *group_models = H.ctx.sig_group + jb.type;       // &cu_sig_coeff_group_model[type] (rdo.c:615)
  const int last_group = last_pos >> 4;
  LevelState st = { (last_pos > 0 && !chroma) ? 2 : 0, 1, 0, 0, 0, 0 };
  double base_cost = 0, uncoded_cost = 0;

  for (int g = last_group; g >= 0; g--) {
    const int first_bp = scan[16 * g];
    const int gx = (first_bp & (width - 1)) >> 2, gy = (first_bp >> lg) >> 2;
    const int gpos = gy * gside + gx;
    const int pattern = neighbour_groups(H.group_sig, gx, gy, gside);
    double s_coded = 0, s_uncoded = 0, s_sig = 0, s_sig0 = 0;
    int nnz_above0 = 0;
    bool any = false;

    for (int k = 15; k >= 0; k--) {
      const int sp = 16 * g + k;
      if (sp > last_pos) continue;
      const int bp = scan[sp];
      const int scaled = A.scaled[sp];
      const int max_level = (int)((uint32_t)(scaled + (1 << (q_bits - 1))) >> q_bits);
      const double zero_cost = A.cost_zero[sp];
      uncoded_cost += zero_cost;
      const bool opening = sp == last_pos;
      int sctx = 0;
      if (!opening) sctx = sig_ctx_inc(pattern, scan_idx, bp & (width - 1), bp >> lg, lg, chroma);

      // kvz_get_coded_level (rdo.c:297-339)
      int level = 0;
      double cost, cost_sig = 0, sig1 = 0;
      bool decided = false;
      if (!opening && max_level < 3) {
        cost_sig = lambda * (double)bin_bits(m.sig[sctx], 0);
        cost = zero_cost + cost_sig;
        decided = max_level == 0;
      } else {
        cost = kBigCost;
      }
      if (!decided) {
        if (!opening) sig1 = lambda * (double)bin_bits(m.sig[sctx], 1);
        const int lowest = max_level > 1 ? max_level - 1 : 1;
        for (int lv = max_level; lv >= lowest; lv--) {
          const double err = (double)(scaled - lv * (1 << q_bits));
          double cand = err * err * err_scale + lambda * (double)level_rate(m, lv, st);
          cand += sig1;
          if (cand < cost) { level = lv; cost = cand; cost_sig = sig1; }
        }
      }
      A.cost[sp] = cost;
      A.cost_sig[sp] = cost_sig;

      if (signhide) {
        int4 h;
        h.z = opening ? 0 : bin_bits(m.sig[sctx], 1) - bin_bits(m.sig[sctx], 0);
        h.w = (scaled - level * (1 << q_bits)) >> (q_bits - 8);
        if (level > 0) {
          const int now = level_rate(m, level, st);
          h.x = level_rate(m, level + 1, st) - now;
          h.y = level_rate(m, level - 1, st) - now;
        } else {
          h.x = bin_bits(m.one[4 * st.ctx_set + st.c1], 0);
          h.y = 0;
        }
        A.hide[sp] = h;
      }
      A.lv[bp] = (int16_t)level;
      base_cost += cost;

      // rdo.c:731-761
      if (level >= st.base_level() && level > (3 << st.rice)) st.rice = min(st.rice + 1, 4);
      if (level >= 1) st.c1_idx++;
      if (level > 1) { st.c1 = 0; st.c2 += st.c2 < 2; st.c2_idx++; }
      else if (st.c1 < 3 && st.c1 > 0 && level) st.c1++;
      if (k == 0 && sp > 0) {
        st.c2 = 0; st.rice = 0; st.c1_idx = 0; st.c2_idx = 0;
        st.ctx_set = (sp == 16 || chroma) ? 0 : 2;
        if (st.c1 == 0) st.ctx_set++;
        st.c1 = 1;
      }

      s_sig += cost_sig;
      if (k == 0) s_sig0 = cost_sig;
      if (level) {
        any = true;
        s_coded += cost - cost_sig;
        s_uncoded += zero_cost;
        if (k != 0) nnz_above0++;
      }
    }
    if (any) H.group_sig[gpos] = 1;

    if (g == 0) { H.group_sig[gpos] = 1; continue; }
    const int gctx = neighbour_groups(H.group_sig, gx, gy, gside) != 0;
    if (!any) {
      H.group_sig_cost[g] = lambda * (double)bin_bits(group_models[gctx], 0);
      base_cost += H.group_sig_cost[g] - s_sig;
    } else if (g < last_group) {
      if (nnz_above0 == 0) { base_cost -= s_sig0; s_sig -= s_sig0; }
      double zeroed = base_cost;
      H.group_sig_cost[g] = lambda * (double)bin_bits(group_models[gctx], 1);
      base_cost += H.group_sig_cost[g];
      zeroed += lambda * (double)bin_bits(group_models[gctx], 0);
      zeroed += s_uncoded;
      zeroed -= s_coded;
      zeroed -= s_sig;
      if (zeroed < base_cost) {
        H.group_sig[gpos] = 0;
        base_cost = zeroed;
        H.group_sig_cost[g] = lambda * (double)bin_bits(group_models[gctx], 0);
        for (int k = 15; k >= 0; k--) {
          const int sp = 16 * g + k, bp = scan[sp];
          if (A.lv[bp]) { A.lv[bp] = 0; A.cost[sp] = A.cost_zero[sp]; A.cost_sig[sp] = 0; }
        }
      }
    }
  }

  // the cheapest place to end the coded part (rdo.c:838-900)
  double best_cost;
  {
    uint8_t cbf;
    if (!jb.cu_is_intra && !chroma) cbf = H.ctx.root_cbf;
    else cbf = chroma ? H.ctx.cbf_chroma[jb.tr_depth] : H.ctx.cbf_luma[!jb.tr_depth];
    best_cost = uncoded_cost + lambda * (double)bin_bits(cbf, 0);
    base_cost += lambda * (double)bin_bits(cbf, 1);
  }
  int coded_end = 0;
  bool stop = false;
  for (int g = last_group; g >= 0 && !stop; g--) {
    const int first_bp = scan[16 * g];
    const int gpos = ((first_bp >> lg) >> 2) * gside + ((first_bp & (width - 1)) >> 2);
    base_cost -= H.group_sig_cost[g];
    if (!H.group_sig[gpos]) continue;
    for (int k = 15; k >= 0; k--) {
      const int sp = 16 * g + k;
      if (sp > last_pos) continue;
      const int bp = scan[sp];
      const int level = A.lv[bp];
      if (!level) { base_cost -= A.cost_sig[sp]; continue; }
      const int y = bp >> lg, x = bp & (width - 1);
      const double lc = scan_idx == 2 ? last_pos_cost(lambda, y, x, H.bits_x, H.bits_y)
                                      : last_pos_cost(lambda, x, y, H.bits_x, H.bits_y);
      const double total = base_cost + lc - A.cost_sig[sp];
      if (total < best_cost) { coded_end = sp + 1; best_cost = total; }
      if (level > 1) { stop = true; break; }
      base_cost -= A.cost[sp];
      base_cost += A.cost_zero[sp];
    }
  }
  H.coded_end = coded_end;
}

// rdo.c:441-539 for one coefficient group g (levels already signed)
__device__ void hide_sign_in_group(const RdoqHead &H, const Arrays &A, int g, int top_group, long long rd_factor)
{
  const uint16_t *sc = H.scan.pos + 16 * g;
  int first = 16, last = -1;
  for (int n = 15; n >= 0; n--) if (A.lv[sc[n]]) { last = n; break; }
  for (int n = 0; n <= last; n++) if (A.lv[sc[n]]) { first = n; break; }
  if (last - first < 4) return;                                       // SBH_THRESHOLD
  const int sign = A.lv[sc[first]] <= 0;
  unsigned sum = 0;
  for (int n = first; n <= last; n++) sum += (unsigned)(int)A.lv[sc[n]];
  if (sign == (int)(sum & 1)) return;

  long long best_cost = 0x7FFFFFFFFFFFFFFFLL;
  int best_pos = 0, best_change = 0;
  for (int n = (g == top_group ? last : 15); n >= 0; n--) {
    const int pos = sc[n];
    const int4 h = A.hide[16 * g + n];
    const long long dist = rd_factor * h.w;
    const int mag = abs((int)A.lv[pos]);
    long long cost;
    int change;
    if (mag != 0) {
      long long up = h.x, down = h.y;
      if (mag == 1) down -= kOneBit + h.z;
      if (g == top_group && n == last && mag == 1) down -= 4 * kOneBit;
      up = -dist + up;
      down = dist + down;
      if (up < down) { change = 1; cost = up; }
      else {
        change = -1; cost = down;
        if (n == first && mag == 1) cost = 0x7FFFFFFFFFFFFFFFLL;
      }
    } else {
      const int bits = kOneBit + h.x + h.z;
      cost = -llabs(dist) + bits;
      change = 1;
      if (n < first && (A.coef[pos] >= 0 ? 0 : 1) != sign) cost = 0x7FFFFFFFFFFFFFFFLL;
    }
    if (cost < best_cost) { best_cost = cost; best_pos = pos; best_change = change; }
  }
  const int cur = A.lv[best_pos];
  if (cur == 32767 || cur == -32768) best_change = -1;
  A.lv[best_pos] = (int16_t)(A.coef[best_pos] >= 0 ? cur + best_change : cur - best_change);
}

__global__ void __launch_bounds__(32)
rdoq_kernel(kvz_cuda_quant_cfg cfg, const kvz_cuda_rdoq_ctx *__restrict__ ctxs, const int16_t *__restrict__ coef_g,
            int16_t *__restrict__ out_g, const kvz_cuda_rdoq_job *__restrict__ jobs, int n_jobs, int skip4)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const kvz_cuda_rdoq_job jb = jobs[blockIdx.x];
  const int width = jb.width, lg = ilog2_pow2(width), nn = width * width;
  if (skip4 && width == 4) return;          // cfg.rdoq_skip: 4x4 blocks keep the scalar quantiser's levels (quant-generic.c:213-215)
  RdoqHead &H = *reinterpret_cast<RdoqHead *>(smem_raw);
  Arrays A;
  {
    unsigned char *p = smem_raw + ((sizeof(RdoqHead) + 15) & ~(size_t)15);
    A.cost = reinterpret_cast<double *>(p);       p += (size_t)nn * 8;
    A.cost_sig = reinterpret_cast<double *>(p);   p += (size_t)nn * 8;
    A.cost_zero = reinterpret_cast<double *>(p);  p += (size_t)nn * 8;
    A.hide = reinterpret_cast<int4 *>(p);         p += (size_t)nn * 16;
    A.scaled = reinterpret_cast<int32_t *>(p);    p += (size_t)nn * 4;
    A.coef = reinterpret_cast<int16_t *>(p);      p += (size_t)nn * 2;
    A.lv = reinterpret_cast<int16_t *>(p);
  }
  const bool chroma = jb.type != 0;
  const int transform_shift = 15 - cfg.bitdepth - lg;
  const int qp_scaled = scaled_qp(jb.type, jb.qp, (cfg.bitdepth - 8) * 6);
  const int q_bits = 14 + qp_scaled / 6 + transform_shift;
  const int q = c_rq_quant_scale[qp_scaled % 6];
  // scalinglist.c:339-356 with the flat list: 2^15 * 2^(-2 * transform_shift) / q / q / 2^(2 (bitdepth - 8))
  const double err_scale = ldexp(1.0, 15 - 2 * transform_shift) / (double)q / (double)q / (double)(1 << (2 * (cfg.bitdepth - 8)));
  const int round = 1 << (q_bits - 1);

  // ---- all lanes: stage the block, the context snapshot and the scan ----
  {
    const int16_t *src = coef_g + jb.coef_offset;
    for (int i = lane; i < nn; i += 32) A.coef[i] = src[i];
    const uint32_t *cs = reinterpret_cast<const uint32_t *>(ctxs + jb.ctx_index);
    uint32_t *cd = reinterpret_cast<uint32_t *>(&H.ctx);
    for (int i = lane; i < (int)(sizeof(kvz_cuda_rdoq_ctx) / 4); i += 32) cd[i] = cs[i];
    for (int i = lane; i < 64; i += 32) { H.group_sig[i] = 0; H.group_sig_cost[i] = 0.0; }
    if (lane == 0) { H.coded_end = 0; H.abs_sum = 0; }
  }
  build_scan_table(H.scan, jb.scan_idx, lg, lane, 32);      // ends with a barrier

  // ---- all lanes: |coef| * scale, cost of dropping the coefficient, start of the scan (rdo.c:632-661) ----
  int last_pos = -1;
  for (int sp = lane; sp < nn; sp += 32) {
    const int bp = H.scan.pos[sp];
    int scaled = abs((int)A.coef[bp]) * q;
    scaled = min(scaled, 0x7FFFFFFF - round);
    A.scaled[sp] = scaled;
    const double e0 = (double)scaled;
    A.cost_zero[sp] = e0 * e0 * err_scale;
    if (((uint32_t)(scaled + round) >> q_bits) > 0) last_pos = sp;      // ascending: the lane's largest survives
    A.lv[bp] = 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) last_pos = max(last_pos, __shfl_xor_sync(0xffffffffu, last_pos, o));
  __syncwarp();

  int16_t *dst = out_g + jb.coef_offset;
  if (last_pos < 0) {                                        // nothing survives: all levels zero (rdo.c:659-661)
    for (int i = lane; i < nn; i += 32) dst[i] = 0;
    return;
  }

  // ---- lane 0: the serial chain ----
  if (lane == 0) {
    last_prefix_bits(chroma ? H.ctx.last_x_chroma : H.ctx.last_x_luma, lg, chroma, H.bits_x);
    last_prefix_bits(chroma ? H.ctx.last_y_chroma : H.ctx.last_y_luma, lg, chroma, H.bits_y);
    serial_pass(H, A, jb, cfg.signhide != 0, lg, q_bits, err_scale, last_pos);
  }
  __syncwarp();
  const int coded_end = H.coded_end;

  // ---- all lanes: signs, uncoded tail, sum of levels (rdo.c:902-912) ----
  uint32_t abs_sum = 0;
  for (int sp = lane; sp <= last_pos; sp += 32) {
    const int bp = H.scan.pos[sp];
    if (sp < coded_end) {
      const int level = A.lv[bp];
      abs_sum += (uint32_t)level;
      A.lv[bp] = (int16_t)(A.coef[bp] < 0 ? -level : level);
    } else {
      A.lv[bp] = 0;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) abs_sum += __shfl_xor_sync(0xffffffffu, abs_sum, o);
  __syncwarp();

  // ---- sign hiding: one lane per coefficient group (rdo.c:405-540) ----
  if (cfg.signhide && abs_sum >= 2) {
    const int iq = c_rq_inv_quant_scale[qp_scaled % 6];
    const long long rd_factor =
        (long long)((double)(iq * iq * (1 << (2 * (qp_scaled / 6)))) / jb.lambda / 16.0 / (double)(1 << (2 * (cfg.bitdepth - 8))) + 0.5);
    const int top_group = (coded_end - 1) >> 4;
    for (int g = lane; g <= top_group; g += 32) hide_sign_in_group(H, A, g, top_group, rd_factor);
    __syncwarp();
  }
  for (int i = lane; i < nn; i += 32) dst[i] = A.lv[i];
}

}  // namespace

int launch_rdoq_batch(const kvz_cuda_quant_cfg &cfg, const kvz_cuda_rdoq_ctx *d_ctx, const int16_t *d_coef,
                      int16_t *d_q_coef, const kvz_cuda_rdoq_job *d_jobs, int n_jobs, int max_width, int skip4,
                      cudaStream_t stream)
{
  if (n_jobs == 0) return 1;
  const size_t smem = rdoq_smem_bytes(max_width * max_width);
  KVZ_CUDA_TRY(cudaFuncSetAttribute(rdoq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  rdoq_kernel<<<n_jobs, 32, smem, stream>>>(cfg, d_ctx, d_coef, d_q_coef, d_jobs, n_jobs, skip4);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
