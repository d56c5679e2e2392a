// kvazaar_b200/csrc/coeff_bits.cu
//
// CABAC-exact bit cost of coefficient blocks (A15's kvz_get_coeff_cost): what
// get_coeff_cabac_cost (rdo.c:158-197) returns after running kvz_encode_coeff_nxn
// (encode_coding_tree.c:103-345) on a copy of the CABAC state in counting mode.  In that mode kvz_cabac_write
// only counts bytes (cabac.c:128-137), so the result -- (23 - bits_left) + 8 * num_buffered_bytes -- is the total
// of the coder's renormalisation shifts: a function of the coder `range` and the context states alone.
//
// One warp per block.  All lanes: stage the block and the context snapshot, build the scan, mark the
// significant coefficient groups, find the last coefficient.  The group flags are final before the walk starts
// (the two groups the walk forces to 1 are the last one, which holds the last coefficient, and group 0, which
// nobody looks at afterwards), so lanes also precompute every position's significance context.  Lane 0 then
// runs the chain the reference runs: the context-coded bins in coding order, each updating `range` and its
// context state; bypass bins only add their count.
#include <cstdlib>
#include "common.cuh"
#include "transform.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

// HEVC table 9-46 (rangeTabLps), cabac.c:52-67
__device__ __constant__ uint8_t c_range_lps[64][4] = {
  {128, 176, 208, 240}, {128, 167, 197, 227}, {128, 158, 187, 216}, {123, 150, 178, 205}, {116, 142, 169, 195},
  {111, 135, 160, 185}, {105, 128, 152, 175}, {100, 122, 144, 166}, { 95, 116, 137, 158}, { 90, 110, 130, 150},
  { 85, 104, 123, 142}, { 81,  99, 117, 135}, { 77,  94, 111, 128}, { 73,  89, 105, 122}, { 69,  85, 100, 116},
  { 66,  80,  95, 110}, { 62,  76,  90, 104}, { 59,  72,  86,  99}, { 56,  69,  81,  94}, { 53,  65,  77,  89},
  { 51,  62,  73,  85}, { 48,  59,  69,  80}, { 46,  56,  66,  76}, { 43,  53,  63,  72}, { 41,  50,  59,  69},
  { 39,  48,  56,  65}, { 37,  45,  54,  62}, { 35,  43,  51,  59}, { 33,  41,  48,  56}, { 32,  39,  46,  53},
  { 30,  37,  43,  50}, { 29,  35,  41,  48}, { 27,  33,  39,  45}, { 26,  31,  37,  43}, { 24,  30,  35,  41},
  { 23,  28,  33,  39}, { 22,  27,  32,  37}, { 21,  26,  30,  35}, { 20,  24,  29,  33}, { 19,  23,  27,  31},
  { 18,  22,  26,  30}, { 17,  21,  25,  28}, { 16,  20,  23,  27}, { 15,  19,  22,  25}, { 14,  18,  21,  24},
  { 14,  17,  20,  23}, { 13,  16,  19,  22}, { 12,  15,  18,  21}, { 12,  14,  17,  20}, { 11,  14,  16,  19},
  { 11,  13,  15,  18}, { 10,  12,  15,  17}, { 10,  12,  14,  16}, {  9,  11,  13,  15}, {  9,  11,  12,  14},
  {  8,  10,  12,  14}, {  8,   9,  11,  13}, {  7,   9,  11,  12}, {  7,   9,  10,  12}, {  7,   8,  10,  11},
  {  6,   8,   9,  11}, {  6,   7,   9,  10}, {  6,   7,   8,   9}, {  2,   2,   2,   2} };
// HEVC table 9-47 (transIdxLps)
__device__ __constant__ uint8_t c_next_lps[64] = { 0, 0, 1, 2, 2, 4, 4, 5, 6, 7, 8, 9, 9, 11, 11, 12, 13, 13, 15, 15, 16, 16,
  18, 18, 19, 19, 21, 21, 22, 22, 23, 24, 24, 25, 26, 26, 27, 27, 28, 29, 29, 30, 30, 30, 31, 32, 32, 33, 33, 33, 34,
  34, 35, 35, 35, 36, 36, 36, 37, 37, 37, 38, 38, 63 };
// kvz_g_auc_renorm_table (cabac.c:69-73)
__device__ __constant__ uint8_t c_renorm[32] = { 6, 5, 4, 4, 3, 3, 3, 3, 2, 2, 2, 2, 2, 2, 2, 2,
                                                 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1 };
__device__ __constant__ uint8_t c_cb_last_group[32] = { 0, 1, 2, 3, 4, 4, 5, 5, 6, 6, 6, 6, 7, 7, 7, 7,   // encoderstate.h:345-349
                                                        8, 8, 8, 8, 8, 8, 8, 8, 9, 9, 9, 9, 9, 9, 9, 9 };
__device__ __constant__ uint8_t c_cb_sig_map4x4[16] = { 0, 1, 4, 5, 2, 3, 4, 5, 6, 6, 8, 8, 7, 7, 8, 8 };   // context.c:357-363

// `range`, the bit count, and where the coder's tables live: constant memory for the warp-per-block kernel (one lane
// reads them), a shared-memory copy for the thread-per-block kernel (32 lanes read 32 different entries, which
// the constant cache would serialise)
struct Counter {
  uint32_t range, bits;
  const uint8_t *lps, *next, *renorm;
};

// kvz_cabac_encode_bin (cabac.c:91-120) without `low`; state = (pStateIdx << 1) | valMps
__device__ __forceinline__ void count_bin(Counter &b, uint8_t &state, int bin)
{
  const uint32_t st = state >> 1, lps = b.lps[st * 4 + ((b.range >> 6) & 3)];
  b.range -= lps;
  if (bin != (int)(state & 1)) {
    const int shift = b.renorm[lps >> 3];
    b.range = lps << shift;
    b.bits += shift;
    state = st == 0 ? (uint8_t)(state ^ 1) : (uint8_t)((b.next[st] << 1) | (state & 1));
  } else {
    if (st < 62) state += 2;
    if (b.range < 256) { b.range <<= 1; b.bits += 1; }
  }
}

// context.c:354-385
__device__ int cb_sig_ctx_inc(int pattern, int scan_idx, int x, int y, int lg, bool chroma, const uint8_t *map4 = c_cb_sig_map4x4)
{
  if (x + y == 0) return 0;
  if (lg == 2) return map4[4 * y + x];
  const int base = lg == 3 ? (scan_idx == 0 ? 9 : 15) : (chroma ? 12 : 21);
  const int xs = x & 3, ys = y & 3;
  int cnt;
  if (pattern == 0) cnt = xs + ys == 0 ? 2 : (xs + ys <= 2 ? 1 : 0);
  else if (pattern == 1) cnt = ys == 0 ? 2 : (ys == 1 ? 1 : 0);
  else if (pattern == 2) cnt = xs == 0 ? 2 : (xs == 1 ? 1 : 0);
  else cnt = 2;
  return ((!chroma && ((x >> 2) + (y >> 2)) > 0) ? 3 : 0) + base + cnt;
}

struct Head {
  ScanTable scan;
  kvz_cuda_rdoq_ctx ctx;
  uint8_t group_sig[64];
};

__host__ __device__ inline size_t coeff_bits_smem(int nn)
{
  return ((sizeof(Head) + 15) & ~(size_t)15) + (size_t)nn * 3;     // coefficients by block position + sig context by scan position
}

__global__ void __launch_bounds__(32)
coeff_bits_kernel(const kvz_cuda_rdoq_ctx *__restrict__ ctxs, const int16_t *__restrict__ coef_g,
                  const kvz_cuda_coeff_job *__restrict__ jobs, int signhide, int trskip_enable, uint32_t *__restrict__ bits_out)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x;
  const kvz_cuda_coeff_job jb = jobs[blockIdx.x];
  const int width = jb.width, lg = ilog2_pow2(width), nn = width * width, side = width >> 2;
  const bool chroma = jb.type != 0;
  Head &H = *reinterpret_cast<Head *>(smem_raw);
  int16_t *coef = reinterpret_cast<int16_t *>(smem_raw + ((sizeof(Head) + 15) & ~(size_t)15));
  uint8_t *sig_ctx = reinterpret_cast<uint8_t *>(coef + nn);

  // ---- all lanes: stage, scan, group flags, last position ----
  {
    const int16_t *src = coef_g + jb.coef_offset;
    for (int i = lane; i < nn; i += 32) coef[i] = src[i];
    const uint32_t *cs = reinterpret_cast<const uint32_t *>(ctxs + jb.ctx_index);
    uint32_t *cd = reinterpret_cast<uint32_t *>(&H.ctx);
    for (int i = lane; i < (int)(sizeof(kvz_cuda_rdoq_ctx) / 4); i += 32) cd[i] = cs[i];
    for (int i = lane; i < 64; i += 32) H.group_sig[i] = 0;
  }
  build_scan_table(H.scan, jb.scan_idx, lg, lane, 32);      // ends with a barrier
  int last_scan = -1;
  for (int sp = lane; sp < nn; sp += 32) {
    const int bp = H.scan.pos[sp];
    if (coef[bp] != 0) {
      last_scan = sp;                                        // ascending: the lane's largest survives
      H.group_sig[((bp >> lg) >> 2) * side + ((bp & (width - 1)) >> 2)] = 1;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) last_scan = max(last_scan, __shfl_xor_sync(0xffffffffu, last_scan, o));
  __syncwarp();
  if (last_scan < 0) {                                       // rdo.c:166-173
    if (lane == 0) bits_out[blockIdx.x] = 0;
    return;
  }
  for (int sp = lane; sp <= last_scan; sp += 32) {
    const int bp = H.scan.pos[sp], py = bp >> lg, px = bp & (width - 1), gx = px >> 2, gy = py >> 2;
    int pattern = -1;
    if (width != 4) {                                        // kvz_context_calc_pattern_sig_ctx (context.c:327-339)
      const int r = gx < side - 1 ? H.group_sig[gy * side + gx + 1] : 0;
      const int d = gy < side - 1 ? H.group_sig[(gy + 1) * side + gx] : 0;
      pattern = r + (d << 1);
    }
    sig_ctx[sp] = (uint8_t)cb_sig_ctx_inc(pattern, jb.scan_idx, px, py, lg, chroma);
  }
  __syncwarp();
  if (lane != 0) return;

  // ---- lane 0: the chain ----
  Counter b = { jb.range, 0, &c_range_lps[0][0], c_next_lps, c_renorm };
  kvz_cuda_rdoq_ctx &ctx = H.ctx;
  const int cg_last = last_scan >> 4, pos_last = H.scan.pos[last_scan];
  if (width == 4 && trskip_enable) count_bin(b, chroma ? ctx.trskip_chroma : ctx.trskip_luma, 0);
  {
    // encode_last_significant_xy (encode_coding_tree.c:49-101)
    int lx = pos_last & (width - 1), ly = pos_last >> lg;
    const int index = lg - 2;
    const int ctx_offset = chroma ? 0 : index * 3 + (index + 1) / 4;
    const int shift = chroma ? index : (index + 3) / 4;
    uint8_t *mx = chroma ? ctx.last_x_chroma : ctx.last_x_luma, *my = chroma ? ctx.last_y_chroma : ctx.last_y_luma;
    if (jb.scan_idx == 2) { const int t = lx; lx = ly; ly = t; }
    const int gx = c_cb_last_group[lx], gy = c_cb_last_group[ly], gmax = c_cb_last_group[width - 1];
    for (int i = 0; i < gx; i++) count_bin(b, mx[ctx_offset + (i >> shift)], 1);
    if (gx < gmax) count_bin(b, mx[ctx_offset + (gx >> shift)], 0);
    for (int i = 0; i < gy; i++) count_bin(b, my[ctx_offset + (i >> shift)], 1);
    if (gy < gmax) count_bin(b, my[ctx_offset + (gy >> shift)], 0);
    if (gx > 3) b.bits += (gx - 2) / 2;
    if (gy > 3) b.bits += (gy - 2) / 2;
  }
  uint8_t *sig_models = chroma ? ctx.sig_chroma : ctx.sig_luma;
  int c1 = 1, scan_pos = last_scan;
  for (int g = cg_last; g >= 0; g--) {
    const int sub_pos = g << 4;
    const int cgx = H.scan.gx[g], cgy = H.scan.gy[g], cg_blk = cgy * side + cgx;
    int abs_coeff[16], n_nz = 0, last_nz = -1, first_nz = 16;
    uint32_t rice = 0;
    if (scan_pos == last_scan) {
      abs_coeff[0] = abs((int)coef[pos_last]);
      n_nz = 1; last_nz = first_nz = scan_pos;
      scan_pos--;
    }
    bool coded = H.group_sig[cg_blk] != 0;
    if (g == cg_last || g == 0) {
      coded = true;
    } else {
      const int r = cgx < side - 1 ? H.group_sig[cgy * side + cgx + 1] : 0;      // kvz_context_get_sig_coeff_group
      const int d = cgy < side - 1 ? H.group_sig[(cgy + 1) * side + cgx] : 0;
      count_bin(b, ctx.sig_group[jb.type + ((r | d) != 0)], coded);
    }
    if (coded) {
      for (; scan_pos >= sub_pos; scan_pos--) {
        const int level = coef[H.scan.pos[scan_pos]];
        const int sig = level != 0;
        if (scan_pos > sub_pos || g == 0 || n_nz) count_bin(b, sig_models[sig_ctx[scan_pos]], sig);
        if (sig) {
          abs_coeff[n_nz++] = abs(level);
          if (last_nz == -1) last_nz = scan_pos;
          first_nz = scan_pos;
        }
      }
    } else {
      scan_pos = sub_pos - 1;
    }
    if (n_nz > 0) {
      const bool sign_hidden = last_nz - first_nz >= 4;
      int ctx_set = (g > 0 && !chroma) ? 2 : 0;
      if (c1 == 0) ctx_set++;
      c1 = 1;
      uint8_t *one = (chroma ? ctx.one_chroma : ctx.one_luma) + 4 * ctx_set;
      const int n_c1 = min(n_nz, 8);                         // C1FLAG_NUMBER
      int first_c2 = -1;
      for (int i = 0; i < n_c1; i++) {
        const int sym = abs_coeff[i] > 1;
        count_bin(b, one[c1], sym);
        if (sym) { c1 = 0; if (first_c2 == -1) first_c2 = i; }
        else if (c1 < 3 && c1 > 0) c1++;
      }
      if (c1 == 0 && first_c2 != -1)
        count_bin(b, (chroma ? ctx.abs_chroma : ctx.abs_luma)[ctx_set], abs_coeff[first_c2] > 2);
      b.bits += (signhide && sign_hidden) ? n_nz - 1 : n_nz;  // coeff_sign_flag, bypass
      if (c1 == 0 || n_nz > 8) {
        int first_coeff2 = 1;
        for (int i = 0; i < n_nz; i++) {
          const int base = i < 8 ? 2 + first_coeff2 : 1;
          if (abs_coeff[i] >= base) {
            // kvz_cabac_write_coeff_remain (cabac.c:317-337): bypass bins only
            const uint32_t symbol = (uint32_t)(abs_coeff[i] - base);
            if (symbol < (3u << rice)) {
              b.bits += (symbol >> rice) + 1 + rice;
            } else {
              uint32_t length = rice, code = symbol - (3u << rice);
              while (code >= (1u << length)) { code -= 1u << length; length++; }
              b.bits += (3 + length + 1 - rice) + length;
            }
            if (abs_coeff[i] > 3 * (1 << rice)) rice = min(rice + 1, 4u);
          }
          if (abs_coeff[i] >= 2) first_coeff2 = 0;
        }
      }
    }
  }
  bits_out[blockIdx.x] = b.bits;
}

// ---- blocks up to 8x8: one THREAD per block ----
// The chain is serial per block whatever the width; a warp per block leaves 31 lanes idle during it.  For 4x4 and
// 8x8 blocks (the bulk of a frame) every lane runs the chain of its own block: coefficients in registers / local
// memory, its private copy of the context snapshot in shared memory (the chain updates it), scan order from a
// constant table filled by the host on first use.
constexpr int kSmallT = 64;
__device__ __constant__ uint8_t c_small_scan[2][3][64];       // [log2 width - 2][scan_idx][scan position] -> block position

__global__ void __launch_bounds__(kSmallT)
coeff_bits_small_kernel(const kvz_cuda_rdoq_ctx *__restrict__ ctxs, const int16_t *__restrict__ coef_g,
                        const kvz_cuda_coeff_job *__restrict__ jobs, int n_jobs, int signhide, int trskip_enable,
                        uint32_t *__restrict__ bits_out)
{
  __shared__ kvz_cuda_rdoq_ctx s_ctx[kSmallT];
  __shared__ uint8_t s_lps[256], s_next[64], s_renorm[32], s_last_group[32], s_map4[16], s_scan[2 * 3 * 64];
  for (int i = threadIdx.x; i < 256; i += kSmallT) s_lps[i] = (&c_range_lps[0][0])[i];
  for (int i = threadIdx.x; i < 64; i += kSmallT) s_next[i] = c_next_lps[i];
  for (int i = threadIdx.x; i < 32; i += kSmallT) { s_renorm[i] = c_renorm[i]; s_last_group[i] = c_cb_last_group[i]; }
  for (int i = threadIdx.x; i < 16; i += kSmallT) s_map4[i] = c_cb_sig_map4x4[i];
  for (int i = threadIdx.x; i < 2 * 3 * 64; i += kSmallT) s_scan[i] = (&c_small_scan[0][0][0])[i];
  __syncthreads();
  const int ji = blockIdx.x * kSmallT + threadIdx.x;
  if (ji >= n_jobs) return;
  const kvz_cuda_coeff_job jb = jobs[ji];
  const int width = jb.width, lg = ilog2_pow2(width), nn = width * width, side = width >> 2;
  const bool chroma = jb.type != 0;
  const uint8_t *scan = s_scan + ((lg - 2) * 3 + jb.scan_idx) * 64;
  alignas(8) int16_t coef[64];
  {
    const int16_t *src16 = coef_g + jb.coef_offset;
    if ((reinterpret_cast<uintptr_t>(src16) & 7) == 0) {       // the usual case: blocks packed back to back
      const uint2 *src = reinterpret_cast<const uint2 *>(src16);
      uint2 *dst = reinterpret_cast<uint2 *>(coef);
      for (int i = 0; i < nn / 4; i++) dst[i] = src[i];
    } else {
      for (int i = 0; i < nn; i++) coef[i] = src16[i];
    }
  }
  unsigned group_sig = 0;                                     // bit gy * side + gx
  int last_scan = -1;
  for (int sp = 0; sp < nn; sp++) {
    const int bp = scan[sp];
    if (coef[bp] != 0) { last_scan = sp; group_sig |= 1u << (((bp >> lg) >> 2) * side + ((bp & (width - 1)) >> 2)); }
  }
  if (last_scan < 0) { bits_out[ji] = 0; return; }
  kvz_cuda_rdoq_ctx &ctx = s_ctx[threadIdx.x];
  {
    const uint32_t *cs = reinterpret_cast<const uint32_t *>(ctxs + jb.ctx_index);
    uint32_t *cd = reinterpret_cast<uint32_t *>(&ctx);
    for (int i = 0; i < (int)(sizeof(kvz_cuda_rdoq_ctx) / 4); i++) cd[i] = cs[i];
  }
  Counter b = { jb.range, 0, s_lps, s_next, s_renorm };
  const int cg_last = last_scan >> 4, pos_last = scan[last_scan];
  if (width == 4 && trskip_enable) count_bin(b, chroma ? ctx.trskip_chroma : ctx.trskip_luma, 0);
  {
    int lx = pos_last & (width - 1), ly = pos_last >> lg;
    const int index = lg - 2;
    const int ctx_offset = chroma ? 0 : index * 3 + (index + 1) / 4;
    const int shift = chroma ? index : (index + 3) / 4;
    uint8_t *mx = chroma ? ctx.last_x_chroma : ctx.last_x_luma, *my = chroma ? ctx.last_y_chroma : ctx.last_y_luma;
    if (jb.scan_idx == 2) { const int t = lx; lx = ly; ly = t; }
    const int gx = s_last_group[lx], gy = s_last_group[ly], gmax = s_last_group[width - 1];
    for (int i = 0; i < gx; i++) count_bin(b, mx[ctx_offset + (i >> shift)], 1);
    if (gx < gmax) count_bin(b, mx[ctx_offset + (gx >> shift)], 0);
    for (int i = 0; i < gy; i++) count_bin(b, my[ctx_offset + (i >> shift)], 1);
    if (gy < gmax) count_bin(b, my[ctx_offset + (gy >> shift)], 0);
    if (gx > 3) b.bits += (gx - 2) / 2;
    if (gy > 3) b.bits += (gy - 2) / 2;
  }
  uint8_t *sig_models = chroma ? ctx.sig_chroma : ctx.sig_luma;
  int c1 = 1, scan_pos = last_scan;
  for (int g = cg_last; g >= 0; g--) {
    const int sub_pos = g << 4, gbp = scan[sub_pos];
    const int cgx = (gbp & (width - 1)) >> 2, cgy = (gbp >> lg) >> 2;
    int abs_coeff[16], n_nz = 0, last_nz = -1, first_nz = 16;
    uint32_t rice = 0;
    if (scan_pos == last_scan) {
      abs_coeff[0] = abs((int)coef[pos_last]);
      n_nz = 1; last_nz = first_nz = scan_pos;
      scan_pos--;
    }
    const int r = cgx < side - 1 ? (group_sig >> (cgy * side + cgx + 1)) & 1 : 0;
    const int d = cgy < side - 1 ? (group_sig >> ((cgy + 1) * side + cgx)) & 1 : 0;
    bool coded = (group_sig >> (cgy * side + cgx)) & 1;
    if (g == cg_last || g == 0) coded = true;
    else count_bin(b, ctx.sig_group[jb.type + ((r | d) != 0)], coded);
    if (coded) {
      const int pattern = width == 4 ? -1 : r + (d << 1);
      for (; scan_pos >= sub_pos; scan_pos--) {
        const int bp = scan[scan_pos], level = coef[bp], sig = level != 0;
        if (scan_pos > sub_pos || g == 0 || n_nz)
          count_bin(b, sig_models[cb_sig_ctx_inc(pattern, jb.scan_idx, bp & (width - 1), bp >> lg, lg, chroma, s_map4)], sig);
        if (sig) {
          abs_coeff[n_nz++] = abs(level);
          if (last_nz == -1) last_nz = scan_pos;
          first_nz = scan_pos;
        }
      }
    } else {
      scan_pos = sub_pos - 1;
    }
    if (n_nz > 0) {
      const bool sign_hidden = last_nz - first_nz >= 4;
      int ctx_set = (g > 0 && !chroma) ? 2 : 0;
      if (c1 == 0) ctx_set++;
      c1 = 1;
      uint8_t *one = (chroma ? ctx.one_chroma : ctx.one_luma) + 4 * ctx_set;
      const int n_c1 = min(n_nz, 8);
      int first_c2 = -1;
      for (int i = 0; i < n_c1; i++) {
        const int sym = abs_coeff[i] > 1;
        count_bin(b, one[c1], sym);
        if (sym) { c1 = 0; if (first_c2 == -1) first_c2 = i; }
        else if (c1 < 3 && c1 > 0) c1++;
      }
      if (c1 == 0 && first_c2 != -1)
        count_bin(b, (chroma ? ctx.abs_chroma : ctx.abs_luma)[ctx_set], abs_coeff[first_c2] > 2);
      b.bits += (signhide && sign_hidden) ? n_nz - 1 : n_nz;
      if (c1 == 0 || n_nz > 8) {
        int first_coeff2 = 1;
        for (int i = 0; i < n_nz; i++) {
          const int base = i < 8 ? 2 + first_coeff2 : 1;
          if (abs_coeff[i] >= base) {
            const uint32_t symbol = (uint32_t)(abs_coeff[i] - base);
            if (symbol < (3u << rice)) {
              b.bits += (symbol >> rice) + 1 + rice;
            } else {
              uint32_t length = rice, code = symbol - (3u << rice);
              while (code >= (1u << length)) { code -= 1u << length; length++; }
              b.bits += (3 + length + 1 - rice) + length;
            }
            if (abs_coeff[i] > 3 * (1 << rice)) rice = min(rice + 1, 4u);
          }
          if (abs_coeff[i] >= 2) first_coeff2 = 0;
        }
      }
    }
  }
  bits_out[ji] = b.bits;
}

// the scan orders of 4x4 and 8x8 blocks, as build_scan_table_t (transform.cuh) makes them
void host_small_scans(uint8_t out[2][3][64])
{
  for (int lgi = 0; lgi < 2; lgi++)
    for (int scan = 0; scan < 3; scan++) {
      const int n = 4 << lgi, gs = n >> 2;
      uint8_t gx[4], gy[4], px[16], py[16];
      auto square = [](int scan_idx, int m, uint8_t *xs, uint8_t *ys) {
        int c = 0;
        if (scan_idx == 0) {
          for (int d = 0; d <= 2 * (m - 1); d++)
            for (int x = (d - (m - 1) > 0 ? d - (m - 1) : 0); x <= (d < m - 1 ? d : m - 1); x++) { xs[c] = (uint8_t)x; ys[c] = (uint8_t)(d - x); c++; }
        } else if (scan_idx == 1) {
          for (int y = 0; y < m; y++) for (int x = 0; x < m; x++) { xs[c] = (uint8_t)x; ys[c] = (uint8_t)y; c++; }
        } else {
          for (int x = 0; x < m; x++) for (int y = 0; y < m; y++) { xs[c] = (uint8_t)x; ys[c] = (uint8_t)y; c++; }
        }
      };
      square(scan, gs, gx, gy);
      square(scan, 4, px, py);
      for (int i = 0; i < n * n; i++) {
        const int g = i >> 4, k = i & 15;
        out[lgi][scan][i] = (uint8_t)((gy[g] * 4 + py[k]) * n + gx[g] * 4 + px[k]);
      }
    }
}

}  // namespace

int launch_coeff_bits(const kvz_cuda_rdoq_ctx *d_ctx, const int16_t *d_coeffs, const kvz_cuda_coeff_job *d_jobs,
                      int n_jobs, int max_width, int signhide, int trskip_enable, uint32_t *d_bits, cudaStream_t stream)
{
  if (n_jobs == 0) return 1;
  if (max_width <= 8 && !getenv("KVZ_COEFF_BITS_WARP")) {
    static bool scans_ready = false;
    if (!scans_ready) {
      uint8_t tab[2][3][64] = {};
      host_small_scans(tab);
      KVZ_CUDA_TRY(cudaMemcpyToSymbol(c_small_scan, tab, sizeof(tab)));
      scans_ready = true;
    }
    coeff_bits_small_kernel<<<(n_jobs + kSmallT - 1) / kSmallT, kSmallT, 0, stream>>>(d_ctx, d_coeffs, d_jobs, n_jobs, signhide,
                                                                                    trskip_enable, d_bits);
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
    return 1;
  }
  const size_t smem = coeff_bits_smem(max_width * max_width);
  KVZ_CUDA_TRY(cudaFuncSetAttribute(coeff_bits_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  coeff_bits_kernel<<<n_jobs, 32, smem, stream>>>(d_ctx, d_coeffs, d_jobs, signhide, trskip_enable, d_bits);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
