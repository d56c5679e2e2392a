/*
 * oracle/kvz_oracle.c -- TEST INFRASTRUCTURE ONLY (see kvz_oracle.h).
 *
 * Plain-C restatement of the arithmetic on Kvazaar's inner encode loop.
 * Parity status: PINNED -- tests/test_oracle_*.py check every function here
 * against the reference's own known-answer vectors and against the unmodified
 * reference compiled into oracle/_ref/ (fixtures committed in tests/golden/).
 *
 * Written from the algorithm, not transliterated: transforms are plain matrix
 * products from a 33-entry cosine table, Hadamard is a natural-order FWHT,
 * edge handling is a per-pixel coordinate clamp, the four FME block filters
 * collapse into one generic fractional-sample interpolator.
 */
#include "kvz_oracle.h"

#include <stdlib.h>
#include <string.h>
#include <limits.h>

#define KMIN(a, b) ((a) < (b) ? (a) : (b))
#define KMAX(a, b) ((a) > (b) ? (a) : (b))
#define KCLIP(lo, hi, v) ((v) < (lo) ? (lo) : ((v) > (hi) ? (hi) : (v)))
#define LCU_W 64

/* ------------------------------------------------------------------ */
/* MV cost model (search_inter.c:235-412)                               */
/* ------------------------------------------------------------------ */

/* search_inter.c:235-253: 2*floor(log2(symbol+2)) without the final "+1" */
static unsigned ep_ex_golomb_bits(unsigned symbol)
{
  unsigned v = symbol + 2, lg = 0;
  while (v >> (lg + 1)) lg++;         /* lg = floor(log2(v)) */
  return 2 * lg;
}

/* search_inter.c:310-323 with CTX_FRAC_BITS = 15: the <<15, +half, >>15
 * round trip is the identity on whole bits. */
static uint32_t mvd_coding_bits(int mvd_x, int mvd_y)
{
  return ep_ex_golomb_bits((unsigned)abs(mvd_x)) + ep_ex_golomb_bits((unsigned)abs(mvd_y));
}

int kvzo_select_mv_cand(const int16_t mv_cand[2][2], int mv_x, int mv_y)
{
  /* search_inter.c:326-370, cost_out == NULL */
  if (mv_cand[0][0] == mv_cand[1][0] && mv_cand[0][1] == mv_cand[1][1]) return 0;
  uint32_t c0 = mvd_coding_bits(mv_x - mv_cand[0][0], mv_y - mv_cand[0][1]);
  uint32_t c1 = mvd_coding_bits(mv_x - mv_cand[1][0], mv_y - mv_cand[1][1]);
  return c1 < c0 ? 1 : 0;
}

uint32_t kvzo_mvd_cost(const kvzo_me_job *j, int x, int y, int mv_shift, uint32_t *bitcost)
{
  /* search_inter.c:373-412 */
  x *= 1 << mv_shift;
  y *= 1 << mv_shift;
  uint32_t bits = 0;
  int merged = 0;
  for (int i = 0; i < j->num_merge_cand; i++) {
    const kvzo_merge_cand *m = &j->merge_cand[i];
    if (m->dir == 3) continue;
    int l = m->dir - 1;
    if (m->mv[l][0] == x && m->mv[l][1] == y && j->ref_LX[l][m->ref[l]] == j->ref_idx) {
      bits = (uint32_t)i;
      merged = 1;
      break;
    }
  }
  if (!merged) {
    uint32_t c0 = mvd_coding_bits(x - j->mv_cand[0][0], y - j->mv_cand[0][1]);
    uint32_t c1 = mvd_coding_bits(x - j->mv_cand[1][0], y - j->mv_cand[1][1]);
    bits = KMIN(c0, c1);
  }
  if (bitcost) *bitcost = bits;
  return bits * (uint32_t)j->lambda_sqrt_i;
}

int kvzo_fracmv_within_tile(const kvzo_me_job *j, int x, int y)
{
  /* search_inter.c:87-172 */
  const int frac_luma = (x % 4 != 0) || (y % 4 != 0);
  const int frac_chroma = (x % 8 != 0) || (y % 8 != 0);
  if (j->owf && j->wpp) {
    int margin = frac_luma ? 4 : (frac_chroma ? 2 : 0);
    if (j->sao_type) margin += 10;            /* SAO_DELAY_PX  global.h:175 */
    else if (j->deblock_enable) margin += 8;  /* DEBLOCK_DELAY_PX global.h:163 */
    const int lcu_x = j->origin_x / LCU_W, lcu_y = j->origin_y / LCU_W;
    const int dx = ((j->origin_x + j->pu_w + margin) * 4 + x) / (LCU_W << 2) - lcu_x;
    const int dy = ((j->origin_y + j->pu_h + margin) * 4 + y) / (LCU_W << 2) - lcu_y;
    if (dy > j->max_ref_lcu_down) return 0;
    if (dx + dy > j->max_ref_lcu_down + j->max_ref_lcu_right) return 0;
  }
  if (j->mv_constraint == 0) return 1;   /* KVZ_MV_CONSTRAIN_NONE */
  int margin = 0;
  if (j->mv_constraint == 4) {           /* KVZ_MV_CONSTRAIN_FRAME_AND_TILE_MARGIN */
    margin = frac_luma ? 16 : (frac_chroma ? 8 : 0);
  }
  const int ax = j->origin_x * 4 + x, ay = j->origin_y * 4 + y;
  const int from_right = (j->width << 2) - (ax + (j->pu_w << 2));
  const int from_bottom = (j->height << 2) - (ay + (j->pu_h << 2));
  return ax >= margin && ay >= margin && from_right >= margin && from_bottom >= margin;
}

/* ------------------------------------------------------------------ */
/* Transforms (dct-generic.c:26-637)                                    */
/* ------------------------------------------------------------------ */

/* round(64*sqrt(2)*cos(m*pi/64)) as standardised for HEVC, m = 0..32,
 * with the m=0 and m=16 entries both 64. */
static const int8_t k_cos_tab[33] = {
  64, 90, 90, 90, 89, 88, 87, 85, 83, 82, 80, 78, 75, 73, 70, 67, 64,
  61, 57, 54, 50, 46, 43, 38, 36, 31, 25, 22, 18, 13, 9, 4, 0 };

static int hevc_cos(int m)   /* period 128 in m */
{
  m &= 127;
  if (m > 64) m = 128 - m;
  return m <= 32 ? k_cos_tab[m] : -k_cos_tab[64 - m];
}

/* basis row k, sample n of the N-point transform */
static int tr_coef(int N, int dst4, int k, int n)
{
  if (dst4) {
    static const int8_t dst[4][4] = {
      { 29, 55, 74, 84 }, { 74, 74, 0, -74 }, { 84, -29, -74, 55 }, { 55, -84, 74, -29 } };
    return dst[k][n];
  }
  return hevc_cos((32 / N) * k * (2 * n + 1));
}

/* one forward pass: out[k][j] = (sum_n M[k][n] in[j][n] + rnd) >> shift, truncated to int16 */
static void fwd_pass(int N, int dst4, const int16_t *in, int16_t *out, int shift)
{
  const int rnd = 1 << (shift - 1);
  for (int j = 0; j < N; j++)
    for (int k = 0; k < N; k++) {
      int32_t s = 0;
      for (int n = 0; n < N; n++) s += tr_coef(N, dst4, k, n) * in[j * N + n];
      out[k * N + j] = (int16_t)((s + rnd) >> shift);
    }
}

/* one inverse pass: out[j][n] = clip16((sum_k M[k][n] in[k][j] + rnd) >> shift) */
static void inv_pass(int N, int dst4, const int16_t *in, int16_t *out, int shift)
{
  const int rnd = 1 << (shift - 1);
  for (int j = 0; j < N; j++)
    for (int n = 0; n < N; n++) {
      int32_t s = 0;
      for (int k = 0; k < N; k++) s += tr_coef(N, dst4, k, n) * in[k * N + j];
      s = (s + rnd) >> shift;
      out[j * N + n] = (int16_t)KCLIP(-32768, 32767, s);
    }
}

static int ilog2(int n) { int l = 0; while ((1 << l) < n) l++; return l; }

void kvzo_fwd_transform(int n, int dst4, int bitdepth, const int16_t *in, int16_t *out)
{
  int16_t tmp[32 * 32];
  const int lg = ilog2(n);
  fwd_pass(n, dst4, in, tmp, lg - 1 + (bitdepth - 8));   /* dct-generic.c:571 */
  fwd_pass(n, dst4, tmp, out, lg + 6);                   /* dct-generic.c:572 */
}

void kvzo_inv_transform(int n, int dst4, int bitdepth, const int16_t *in, int16_t *out)
{
  int16_t tmp[32 * 32];
  inv_pass(n, dst4, in, tmp, 7);                         /* dct-generic.c:582 */
  inv_pass(n, dst4, tmp, out, 12 - (bitdepth - 8));      /* dct-generic.c:583 */
}

void kvzo_transformskip(int bitdepth, const int16_t *block, int16_t *coeff, int n)
{
  const int shift = 15 - bitdepth - ilog2(n);
  for (int i = 0; i < n * n; i++) coeff[i] = (int16_t)(block[i] << shift);
}

void kvzo_itransformskip(int bitdepth, int16_t *block, const int16_t *coeff, int n)
{
  const int shift = 15 - bitdepth - ilog2(n);
  const int off = 1 << (shift - 1);
  for (int i = 0; i < n * n; i++) block[i] = (int16_t)((coeff[i] + off) >> shift);
}

/* ------------------------------------------------------------------ */
/* Coefficient scans, quantisation (tables.c:13-71, quant-generic.c)    */
/* ------------------------------------------------------------------ */

/* position list of a (1<<lg)^2 square in scan order; idx 0 up-right diagonal
 * (bottom-left to top-right), 1 horizontal, 2 vertical */
static int square_scan(int scan_idx, int lg, int *xs, int *ys)
{
  const int n = 1 << lg;
  int c = 0;
  if (scan_idx == 0) {
    for (int d = 0; d <= 2 * (n - 1); d++)
      for (int x = KMAX(0, d - (n - 1)); x <= KMIN(d, n - 1); x++) { xs[c] = x; ys[c] = d - x; c++; }
  } else if (scan_idx == 1) {
    for (int y = 0; y < n; y++) for (int x = 0; x < n; x++) { xs[c] = x; ys[c] = y; c++; }
  } else {
    for (int x = 0; x < n; x++) for (int y = 0; y < n; y++) { xs[c] = x; ys[c] = y; c++; }
  }
  return c;
}

void kvzo_scan_order(int scan_idx, int log2_size, uint32_t *out)
{
  const int n = 1 << log2_size;
  int cx[64], cy[64], px[16], py[16];
  if (log2_size < 2) {         /* 2x2: a single group scanned directly */
    int c = square_scan(scan_idx, log2_size, px, py);
    for (int i = 0; i < c; i++) out[i] = (uint32_t)(py[i] * n + px[i]);
    return;
  }
  const int ncg = square_scan(scan_idx, log2_size - 2, cx, cy);
  square_scan(scan_idx, 2, px, py);
  int c = 0;
  for (int g = 0; g < ncg; g++)
    for (int i = 0; i < 16; i++)
      out[c++] = (uint32_t)((cy[g] * 4 + py[i]) * n + cx[g] * 4 + px[i]);
}

int32_t kvzo_scaled_qp(int type, int qp, int qp_offset)
{
  /* transform.c:129-143; chroma QP mapping table of the HEVC spec (8-5) */
  static const uint8_t chroma_map[58] = {
    0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25,
    26, 27, 28, 29, 29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37, 38, 39, 40, 41, 42,
    43, 44, 45, 46, 47, 48, 49, 50, 51 };
  if (type == 0) return qp + qp_offset;
  int q = KCLIP(-qp_offset, 57, qp);
  return q < 0 ? q + qp_offset : chroma_map[q] + qp_offset;
}

static const int32_t k_quant_scale[6] = { 26214, 23302, 20560, 18396, 16384, 14564 };  /* scalinglist.c:66 */
static const int32_t k_inv_quant_scale[6] = { 40, 45, 51, 57, 64, 72 };                /* scalinglist.c:67 */

void kvzo_quant(const kvzo_q_cfg *c, const int16_t *coef, int16_t *q_coef, int width, int type, int scan_idx)
{
  /* quant-generic.c:37-163 with the flat (disabled) scaling list:
   * quant_coeff[n] == k_quant_scale[qp % 6] for every n (scalinglist.c:374) */
  const int lg = ilog2(width);
  const int nn = width * width;
  const int qp_scaled = kvzo_scaled_qp(type, c->qp, (c->bitdepth - 8) * 6);
  const int64_t qc = k_quant_scale[qp_scaled % 6];
  const int q_bits = 14 + qp_scaled / 6 + (15 - c->bitdepth - lg);
  const int32_t add = (c->slice_is_intra ? 171 : 85) << (q_bits - 9);
  uint32_t ac_sum = 0;

  for (int n = 0; n < nn; n++) {
    int32_t a = abs(coef[n]);
    int32_t level = (int32_t)((a * qc + add) >> q_bits);
    ac_sum += (uint32_t)level;
    if (coef[n] < 0) level = -level;
    q_coef[n] = (int16_t)KCLIP(-32768, 32767, level);
  }
  if (!c->signhide || ac_sum < 2) return;

  /* sign-bit hiding, one 16-coefficient group at a time in reverse scan order */
  uint32_t scan[32 * 32];
  int32_t delta[32 * 32];
  kvzo_scan_order(scan_idx, lg, scan);
  for (int n = 0; n < nn; n++) {
    int64_t prod = (int64_t)abs(coef[n]) * qc;
    int32_t level = (int32_t)((prod + add) >> q_bits);
    delta[n] = (int32_t)((prod - ((int64_t)level << q_bits)) >> (q_bits - 8));
  }
  int seen_last_group = 0;
  for (int g = (nn - 1) >> 4; g >= 0; g--) {
    const uint32_t *sc = scan + (g << 4);
    int first = 16, last = -1, sum = 0;
    for (int n = 15; n >= 0; n--) if (q_coef[sc[n]]) { last = n; break; }
    for (int n = 0; n < 16; n++) if (q_coef[sc[n]]) { first = n; break; }
    for (int n = first; n <= last; n++) sum += q_coef[sc[n]];
    const int is_last_group = (last >= 0 && !seen_last_group);
    if (last >= 0) seen_last_group = 1;
    if (last - first < 4) continue;
    const int signbit = q_coef[sc[first]] > 0 ? 0 : 1;
    if (signbit == (sum & 1)) continue;

    int32_t best = INT_MAX, pos = -1;
    int change = 0, cur_change = 0;
    for (int n = is_last_group ? last : 15; n >= 0; n--) {
      const uint32_t p = sc[n];
      int32_t cost = INT_MAX;
      /* note: cur_change is deliberately NOT reset per coefficient, like the
       * reference (quant-generic.c:117-146) */
      if (q_coef[p] != 0) {
        if (delta[p] > 0) { cost = -delta[p]; cur_change = 1; }
        else if (n == first && abs(q_coef[p]) == 1) { cost = INT_MAX; }
        else { cost = delta[p]; cur_change = -1; }
      } else if (n < first && ((coef[p] >= 0) ? 0 : 1) != signbit) {
        cost = INT_MAX;
      } else {
        cost = -delta[p]; cur_change = 1;
      }
      if (cost < best) { best = cost; change = cur_change; pos = (int32_t)p; }
    }
    if (q_coef[pos] == 32767 || q_coef[pos] == -32768) change = -1;
    if (coef[pos] >= 0) q_coef[pos] = (int16_t)(q_coef[pos] + change);
    else q_coef[pos] = (int16_t)(q_coef[pos] - change);
  }
}

void kvzo_dequant(const kvzo_q_cfg *c, const int16_t *q_coef, int16_t *coef, int width, int type)
{
  /* quant-generic.c:279-321, scaling list disabled branch */
  const int lg = ilog2(width);
  const int transform_shift = 15 - c->bitdepth - lg;
  const int qp_scaled = kvzo_scaled_qp(type, c->qp, (c->bitdepth - 8) * 6);
  const int shift = 20 - 14 - transform_shift;
  const int32_t scale = k_inv_quant_scale[qp_scaled % 6] << (qp_scaled / 6);
  const int32_t add = 1 << (shift - 1);
  for (int n = 0; n < width * width; n++) {
    int32_t v = (q_coef[n] * scale + add) >> shift;
    coef[n] = (int16_t)KCLIP(-32768, 32767, v);
  }
}

uint32_t kvzo_coeff_abs_sum(const int16_t *c, uint32_t n)
{
  uint32_t s = 0;
  for (uint32_t i = 0; i < n; i++) s += (uint32_t)abs(c[i]);
  return s;
}

/* ------------------------------------------------------------------ */
/* pixel-typed half, instantiated for 8- and 10-bit builds              */
/* ------------------------------------------------------------------ */
#define PIX uint8_t
#define BD 8
#define FN(name) name##_8
#include "kvz_oracle_px.inc"
#undef PIX
#undef BD
#undef FN

#define PIX uint16_t
#define BD 10
#define FN(name) name##_10
#include "kvz_oracle_px.inc"
#undef PIX
#undef BD
#undef FN
