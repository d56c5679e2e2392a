/*
 * oracle/kvz_oracle_rdoq.c -- TEST INFRASTRUCTURE ONLY (see kvz_oracle.h).
 *
 * CPU restatement of Kvazaar's rate-distortion optimised quantisation:
 *   kvz_rdoq               rdo.c:548-885
 *   kvz_get_coded_level    rdo.c:297-339
 *   kvz_get_ic_rate        rdo.c:233-278
 *   calc_last_bits         rdo.c:366-395, get_rate_last rdo.c:350-364
 *   kvz_rdoq_sign_hiding   rdo.c:405-540
 *   context selection      context.c:303-385
 * Parity status: PINNED -- tests/test_oracle_rdoq.py compares every output level with the
 * unmodified reference (oracle/_ref, ref_shim_rdoq_ctx) for all sizes, scans, luma/chroma,
 * intra/inter, sign hiding on/off and random context states.
 *
 * The algorithm is a serial chain in double precision; the ORDER of every floating-point
 * operation below is the reference's, because additions do not associate.  What is restated
 * differently: the pass is organised as three phases over explicit per-position records
 * (level choice per group, group zeroing, last-position search), the scan tables are generated
 * (kvzo_scan_order) instead of stored, and the flat scaling list is folded into two scalars.
 */
#include "kvz_oracle.h"

#include <stdlib.h>
#include <string.h>
#include <limits.h>

#define ONE_BIT 32768            /* 1 << CTX_FRAC_BITS, rdo.h:65-66 */
#define BIG_COST 1.7e+308        /* MAX_DOUBLE, global.h:269 */

/* Fractional bit cost of coding a bin in a given context state, indexed by state ^ bin.  These are
 * the entropy estimates of the HEVC test model (HM 12.0) that rdo.c:53-63 tabulates. */
static const uint32_t k_entropy_bits[128] = {
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
static int32_t bin_bits(uint8_t state, int bin) { return (int32_t)k_entropy_bits[state ^ bin]; }

static const int32_t k_quant_scale[6] = { 26214, 23302, 20560, 18396, 16384, 14564 };  /* scalinglist.c:66 */
static const int32_t k_inv_quant_scale[6] = { 40, 45, 51, 57, 64, 72 };                /* scalinglist.c:67 */
/* last-position prefix group of a coordinate (encoderstate.h:345-349) */
static const uint8_t k_last_group[32] = { 0, 1, 2, 3, 4, 4, 5, 5, 6, 6, 6, 6, 7, 7, 7, 7,
                                          8, 8, 8, 8, 8, 8, 8, 8, 9, 9, 9, 9, 9, 9, 9, 9 };

/* the context models of one colour channel */
typedef struct {
  const uint8_t *sig, *one, *abs, *last_x, *last_y;
} chan_models;

/* state of the greater-than-1 / greater-than-2 / Rice derivation inside a coefficient group */
typedef struct {
  int ctx_set, c1, c2, rice;
  unsigned c1_idx, c2_idx;
} level_state;

/* rdo.c:233-278: bits (x 2^15) of coding |level|, sign bin included */
static int32_t level_rate(const chan_models *m, uint32_t level, const level_state *s)
{
  const int one_ctx = 4 * s->ctx_set + s->c1, abs_ctx = s->ctx_set + s->c2;
  const int g1_coded = s->c1_idx < 8, g2_coded = g1_coded && s->c2_idx < 1;   /* C1FLAG_NUMBER, C2FLAG_NUMBER */
  const uint32_t base = g1_coded ? (s->c2_idx < 1 ? 3u : 2u) : 1u;
  int32_t rate = ONE_BIT;
  if (level >= base) {
    int32_t rem = (int32_t)(level - base);
    const int k = s->rice;
    if (rem < (3 << k)) {
      rate += ((rem >> k) + 1 + k) * ONE_BIT;                   /* unary prefix + k suffix bits */
    } else {
      int len = k;
      rem -= 3 << k;
      while (rem >= (1 << len)) { rem -= 1 << len; len++; }     /* Exp-Golomb escape */
      rate += (3 + len + 1 - k + len) * ONE_BIT;
    }
    if (g1_coded) {
      rate += bin_bits(m->one[one_ctx], 1);
      if (g2_coded) rate += bin_bits(m->abs[abs_ctx], 1);
    }
  } else if (level == 1) {
    rate += bin_bits(m->one[one_ctx], 0);
  } else if (level == 2) {
    rate += bin_bits(m->one[one_ctx], 1);
    rate += bin_bits(m->abs[abs_ctx], 0);
  }
  return rate;
}

/* rdo.c:297-339: the better of {max_level, max_level - 1, (0)} for one position.
 * scaled: |coef| * quant scale; is_last: the position opening the scan (no significance bin). */
static uint32_t choose_level(const chan_models *m, double lambda, const level_state *s, int32_t scaled,
                             uint32_t max_level, int sig_ctx, int q_bits, double err_scale, int is_last,
                             double cost_zero, double *cost, double *cost_sig)
{
  uint32_t best = 0;
  double sig1 = 0;
  if (!is_last && max_level < 3) {
    *cost_sig = lambda * bin_bits(m->sig[sig_ctx], 0);
    *cost = cost_zero + *cost_sig;
    if (max_level == 0) return 0;
  } else {
    *cost = BIG_COST;
  }
  if (!is_last) sig1 = lambda * bin_bits(m->sig[sig_ctx], 1);
  const int32_t lowest = max_level > 1 ? (int32_t)max_level - 1 : 1;
  for (int32_t lv = (int32_t)max_level; lv >= lowest; lv--) {
    const double err = (double)(scaled - lv * (1 << q_bits));
    double cand = err * err * err_scale + lambda * level_rate(m, (uint32_t)lv, s);
    cand += sig1;
    if (cand < *cost) { best = (uint32_t)lv; *cost = cand; *cost_sig = sig1; }
  }
  return best;
}

/* context.c:303-339: are the groups right of / below (gx, gy) significant? bit 0 right, bit 1 below */
static int neighbour_groups(const uint8_t *group_sig, int gx, int gy, int groups_per_side)
{
  int r = 0, b = 0;
  if (gx < groups_per_side - 1) r = group_sig[gy * groups_per_side + gx + 1] != 0;
  if (gy < groups_per_side - 1) b = group_sig[(gy + 1) * groups_per_side + gx] != 0;
  return r | (b << 1);
}

/* context.c:354-385 */
static int sig_ctx_inc(int pattern, int scan_idx, int x, int y, int log2_size, int is_chroma)
{
  static const uint8_t map4x4[16] = { 0, 1, 4, 5, 2, 3, 4, 5, 6, 6, 8, 8, 7, 7, 8, 8 };
  if (x + y == 0) return 0;
  if (log2_size == 2) return map4x4[4 * y + x];
  const int base = log2_size == 3 ? (scan_idx == 0 ? 9 : 15) : (is_chroma ? 12 : 21);
  const int xs = x & 3, ys = y & 3;
  int cnt;
  switch (pattern) {
    case 0: cnt = xs + ys == 0 ? 2 : (xs + ys <= 2 ? 1 : 0); break;
    case 1: cnt = ys == 0 ? 2 : (ys == 1 ? 1 : 0); break;
    case 2: cnt = xs == 0 ? 2 : (xs == 1 ? 1 : 0); break;
    default: cnt = 2; break;
  }
  return ((!is_chroma && ((x >> 2) + (y >> 2)) > 0) ? 3 : 0) + base + cnt;
}

/* what the sign-hiding pass needs to know about a position (struct sh_rates_t, rdo.c:131-144) */
typedef struct {
  int32_t inc, dec, sig_inc, delta;
} hide_rec;

/* rdo.c:405-540: where a group's parity disagrees with the sign to hide, move one level by one */
static void hide_signs(double lambda, int bitdepth, int qp_scaled, const uint32_t *scan, const hide_rec *hr,
                       int coded_end, const int16_t *coef, int16_t *lv)
{
  const int iq = k_inv_quant_scale[qp_scaled % 6];
  const int64_t rd_factor =
      (int64_t)(iq * iq * (1 << (2 * (qp_scaled / 6))) / lambda / 16 / (1 << (2 * (bitdepth - 8))) + 0.5);
  const int top_group = (coded_end - 1) >> 4;
  for (int g = top_group; g >= 0; g--) {
    const uint32_t *sc = scan + 16 * g;
    int first = 16, last = -1;
    for (int n = 15; n >= 0; n--) if (lv[sc[n]]) { last = n; break; }
    for (int n = 0; n <= last; n++) if (lv[sc[n]]) { first = n; break; }
    if (last - first < 4) continue;                                   /* SBH_THRESHOLD */
    const int sign = lv[sc[first]] <= 0;
    unsigned sum = 0;
    for (int n = first; n <= last; n++) sum += (unsigned)lv[sc[n]];
    if (sign == (int)(sum & 1)) continue;

    int64_t best_cost = INT64_MAX;
    int best_pos = 0, best_change = 0;
    for (int n = (g == top_group ? last : 15); n >= 0; n--) {
      const int pos = (int)sc[n];
      const int64_t dist = rd_factor * hr[pos].delta;
      const int mag = abs(lv[pos]);
      int64_t cost;
      int change;
      if (mag != 0) {
        int64_t up = hr[pos].inc, down = hr[pos].dec;
        if (mag == 1) down -= ONE_BIT + hr[pos].sig_inc;              /* the sign and significance go away */
        if (g == top_group && n == last && mag == 1) down -= 4 * ONE_BIT;
        up = -dist + up;                                              /* PRECISION_INC == 0 */
        down = dist + down;
        if (up < down) { change = 1; cost = up; }
        else {
          change = -1; cost = down;
          if (n == first && mag == 1) cost = INT64_MAX;
        }
      } else {
        const int bits = ONE_BIT + hr[pos].inc + hr[pos].sig_inc;
        cost = -llabs(dist) + bits;
        change = 1;
        if (n < first && (coef[pos] >= 0 ? 0 : 1) != sign) cost = INT64_MAX;
      }
      if (cost < best_cost) { best_cost = cost; best_pos = pos; best_change = change; }
    }
    if (lv[best_pos] == 32767 || lv[best_pos] == -32768) best_change = -1;
    if (coef[best_pos] >= 0) lv[best_pos] = (int16_t)(lv[best_pos] + best_change);
    else lv[best_pos] = (int16_t)(lv[best_pos] - best_change);
  }
}

/* rdo.c:366-395: cumulative bits of the truncated-unary last-position prefix, per prefix group */
static void last_prefix_bits(const uint8_t *models, int log2_size, int is_chroma, int32_t *bits)
{
  const int cb = log2_size - 2;
  const int offset = is_chroma ? 0 : cb * 3 + ((cb + 1) >> 2);
  const int shift = is_chroma ? cb : (cb + 3) >> 2;
  const int groups = k_last_group[(1 << log2_size) - 1];
  int32_t run = 0;
  for (int g = 0; g < groups; g++) {
    const uint8_t st = models[offset + (g >> shift)];
    bits[g] = run + bin_bits(st, 0);
    run += bin_bits(st, 1);
  }
  bits[groups] = run;
}

/* rdo.c:350-364 */
static double last_pos_cost(double lambda, int a, int b, const int32_t *bits_x, const int32_t *bits_y)
{
  const unsigned ga = k_last_group[a], gb = k_last_group[b];
  double bits = bits_x[ga] + bits_y[gb];
  if (ga > 3) bits += ONE_BIT * ((ga - 2) >> 1);
  if (gb > 3) bits += ONE_BIT * ((gb - 2) >> 1);
  return lambda * bits;
}

void kvzo_rdoq(const kvzo_q_cfg *c, double lambda, const kvzo_rdoq_ctx *ctx, const int16_t *coef,
               int16_t *dest, int width, int type, int scan_idx, int tr_depth)
{
  int lg = 0;
  while ((1 << lg) < width) lg++;
  const int nn = width * width, gside = width >> 2;
  const int is_chroma = type != 0;
  const int transform_shift = 15 - c->bitdepth - lg;
  const int qp_scaled = kvzo_scaled_qp(type, c->qp, (c->bitdepth - 8) * 6);
  const int q_bits = 14 + qp_scaled / 6 + transform_shift;
  /* flat scaling list: one quantiser scale and one error scale for the block
   * (scalinglist.c:339-356, 374) */
  const int32_t q = k_quant_scale[qp_scaled % 6];
  double err_scale = (double)ONE_BIT;
  for (int i = 0; i < 2 * transform_shift; i++) err_scale *= 0.5;      /* * 2^(-2 * transform_shift), exact */
  for (int i = 0; i > 2 * transform_shift; i--) err_scale *= 2.0;
  err_scale = err_scale / q / q / (1 << (2 * (c->bitdepth - 8)));
  const int32_t round = 1 << (q_bits - 1);

  const chan_models m = {
    is_chroma ? ctx->sig_chroma : ctx->sig_luma, is_chroma ? ctx->one_chroma : ctx->one_luma,
    is_chroma ? ctx->abs_chroma : ctx->abs_luma, is_chroma ? ctx->last_x_chroma : ctx->last_x_luma,
    is_chroma ? ctx->last_y_chroma : ctx->last_y_luma };
  const uint8_t *group_models = ctx->sig_group + type;                /* &cu_sig_coeff_group_model[type], rdo.c:615 */

  uint32_t scan[32 * 32];
  kvzo_scan_order(scan_idx, lg, scan);

  static __thread double cost[32 * 32], cost_sig[32 * 32], cost_zero[32 * 32];   /* by scan position */
  static __thread hide_rec hr[32 * 32];                                          /* by block position */
  double group_sig_cost[64];
  uint8_t group_sig[64];
  memset(group_sig, 0, sizeof(group_sig));

  /* ---- where does the scan start? (rdo.c:632-661) ---- */
  int last_pos = -1;
  for (int sp = nn - 1; sp >= 0; sp--) {
    const uint32_t bp = scan[sp];
    int32_t scaled = abs(coef[bp]) * q;
    if (scaled > INT_MAX - round) scaled = INT_MAX - round;
    if (((uint32_t)(scaled + round) >> q_bits) > 0) { last_pos = sp; hr[bp].sig_inc = 0; break; }
    dest[bp] = 0;
  }
  if (last_pos < 0) return;
  const int last_group = last_pos >> 4;
  level_state st = { (last_pos > 0 && !is_chroma) ? 2 : 0, 1, 0, 0, 0, 0 };
  for (int g = last_group; g >= 0; g--) group_sig_cost[g] = 0;

  int32_t bits_x[32], bits_y[32];
  last_prefix_bits(m.last_x, lg, is_chroma, bits_x);
  last_prefix_bits(m.last_y, lg, is_chroma, bits_y);

  double base_cost = 0, uncoded_cost = 0;

  /* ---- phase 1+2: levels of each group, then keep or zero the group (rdo.c:671-836) ---- */
  for (int g = last_group; g >= 0; g--) {
    const uint32_t first_bp = scan[16 * g];
    const int gx = (int)(first_bp & (width - 1)) >> 2, gy = (int)(first_bp >> lg) >> 2;
    const int gpos = gy * gside + gx;
    const int pattern = neighbour_groups(group_sig, gx, gy, gside);
    double s_coded = 0, s_uncoded = 0, s_sig = 0, s_sig0 = 0;
    int nnz_above0 = 0;

    for (int k = 15; k >= 0; k--) {
      const int sp = 16 * g + k;
      if (sp > last_pos) continue;
      const uint32_t bp = scan[sp];
      int32_t scaled = abs(coef[bp]) * q;
      if (scaled > INT_MAX - round) scaled = INT_MAX - round;
      const uint32_t max_level = (uint32_t)(scaled + round) >> q_bits;
      const double e0 = (double)scaled;
      cost_zero[sp] = e0 * e0 * err_scale;
      uncoded_cost += cost_zero[sp];

      uint32_t level;
      int sctx = 0;
      if (sp == last_pos) {
        level = choose_level(&m, lambda, &st, scaled, max_level, 0, q_bits, err_scale, 1, cost_zero[sp],
                             &cost[sp], &cost_sig[sp]);
      } else {
        const int y = (int)(bp >> lg), x = (int)(bp & (width - 1));
        sctx = sig_ctx_inc(pattern, scan_idx, x, y, lg, is_chroma);
        level = choose_level(&m, lambda, &st, scaled, max_level, sctx, q_bits, err_scale, 0, cost_zero[sp],
                             &cost[sp], &cost_sig[sp]);
        if (c->signhide) hr[bp].sig_inc = bin_bits(m.sig[sctx], 1) - bin_bits(m.sig[sctx], 0);
      }
      if (c->signhide) {
        hr[bp].delta = (scaled - (int32_t)level * (1 << q_bits)) >> (q_bits - 8);
        if (level > 0) {
          const int32_t now = level_rate(&m, level, &st);
          hr[bp].inc = level_rate(&m, level + 1, &st) - now;
          hr[bp].dec = level_rate(&m, level - 1, &st) - now;
        } else {
          hr[bp].inc = bin_bits(m.one[4 * st.ctx_set + st.c1], 0);
        }
      }
      dest[bp] = (int16_t)level;
      base_cost += cost[sp];

      /* Rice / greater-than-1 / greater-than-2 bookkeeping (rdo.c:731-761) */
      const uint32_t base = st.c1_idx < 8 ? (st.c2_idx < 1 ? 3u : 2u) : 1u;
      if (level >= base && level > (3u << st.rice)) st.rice = st.rice + 1 < 4 ? st.rice + 1 : 4;
      if (level >= 1) st.c1_idx++;
      if (level > 1) { st.c1 = 0; st.c2 += st.c2 < 2; st.c2_idx++; }
      else if (st.c1 < 3 && st.c1 > 0 && level) st.c1++;
      if (k == 0 && sp > 0) {
        st.c2 = 0; st.rice = 0; st.c1_idx = 0; st.c2_idx = 0;
        st.ctx_set = (sp == 16 || is_chroma) ? 0 : 2;
        if (st.c1 == 0) st.ctx_set++;
        st.c1 = 1;
      }

      s_sig += cost_sig[sp];
      if (k == 0) s_sig0 = cost_sig[sp];
      if (dest[bp]) {
        group_sig[gpos] = 1;
        s_coded += cost[sp] - cost_sig[sp];
        s_uncoded += cost_zero[sp];
        if (k != 0) nnz_above0++;
      }
    }

    if (g == 0) { group_sig[gpos] = 1; continue; }
    const int gctx = neighbour_groups(group_sig, gx, gy, gside) != 0;
    if (!group_sig[gpos]) {
      group_sig_cost[g] = lambda * bin_bits(group_models[gctx], 0);
      base_cost += group_sig_cost[g] - s_sig;
    } else if (g < last_group) {
      if (nnz_above0 == 0) { base_cost -= s_sig0; s_sig -= s_sig0; }
      double zeroed = base_cost;
      group_sig_cost[g] = lambda * bin_bits(group_models[gctx], 1);
      base_cost += group_sig_cost[g];
      zeroed += lambda * bin_bits(group_models[gctx], 0);
      zeroed += s_uncoded;
      zeroed -= s_coded;
      zeroed -= s_sig;
      if (zeroed < base_cost) {
        group_sig[gpos] = 0;
        base_cost = zeroed;
        group_sig_cost[g] = lambda * bin_bits(group_models[gctx], 0);
        for (int k = 15; k >= 0; k--) {
          const int sp = 16 * g + k;
          if (dest[scan[sp]]) { dest[scan[sp]] = 0; cost[sp] = cost_zero[sp]; cost_sig[sp] = 0; }
        }
      }
    }
  }

  /* ---- phase 3: the cheapest place to end the coded part (rdo.c:838-900) ---- */
  double best_cost;
  if (!c->cu_is_intra && !is_chroma) {
    best_cost = uncoded_cost + lambda * bin_bits(ctx->root_cbf, 0);
    base_cost += lambda * bin_bits(ctx->root_cbf, 1);
  } else {
    const uint8_t cbf = is_chroma ? ctx->cbf_chroma[tr_depth] : ctx->cbf_luma[!tr_depth];
    best_cost = uncoded_cost + lambda * bin_bits(cbf, 0);
    base_cost += lambda * bin_bits(cbf, 1);
  }
  int coded_end = 0, stop = 0;
  for (int g = last_group; g >= 0 && !stop; g--) {
    const uint32_t first_bp = scan[16 * g];
    const int gpos = ((int)(first_bp >> lg) >> 2) * gside + ((int)(first_bp & (width - 1)) >> 2);
    base_cost -= group_sig_cost[g];
    if (!group_sig[gpos]) continue;
    for (int k = 15; k >= 0; k--) {
      const int sp = 16 * g + k;
      if (sp > last_pos) continue;
      const uint32_t bp = scan[sp];
      if (!dest[bp]) { base_cost -= cost_sig[sp]; continue; }
      const int y = (int)(bp >> lg), x = (int)(bp & (width - 1));
      const double lc = scan_idx == 2 ? last_pos_cost(lambda, y, x, bits_x, bits_y)
                                      : last_pos_cost(lambda, x, y, bits_x, bits_y);
      const double total = base_cost + lc - cost_sig[sp];
      if (total < best_cost) { coded_end = sp + 1; best_cost = total; }
      if (dest[bp] > 1) { stop = 1; break; }
      base_cost -= cost[sp];
      base_cost += cost_zero[sp];
    }
  }

  uint32_t abs_sum = 0;
  for (int sp = 0; sp < coded_end; sp++) {
    const uint32_t bp = scan[sp];
    const int level = dest[bp];
    abs_sum += (uint32_t)level;
    dest[bp] = (int16_t)(coef[bp] < 0 ? -level : level);
  }
  for (int sp = coded_end; sp <= last_pos; sp++) dest[scan[sp]] = 0;
  if (c->signhide && abs_sum >= 2)
    hide_signs(lambda, c->bitdepth, qp_scaled, scan, hr, coded_end, coef, dest);
}

/* ======================================================================================
 * kvz_get_coeff_cost's CABAC path (rdo.c:158-197): the bits kvz_encode_coeff_nxn
 * (encode_coding_tree.c:103-345) spends on a block when run on a copy of the CABAC state in counting mode.
 * In that mode kvz_cabac_write only counts bytes (cabac.c:128-137), so
 *   (23 - bits_left) + 8 * num_buffered_bytes  ==  the total of all renormalisation shifts,
 * a function of the coder's `range` and the context states only (`low` never matters).
 * ====================================================================================== */

/* HEVC table 9-46 (rangeTabLps), cabac.c:52-67 */
static const uint8_t k_range_lps[64][4] = {
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
/* HEVC table 9-47 (transIdxLps); the MPS transition is state + 1 up to 62 */
static const uint8_t k_next_lps[64] = { 0, 0, 1, 2, 2, 4, 4, 5, 6, 7, 8, 9, 9, 11, 11, 12, 13, 13, 15, 15, 16, 16, 18, 18,
  19, 19, 21, 21, 22, 22, 23, 24, 24, 25, 26, 26, 27, 27, 28, 29, 29, 30, 30, 30, 31, 32, 32, 33, 33, 33, 34, 34, 35,
  35, 35, 36, 36, 36, 37, 37, 37, 38, 38, 63 };
static const uint8_t k_min_in_group[10] = { 0, 1, 2, 3, 4, 6, 8, 12, 16, 24 };      /* encoderstate.h:351 */

typedef struct { uint32_t range; uint32_t bits; } bit_counter;

static int log2_of(int n) { int l = 0; while ((1 << l) < n) l++; return l; }

/* kvz_cabac_encode_bin (cabac.c:91-120) without `low`; state = (pStateIdx << 1) | valMps as in cabac_ctx_t */
static void count_bin(bit_counter *b, uint8_t *state, int bin)
{
  const uint32_t lps = k_range_lps[*state >> 1][(b->range >> 6) & 3];
  b->range -= lps;
  if ((bin ? 1 : 0) != (*state & 1)) {
    /* kvz_g_auc_renorm_table[lps >> 3] (cabac.c:69-73): the shift that brings lps back to >= 256 (state 63's
     * lps of 2 is only ever used by the terminating bin, its table entry is 6 like any lps < 8) */
    static const uint8_t renorm[32] = { 6, 5, 4, 4, 3, 3, 3, 3, 2, 2, 2, 2, 2, 2, 2, 2,
                                        1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1 };
    const int shift = renorm[lps >> 3];
    b->range = lps << shift;
    b->bits += shift;
    const int st = *state >> 1;
    *state = st == 0 ? (uint8_t)(*state ^ 1) : (uint8_t)((k_next_lps[st] << 1) | (*state & 1));
  } else {
    if ((*state >> 1) < 62) *state += 2;
    if (b->range < 256) { b->range <<= 1; b->bits += 1; }
  }
}

static void count_remaining(bit_counter *b, uint32_t symbol, uint32_t rice)
{
  /* kvz_cabac_write_coeff_remain (cabac.c:317-337): bypass bins only */
  if (symbol < (3u << rice)) {
    b->bits += (symbol >> rice) + 1 + rice;
  } else {
    uint32_t length = rice, code = symbol - (3u << rice);
    while (code >= (1u << length)) { code -= 1u << length; length++; }
    b->bits += (3 + length + 1 - rice) + length;
  }
}

uint32_t kvzo_coeff_cabac_bits(const kvzo_rdoq_ctx *ctx_in, uint32_t range, const int16_t *coeff, int width, int type,
                               int scan_mode, int signhide, int trskip_enable)
{
  const int nn = width * width, lg = log2_of(width), side = width >> 2, is_chroma = type != 0;
  int any = 0;
  for (int i = 0; i < nn; i++) any |= coeff[i] != 0;
  if (!any) return 0;                                         /* rdo.c:166-173 */

  kvzo_rdoq_ctx ctx = *ctx_in;                                /* the coder updates its copy of the contexts */
  bit_counter b = { range, 0 };
  uint32_t scan[32 * 32], scan_cg[64];
  kvzo_scan_order(scan_mode, lg, scan);
  for (int g = 0; g < side * side; g++) {                     /* group order = where each 16-run of the scan lies */
    const uint32_t bp = scan[g * 16];
    scan_cg[g] = ((bp >> lg) >> 2) * side + ((bp & (width - 1)) >> 2);
  }
  uint8_t group_sig[64];
  memset(group_sig, 0, sizeof(group_sig));
  for (int i = 0; i < nn; i++)
    if (coeff[i]) group_sig[((i >> lg) >> 2) * side + ((i & (width - 1)) >> 2)] = 1;

  int cg_last = side * side - 1;
  while (!group_sig[scan_cg[cg_last]]) cg_last--;
  int pos_last_scan = cg_last * 16 + 15;
  while (!coeff[scan[pos_last_scan]]) pos_last_scan--;
  const int pos_last = scan[pos_last_scan];

  if (width == 4 && trskip_enable) count_bin(&b, is_chroma ? &ctx.trskip_chroma : &ctx.trskip_luma, 0);

  {
    /* encode_last_significant_xy (encode_coding_tree.c:49-101) */
    int lx = pos_last & (width - 1), ly = pos_last >> lg;
    const int index = lg - 2;
    const int ctx_offset = is_chroma ? 0 : index * 3 + (index + 1) / 4;
    const int shift = is_chroma ? index : (index + 3) / 4;
    uint8_t *mx = is_chroma ? ctx.last_x_chroma : ctx.last_x_luma, *my = is_chroma ? ctx.last_y_chroma : ctx.last_y_luma;
    if (scan_mode == 2) { const int t = lx; lx = ly; ly = t; }
    const int gx = k_last_group[lx], gy = k_last_group[ly], gmax = k_last_group[width - 1];
    for (int i = 0; i < gx; i++) count_bin(&b, &mx[ctx_offset + (i >> shift)], 1);
    if (gx < gmax) count_bin(&b, &mx[ctx_offset + (gx >> shift)], 0);
    for (int i = 0; i < gy; i++) count_bin(&b, &my[ctx_offset + (i >> shift)], 1);
    if (gy < gmax) count_bin(&b, &my[ctx_offset + (gy >> shift)], 0);
    if (gx > 3) b.bits += (gx - 2) / 2;
    if (gy > 3) b.bits += (gy - 2) / 2;
    (void)k_min_in_group;                                     /* the suffix VALUE does not change the bit count */
  }

  uint8_t *sig_models = is_chroma ? ctx.sig_chroma : ctx.sig_luma;
  int c1 = 1, scan_pos = pos_last_scan;
  for (int g = cg_last; g >= 0; g--) {
    const int sub_pos = g << 4, cg_blk = scan_cg[g], cgy = cg_blk / side, cgx = cg_blk - cgy * side;
    int abs_coeff[16], n_nz = 0, last_nz = -1, first_nz = 16;
    uint32_t rice = 0;
    if (scan_pos == pos_last_scan) {
      abs_coeff[0] = abs(coeff[pos_last]);
      n_nz = 1; last_nz = first_nz = scan_pos;
      scan_pos--;
    }
    if (g == cg_last || g == 0) {
      group_sig[cg_blk] = 1;
    } else {
      const int nb = neighbour_groups(group_sig, cgx, cgy, side);
      count_bin(&b, &ctx.sig_group[type + (nb != 0)], group_sig[cg_blk] != 0);   /* &cu_sig_coeff_group_model[type] */
    }
    if (group_sig[cg_blk]) {
      const int pattern = width == 4 ? -1 : neighbour_groups(group_sig, cgx, cgy, side);
      for (; scan_pos >= sub_pos; scan_pos--) {
        const uint32_t bp = scan[scan_pos];
        const int py = bp >> lg, px = bp - (py << lg), sig = coeff[bp] != 0;
        if (scan_pos > sub_pos || g == 0 || n_nz)
          count_bin(&b, &sig_models[sig_ctx_inc(pattern, scan_mode, px, py, lg, is_chroma)], sig);
        if (sig) {
          abs_coeff[n_nz++] = abs(coeff[bp]);
          if (last_nz == -1) last_nz = scan_pos;
          first_nz = scan_pos;
        }
      }
    } else {
      scan_pos = sub_pos - 1;
    }
    if (n_nz > 0) {
      const int sign_hidden = last_nz - first_nz >= 4;
      int ctx_set = (g > 0 && !is_chroma) ? 2 : 0;
      if (c1 == 0) ctx_set++;
      c1 = 1;
      uint8_t *one = (is_chroma ? ctx.one_chroma : ctx.one_luma) + 4 * ctx_set;
      const int n_c1 = n_nz < 8 ? n_nz : 8;                   /* C1FLAG_NUMBER */
      int first_c2 = -1;
      for (int i = 0; i < n_c1; i++) {
        const int sym = abs_coeff[i] > 1;
        count_bin(&b, &one[c1], sym);
        if (sym) { c1 = 0; if (first_c2 == -1) first_c2 = i; }
        else if (c1 < 3 && c1 > 0) c1++;
      }
      if (c1 == 0 && first_c2 != -1)
        count_bin(&b, &(is_chroma ? ctx.abs_chroma : ctx.abs_luma)[ctx_set], abs_coeff[first_c2] > 2);
      b.bits += (signhide && sign_hidden) ? n_nz - 1 : n_nz;  /* coeff_sign_flag, bypass */
      if (c1 == 0 || n_nz > 8) {
        int first_coeff2 = 1;
        for (int i = 0; i < n_nz; i++) {
          const int base = i < 8 ? 2 + first_coeff2 : 1;
          if (abs_coeff[i] >= base) {
            count_remaining(&b, abs_coeff[i] - base, rice);
            if (abs_coeff[i] > 3 * (1 << rice)) rice = rice + 1 < 4 ? rice + 1 : 4;
          }
          if (abs_coeff[i] >= 2) first_coeff2 = 0;
        }
      }
    }
  }
  return b.bits;
}
