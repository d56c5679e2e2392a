/*
 * oracle/kvz_oracle_cand.c -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * CPU restatement of the reference's inter candidate derivation:
 *   kvz_inter_get_merge_cand (inter.c:1314-1446), kvz_inter_get_mv_cand (inter.c:1209-1228) with
 *   get_mv_cand_from_candidates (:1102-1194), the spatial (:799-878) and temporal (:713-797) candidate
 *   positions, is_a0_cand_coded / is_b0_cand_coded (:566-697), MV scaling (:955-1009), add_temporal_candidate
 *   (:1022-1070), and the colocated start vector of search_pu_inter_ref (search_inter.c:1188-1206),
 * on a frame-wide array of SCU records instead of the working lcu_t (the positions the reference reads through
 * LCU_GET_CU_AT_PX / LCU_GET_TOP_RIGHT_CU are the same picture positions).
 * Pinned on the compiled reference through oracle/ref_shim.c (tests/test_oracle_cand.py).
 */
#include <stdlib.h>
#include <string.h>
#include "kvz_oracle.h"

#define KCLIP(lo, hi, v) ((v) < (lo) ? (lo) : ((v) > (hi) ? (hi) : (v)))

typedef struct { const kvzo_scu *a[2], *b[3], *c3, *h; } cand_set;

static const kvzo_scu *scu_at(const kvzo_cand_frame *f, int x, int y)
{
  return &f->scus[(y >> 2) * (f->width >> 2) + (x >> 2)];
}

static int a0_coded(int x, int y, int width, int height)
{
  /* inter.c:566-627 */
  int size = (width & -width) < (height & -height) ? (width & -width) : (height & -height);
  if (height != size) y = y + height - size;
  while (size < 64) {
    const int parent = 2 * size;
    switch ((x % parent != 0) + 2 * (y % parent != 0)) {
      case 0: return 1;
      case 1: return 0;
      case 2: y -= size; size = parent; break;
      default: return 0;
    }
  }
  return 0;
}

static int b0_coded(int x, int y, int width, int height)
{
  /* inter.c:638-697 */
  int size = (width & -width) < (height & -height) ? (width & -width) : (height & -height);
  if (width != size) x = x + width - size;
  while (size < 64) {
    const int parent = 2 * size;
    switch ((x % parent != 0) + 2 * (y % parent != 0)) {
      case 0: return 1;
      case 1: x -= size; size = parent; break;
      case 2: return 1;
      default: return 0;
    }
  }
  return 1;
}

static void spatial(const kvzo_cand_frame *f, int x, int y, int w, int h, cand_set *c)
{
  /* inter.c:799-878 (and its cu_array twin :896-953) */
  const int xl = x & 63, yl = y & 63;
  if (x != 0) {
    const kvzo_scu *a1 = scu_at(f, x - 1, y + h - 1);
    if (a1->type == 2) c->a[1] = a1;
    if (yl + h < 64 && y + h < f->height) {
      const kvzo_scu *a0 = scu_at(f, x - 1, y + h);
      if (a0->type == 2 && a0_coded(x, y, w, h)) c->a[0] = a0;
    }
  }
  if (y != 0) {
    if (x + w < f->width && (xl + w < 64 || yl == 0)) {
      const kvzo_scu *b0 = scu_at(f, x + w, y - 1);
      if (b0->type == 2 && b0_coded(x, y, w, h)) c->b[0] = b0;
    }
    const kvzo_scu *b1 = scu_at(f, x + w - 1, y - 1);
    if (b1->type == 2) c->b[1] = b1;
    if (x != 0) {
      const kvzo_scu *b2 = scu_at(f, x - 1, y - 1);
      if (b2->type == 2) c->b[2] = b2;
    }
  }
}

static void temporal(const kvzo_cand_frame *f, int x, int y, int w, int h, cand_set *c)
{
  /* inter.c:713-797, always asked for list L0, index 0 */
  c->c3 = c->h = NULL;
  if (!f->n_refs || f->ref_LX_size[0] == 0) return;
  const kvzo_scu *col = f->ref_scus[f->ref_LX[0][0]];
  const int per_row = f->ref_scu_pitch;
  const unsigned xbr = x + w, ybr = y + h;
  if (xbr < (unsigned)f->in_width && ybr < (unsigned)f->in_height && ybr % 64 != 0) {
    const kvzo_scu *s = &col[((xbr >> 4) << 4) / 4 + (((ybr >> 4) << 4) / 4) * per_row];
    if (s->type == 2) c->h = s;
  }
  const unsigned xc = x + w / 2, yc = y + h / 2;
  if (xc < (unsigned)f->in_width && yc < (unsigned)f->in_height) {
    const kvzo_scu *s = &col[((xc >> 4) << 4) / 4 + (((yc >> 4) << 4) / 4) * per_row];
    if (s->type == 2) c->c3 = s;
  }
}

static int16_t scaled_mv(int16_t mv, int scale)
{
  const int32_t s = scale * mv;
  return (int16_t)KCLIP(-32768, 32767, (s + 127 + (s < 0)) >> 8);
}

static void scale_by_pocs(int cur_poc, int cur_ref_poc, int nb_poc, int nb_ref_poc, int16_t mv[2])
{
  /* inter.c:961-980 */
  int dc = cur_poc - cur_ref_poc, dn = nb_poc - nb_ref_poc;
  if (dc == dn) return;
  dc = KCLIP(-128, 127, dc); dn = KCLIP(-128, 127, dn);
  const int scale = KCLIP(-4096, 4095, (dc * ((0x4000 + (abs(dn) >> 1)) / dn) + 32) >> 6);
  mv[0] = scaled_mv(mv[0], scale);
  mv[1] = scaled_mv(mv[1], scale);
}

static int temporal_mv(const kvzo_cand_frame *f, int current_ref, const kvzo_scu *col, int reflist, int16_t out[2])
{
  /* inter.c:1022-1070 */
  if (!col || f->ref_LX_size[0] == 0) return 0;
  const int col_ref = f->ref_LX[0][0];
  int col_list = reflist;
  for (int i = 0; i < f->n_refs; i++)
    if (f->ref_pocs[i] > f->poc) { col_list = 1; break; }
  if ((col->mv_dir & (col_list + 1)) == 0) col_list = 1 - col_list;
  out[0] = col->mv[col_list][0]; out[1] = col->mv[col_list][1];
  scale_by_pocs(f->poc, f->ref_pocs[current_ref], f->ref_pocs[col_ref], f->col_ref_pocs[col_list][col->mv_ref[col_list] & 15], out);
  return 1;
}

static int mvp(const kvzo_cand_frame *f, int cur_idx, const kvzo_scu *cand, int reflist, int scaling, int16_t out[2])
{
  /* inter.c:1072-1108 */
  if (!cand) return 0;
  for (int i = 0; i < 2; i++) {
    const int cl = i == 0 ? reflist : !reflist;
    if ((cand->mv_dir & (1 << cl)) == 0) continue;
    if (scaling) {
      out[0] = cand->mv[cl][0]; out[1] = cand->mv[cl][1];
      scale_by_pocs(f->poc, f->ref_pocs[f->ref_LX[reflist][cur_idx]], f->poc, f->ref_pocs[f->ref_LX[cl][cand->mv_ref[cl] & 15]], out);
      return 1;
    }
    if (f->ref_LX[cl][cand->mv_ref[cl] & 15] == f->ref_LX[reflist][cur_idx]) {
      out[0] = cand->mv[cl][0]; out[1] = cand->mv[cl][1];
      return 1;
    }
  }
  return 0;
}

void kvzo_inter_mv_cand(const kvzo_cand_frame *f, int x, int y, int w, int h, int reflist, int lx_idx, int16_t mv_cand[2][2])
{
  /* kvz_inter_get_mv_cand (inter.c:1209-1228) + get_mv_cand_from_candidates (:1102-1194) */
  cand_set c;
  memset(&c, 0, sizeof(c));
  spatial(f, x, y, w, h, &c);
  temporal(f, x, y, w, h, &c);
  int n = 0, nb = 0;
  for (int i = 0; i < 2; i++) if (mvp(f, lx_idx, c.a[i], reflist, 0, mv_cand[n])) { n++; break; }
  if (n == 0) for (int i = 0; i < 2; i++) if (mvp(f, lx_idx, c.a[i], reflist, 1, mv_cand[n])) { n++; break; }
  for (int i = 0; i < 3; i++) if (mvp(f, lx_idx, c.b[i], reflist, 0, mv_cand[n])) { nb++; break; }
  n += nb;
  if (c.a[0] || c.a[1]) nb = 1; else if (n != 2) nb = 0;
  if (!nb) for (int i = 0; i < 3; i++) if (mvp(f, lx_idx, c.b[i], reflist, 1, mv_cand[n])) { n++; break; }
  if (n == 2 && mv_cand[0][0] == mv_cand[1][0] && mv_cand[0][1] == mv_cand[1][1]) n = 1;
  if (f->tmvp_enable && f->poc > 1 && f->n_refs && n < 2 && (c.h || c.c3) &&
      temporal_mv(f, f->ref_LX[reflist][lx_idx], c.h ? c.h : c.c3, reflist, mv_cand[n])) n++;
  while (n < 2) { mv_cand[n][0] = 0; mv_cand[n][1] = 0; n++; }
}

static int same_motion(const kvzo_scu *p, const kvzo_scu *q)
{
  /* inter.c:1262-1279 */
  if (!q) return 0;
  if (p->mv_dir != q->mv_dir) return 0;
  for (int l = 0; l < 2; l++)
    if ((p->mv_dir & (1 << l)) && (p->mv[l][0] != q->mv[l][0] || p->mv[l][1] != q->mv[l][1] || p->mv_ref[l] != q->mv_ref[l])) return 0;
  return 1;
}

static int put_merge(const kvzo_scu *cand, const kvzo_scu *d1, const kvzo_scu *d2, kvzo_merge_cand *out)
{
  /* inter.c:1281-1300; the unused list reads as inter_clear_cu_unused (inter.c:546-555) left it: mv 0, ref 255 */
  if (!cand || same_motion(cand, d1) || same_motion(cand, d2)) return 0;
  for (int l = 0; l < 2; l++) {
    const int used = (cand->mv_dir >> l) & 1;
    out->mv[l][0] = used ? cand->mv[l][0] : 0;
    out->mv[l][1] = used ? cand->mv[l][1] : 0;
    out->ref[l] = used ? cand->mv_ref[l] : 255;
  }
  out->dir = cand->mv_dir;
  return 1;
}

int kvzo_inter_merge_cand(const kvzo_cand_frame *f, int x, int y, int w, int h, int use_a1, int use_b1, kvzo_merge_cand out[5])
{
  /* kvz_inter_get_merge_cand (inter.c:1314-1446) */
  cand_set c;
  memset(&c, 0, sizeof(c));
  memset(out, 0, 5 * sizeof(out[0]));
  spatial(f, x, y, w, h, &c);
  if (!use_a1) c.a[1] = NULL;
  if (!use_b1) c.b[1] = NULL;
  int n = 0;
  if (put_merge(c.a[1], NULL, NULL, &out[n])) n++;
  if (put_merge(c.b[1], c.a[1], NULL, &out[n])) n++;
  if (put_merge(c.b[0], c.b[1], NULL, &out[n])) n++;
  if (put_merge(c.a[0], c.a[1], NULL, &out[n])) n++;
  if (n < 4 && put_merge(c.b[2], c.a[1], c.b[1], &out[n])) n++;

  if (f->tmvp_enable && n < 5 && f->n_refs) {
    out[n].dir = 0;
    for (int l = 0; l <= (f->slice_b ? 1 : 0); l++) {
      temporal(f, x, y, w, h, &c);
      if (temporal_mv(f, f->ref_LX[l][0], c.h ? c.h : c.c3, l, out[n].mv[l])) {
        out[n].ref[l] = 0;
        out[n].dir |= 1 << l;
      }
    }
    if (out[n].dir != 0) n++;
  }

  if (n < 5 && f->slice_b) {
    static const uint8_t p0[12] = { 0, 1, 0, 2, 1, 2, 0, 3, 1, 3, 2, 3 }, p1[12] = { 1, 0, 2, 0, 2, 1, 3, 0, 3, 1, 3, 2 };
    const int cutoff = n;
    for (int idx = 0; idx < cutoff * (cutoff - 1) && n != 5; idx++) {
      const int i = p0[idx], j = p1[idx];
      if (i >= n || j >= n) break;
      if ((out[i].dir & 1) && (out[j].dir & 2)) {
        out[n].dir = 3;
        out[n].mv[0][0] = out[i].mv[0][0]; out[n].mv[0][1] = out[i].mv[0][1];
        out[n].mv[1][0] = out[j].mv[1][0]; out[n].mv[1][1] = out[j].mv[1][1];
        out[n].ref[0] = out[i].ref[0]; out[n].ref[1] = out[j].ref[1];
        if (!(f->ref_LX[0][out[i].ref[0] & 15] == f->ref_LX[1][out[j].ref[1] & 15] &&
              out[i].mv[0][0] == out[j].mv[1][0] && out[i].mv[0][1] == out[j].mv[1][1])) n++;
      }
    }
  }

  int num_ref = f->n_refs;
  if (n < 5 && f->slice_b) {
    int neg = 0, pos = 0;
    for (int j = 0; j < f->n_refs; j++) { if (f->ref_pocs[j] < f->poc) neg++; else pos++; }
    num_ref = neg < pos ? neg : pos;
  }
  int zero_idx = 0;
  while (n != 5) {
    memset(&out[n], 0, sizeof(out[n]));
    out[n].ref[0] = (zero_idx >= num_ref - 1) ? 0 : (uint8_t)zero_idx;
    out[n].ref[1] = out[n].ref[0];
    out[n].dir = f->slice_b ? 3 : 1;
    zero_idx++;
    n++;
  }
  return n;
}

void kvzo_colocated_start_mv(const kvzo_cand_frame *f, int x, int y, int w, int h, int ref_idx, int16_t mv[2])
{
  /* search_inter.c:1188-1206 */
  mv[0] = mv[1] = 0;
  const int mx = f->tile_offset_x + x + (w >> 1), my = f->tile_offset_y + y + (h >> 1);
  const kvzo_scu *s = &f->ref_scus[ref_idx][(my >> 2) * f->ref_scu_pitch + (mx >> 2)];
  if (s->type == 2) {
    const int l = (s->mv_dir & 1) ? 0 : 1;
    mv[0] = s->mv[l][0]; mv[1] = s->mv[l][1];
  }
}
