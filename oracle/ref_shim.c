/*
 * oracle/ref_shim.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Compiled INTO oracle/_ref/libkvazaar_ref.so in place of the reference's
 * src/search_inter.c: it #includes that file from where it lies under
 * /root/reference (nothing is copied) so that the file-static search
 * functions (check_mv_cost, search_mv_full, hexagon_search, search_frac,
 * calc_mvd_cost ... search_inter.c:195-1128) become callable from the parity
 * tests through plain-C wrappers.  The wrappers only build the minimal
 * encoder_state_t / encoder_control_t the reference functions read and then
 * call the UNMODIFIED reference code.
 *
 * Also wraps the strategies whose signatures take encoder_state_t*
 * (quant / dequant / quantize_residual, strategies-quant.h:36-48) and the
 * ipol functions (strategies-ipol.h:36-48) behind plain-pointer signatures.
 */
#include "search_inter.c"   /* resolved through -I$(REF)/src */

#include <string.h>
#include "strategyselector.h"
#include "scalinglist.h"
#include "context.h"
#include "transform.h"
#include "intra.h"
#include "strategies/strategies-quant.h"
#include "strategies/strategies-dct.h"
#include "strategies/strategies-intra.h"

/* ---- plain-C job description shared with tests (mirrored in python) ---- */
typedef struct {
  /* frame */
  int32_t width, height;
  const kvz_pixel *pic_y;     /* source luma, stride == width */
  const kvz_pixel *ref_y;     /* reference luma, stride == width */
  /* PU */
  int32_t origin_x, origin_y, pu_w, pu_h;
  int32_t tile_off_x, tile_off_y;
  /* candidates */
  int16_t mv_cand[2][2];
  int32_t num_merge_cand;
  uint8_t merge_dir[5];
  uint8_t merge_ref[5][2];
  int16_t merge_mv[5][2][2];
  uint8_t ref_LX[2][16];
  int32_t ref_idx;
  int32_t extra_mv[2];        /* quarter-pel, as search_pu_inter_ref:1188 */
  /* config */
  double  lambda_sqrt;
  int32_t owf, wpp, sao_type, deblock_enable, mv_constraint;
  int32_t max_ref_lcu_right, max_ref_lcu_down;
  int32_t fme_level, me_early_termination, me_max_steps;
  int32_t ime_algorithm;      /* enum kvz_ime_algorithm */
  /* in/out */
  int32_t best_mv[2];         /* quarter-pel */
  uint32_t best_cost, best_bitcost;
} ref_me_job;

typedef struct {
  encoder_state_t state;
  encoder_control_t ctrl;
  encoder_state_config_frame_t frame;
  encoder_state_config_tile_t tile;
  videoframe_t vframe;
  kvz_picture pic, ref;
} fake_env;

static int g_shim_init_done = 0;

int ref_shim_init(int cpuid)
{
  /* kvazaar.c:81 -- fills the global strategy pointers */
  int ok = kvz_strategyselector_init(cpuid, KVZ_BIT_DEPTH);
  g_shim_init_done = ok;
  return ok;
}

int ref_shim_bitdepth(void) { return KVZ_BIT_DEPTH; }

static void env_from_job(fake_env *e, const ref_me_job *j, inter_search_info_t *info_out)
{
  memset(e, 0, sizeof(*e));
  e->state.encoder_control = &e->ctrl;
  e->state.frame = &e->frame;
  e->state.tile = &e->tile;
  e->state.lambda_sqrt = j->lambda_sqrt;
  e->tile.frame = &e->vframe;
  e->tile.offset_x = j->tile_off_x;
  e->tile.offset_y = j->tile_off_y;
  e->vframe.width = j->width;
  e->vframe.height = j->height;
  e->vframe.source = &e->pic;
  e->pic.y = (kvz_pixel *)j->pic_y; e->pic.width = j->width; e->pic.height = j->height; e->pic.stride = j->width;
  e->ref.y = (kvz_pixel *)j->ref_y; e->ref.width = j->width; e->ref.height = j->height; e->ref.stride = j->width;
  memcpy(e->frame.ref_LX, j->ref_LX, sizeof(e->frame.ref_LX));
  e->ctrl.cfg.owf = j->owf;
  e->ctrl.cfg.wpp = j->wpp;
  e->ctrl.cfg.sao_type = j->sao_type;
  e->ctrl.cfg.deblock_enable = j->deblock_enable;
  e->ctrl.cfg.mv_constraint = j->mv_constraint;
  e->ctrl.cfg.fme_level = j->fme_level;
  e->ctrl.cfg.me_early_termination = j->me_early_termination;
  e->ctrl.cfg.me_max_steps = j->me_max_steps;
  e->ctrl.cfg.ime_algorithm = j->ime_algorithm;
  e->ctrl.cfg.mv_rdo = 0;
  e->ctrl.max_inter_ref_lcu.right = j->max_ref_lcu_right;
  e->ctrl.max_inter_ref_lcu.down = j->max_ref_lcu_down;
  e->ctrl.bitdepth = KVZ_BIT_DEPTH;

  inter_search_info_t info = {
    .state = &e->state,
    .pic = &e->pic,
    .ref = &e->ref,
    .ref_idx = j->ref_idx,
    .origin = { j->origin_x, j->origin_y },
    .width = j->pu_w,
    .height = j->pu_h,
    .mvd_cost_func = calc_mvd_cost,
  };
  memcpy(info.mv_cand, j->mv_cand, sizeof(info.mv_cand));
  info.num_merge_cand = j->num_merge_cand;
  for (int i = 0; i < 5; i++) {
    info.merge_cand[i].dir = j->merge_dir[i];
    info.merge_cand[i].ref[0] = j->merge_ref[i][0];
    info.merge_cand[i].ref[1] = j->merge_ref[i][1];
    memcpy(info.merge_cand[i].mv, j->merge_mv[i], sizeof(info.merge_cand[i].mv));
  }
  memcpy(info_out, &info, sizeof(info));
}

/* Integer motion estimation exactly as search_pu_inter_ref:1208-1239. */
int ref_shim_ime(ref_me_job *j)
{
  fake_env *e = malloc(sizeof(fake_env));
  inter_search_info_t info;
  env_from_job(e, j, &info);
  vector2d_t mv = { j->extra_mv[0], j->extra_mv[1] };
  int search_range = 32;
  switch (j->ime_algorithm) {
    case KVZ_IME_FULL64: search_range = 64; break;
    case KVZ_IME_FULL32: search_range = 32; break;
    case KVZ_IME_FULL16: search_range = 16; break;
    case KVZ_IME_FULL8:  search_range = 8;  break;
    default: break;
  }
  info.best_cost = UINT32_MAX;
  switch (j->ime_algorithm) {
    case KVZ_IME_TZ: tz_search(&info, mv); break;
    case KVZ_IME_FULL64: case KVZ_IME_FULL32: case KVZ_IME_FULL16:
    case KVZ_IME_FULL8: case KVZ_IME_FULL:
      search_mv_full(&info, search_range, mv); break;
    case KVZ_IME_DIA: diamond_search(&info, mv, j->me_max_steps); break;
    default: hexagon_search(&info, mv, j->me_max_steps); break;
  }
  j->best_mv[0] = info.best_mv.x; j->best_mv[1] = info.best_mv.y;
  j->best_cost = info.best_cost; j->best_bitcost = info.best_bitcost;
  free(e);
  return 1;
}

/* Fractional refinement: search_frac (search_inter.c:965) from best_mv. */
int ref_shim_fme(ref_me_job *j)
{
  fake_env *e = malloc(sizeof(fake_env));
  inter_search_info_t info;
  env_from_job(e, j, &info);
  info.best_mv.x = j->best_mv[0]; info.best_mv.y = j->best_mv[1];
  info.best_cost = j->best_cost; info.best_bitcost = j->best_bitcost;
  search_frac(&info);
  j->best_mv[0] = info.best_mv.x; j->best_mv[1] = info.best_mv.y;
  j->best_cost = info.best_cost; j->best_bitcost = info.best_bitcost;
  free(e);
  return 1;
}

/* SATD at the integer MV: the else-branch of search_pu_inter_ref:1244-1256 */
int ref_shim_int_satd(ref_me_job *j)
{
  fake_env *e = malloc(sizeof(fake_env));
  inter_search_info_t info;
  env_from_job(e, j, &info);
  uint32_t c = kvz_image_calc_satd(&e->pic, &e->ref, j->origin_x, j->origin_y,
      j->tile_off_x + j->origin_x + (j->best_mv[0] >> 2),
      j->tile_off_y + j->origin_y + (j->best_mv[1] >> 2), j->pu_w, j->pu_h);
  c += j->best_bitcost * (int)(j->lambda_sqrt + 0.5);
  j->best_cost = c;
  free(e);
  return 1;
}

/* calc_mvd_cost (search_inter.c:373) for one mv. */
uint32_t ref_shim_mvd_cost(ref_me_job *j, int x, int y, int mv_shift, uint32_t *bitcost)
{
  fake_env *e = malloc(sizeof(fake_env));
  inter_search_info_t info;
  env_from_job(e, j, &info);
  uint32_t c = calc_mvd_cost(&e->state, x, y, mv_shift, info.mv_cand, info.merge_cand,
                             info.num_merge_cand, info.ref_idx, bitcost);
  free(e);
  return c;
}

int ref_shim_fracmv_within_tile(ref_me_job *j, int x, int y)
{
  fake_env *e = malloc(sizeof(fake_env));
  inter_search_info_t info;
  env_from_job(e, j, &info);
  int r = fracmv_within_tile(&info, x, y);
  free(e);
  return r;
}

unsigned ref_shim_image_calc_sad(const kvz_pixel *pic, const kvz_pixel *ref, int w, int h,
                                 int pic_x, int pic_y, int ref_x, int ref_y, int bw, int bh)
{
  kvz_picture p = {0}, r = {0};
  p.y = (kvz_pixel *)pic; p.width = w; p.height = h; p.stride = w;
  r.y = (kvz_pixel *)ref; r.width = w; r.height = h; r.stride = w;
  return kvz_image_calc_sad(&p, &r, pic_x, pic_y, ref_x, ref_y, bw, bh);
}

unsigned ref_shim_image_calc_satd(const kvz_pixel *pic, const kvz_pixel *ref, int w, int h,
                                  int pic_x, int pic_y, int ref_x, int ref_y, int bw, int bh)
{
  kvz_picture p = {0}, r = {0};
  p.y = (kvz_pixel *)pic; p.width = w; p.height = h; p.stride = w;
  r.y = (kvz_pixel *)ref; r.width = w; r.height = h; r.stride = w;
  return kvz_image_calc_satd(&p, &r, pic_x, pic_y, ref_x, ref_y, bw, bh);
}

/* ---- ipol wrappers (encoder argument is unused by the strategies) ---- */
/* Runs the FME filter ladder of search_frac:1046-1060 up to and including
 * `step`, with the given half-pel offsets for the qpel steps, on an extended
 * block fetched like search_frac:1008.  Writes filtered[4][64*64]. */
int ref_shim_fme_filter(const kvz_pixel *ref_y, int ref_w, int ref_h,
                        int x, int y, int mvx, int mvy, int w, int h,
                        int step, int fme_level, int hpel_off_x, int hpel_off_y,
                        kvz_pixel *filtered_out /* [4][64*64] */)
{
  encoder_control_t *ctrl = calloc(1, sizeof(*ctrl));
  kvz_extended_block src = { 0, 0, 0, 0 };
  kvz_pixel (*filtered)[LCU_WIDTH * LCU_WIDTH] = (void *)filtered_out;
  int16_t (*intermediate)[(KVZ_EXT_BLOCK_W_LUMA + 1) * LCU_WIDTH] =
      malloc(sizeof(int16_t) * 5 * (KVZ_EXT_BLOCK_W_LUMA + 1) * LCU_WIDTH);
  int16_t hor_first_cols[5][KVZ_EXT_BLOCK_W_LUMA + 1];
  const int iw = ((w + 7) >> 3) << 3, ih = ((h + 7) >> 3) << 3;
  kvz_get_extended_block(x, y, mvx - 1, mvy - 1, 0, 0, (kvz_pixel *)ref_y, ref_w, ref_h,
                         KVZ_LUMA_FILTER_TAPS, iw + 1, ih + 1, &src);
  ipol_blocks_func *steps[4] = {
    kvz_filter_hpel_blocks_hor_ver_luma, kvz_filter_hpel_blocks_diag_luma,
    kvz_filter_qpel_blocks_hor_ver_luma, kvz_filter_qpel_blocks_diag_luma,
  };
  for (int s = 0; s <= step; s++) {
    int ox = s >= 2 ? hpel_off_x : 0, oy = s >= 2 ? hpel_off_y : 0;
    steps[s](ctrl, src.orig_topleft, src.stride, iw, ih, filtered, intermediate,
             fme_level, hor_first_cols, ox, oy);
  }
  if (src.malloc_used) free(src.buffer);
  free(intermediate);
  free(ctrl);
  return 1;
}

/* Final MC sample functions, called like inter.c:43-140 (luma) with an
 * extended block so that edge clamping matches the encoder's use. */
int ref_shim_mc_luma(const kvz_pixel *ref_y, int ref_w, int ref_h, int x, int y,
                     int mvx_q, int mvy_q, int w, int h, kvz_pixel *dst, int dst_stride,
                     int16_t *dst14 /* may be NULL */)
{
  encoder_control_t *ctrl = calloc(1, sizeof(*ctrl));
  kvz_extended_block src = { 0, 0, 0, 0 };
  int16_t mv[2] = { (int16_t)mvx_q, (int16_t)mvy_q };
  kvz_get_extended_block(x, y, mvx_q >> 2, mvy_q >> 2, 0, 0, (kvz_pixel *)ref_y, ref_w, ref_h,
                         KVZ_LUMA_FILTER_TAPS, w, h, &src);
  if (dst14)
    kvz_sample_14bit_quarterpel_luma(ctrl, src.orig_topleft, src.stride, w, h, dst14, dst_stride,
                                     mv[0] & 3, mv[1] & 3, mv);
  else
    kvz_sample_quarterpel_luma(ctrl, src.orig_topleft, src.stride, w, h, dst, dst_stride,
                               mv[0] & 3, mv[1] & 3, mv);
  if (src.malloc_used) free(src.buffer);
  free(ctrl);
  return 1;
}

/* chroma: x,y,w,h in chroma samples, mv in luma quarter-pel (= chroma 1/8) */
int ref_shim_mc_chroma(const kvz_pixel *ref_c, int ref_w, int ref_h, int x, int y,
                       int mvx_q, int mvy_q, int w, int h, kvz_pixel *dst, int dst_stride,
                       int16_t *dst14)
{
  encoder_control_t *ctrl = calloc(1, sizeof(*ctrl));
  kvz_extended_block src = { 0, 0, 0, 0 };
  int16_t mv[2] = { (int16_t)mvx_q, (int16_t)mvy_q };
  kvz_get_extended_block(x, y, mvx_q >> 3, mvy_q >> 3, 0, 0, (kvz_pixel *)ref_c, ref_w, ref_h,
                         KVZ_CHROMA_FILTER_TAPS, w, h, &src);
  if (dst14)
    kvz_sample_14bit_octpel_chroma(ctrl, src.orig_topleft, src.stride, w, h, dst14, dst_stride,
                                   mv[0] & 7, mv[1] & 7, mv);
  else
    kvz_sample_octpel_chroma(ctrl, src.orig_topleft, src.stride, w, h, dst, dst_stride,
                             mv[0] & 7, mv[1] & 7, mv);
  if (src.malloc_used) free(src.buffer);
  free(ctrl);
  return 1;
}

/* ---- quant family ---- */
typedef struct {
  int32_t qp;            /* state->qp */
  int32_t slice_is_intra;/* frame->slicetype == KVZ_SLICE_I */
  int32_t signhide;      /* cfg.signhide_enable */
  int32_t rdoq;          /* cfg.rdoq_enable */
  int32_t rdoq_skip;
  int32_t cu_is_intra;   /* block_type == CU_INTRA */
  double  lambda;
} ref_q_cfg;

typedef struct {
  encoder_state_t state;
  encoder_control_t ctrl;
  encoder_state_config_frame_t frame;
} fake_qenv;

static fake_qenv *qenv_make(const ref_q_cfg *c)
{
  fake_qenv *e = calloc(1, sizeof(*e));
  e->state.encoder_control = &e->ctrl;
  e->state.frame = &e->frame;
  e->state.qp = (int8_t)c->qp;
  e->state.lambda = c->lambda;
  e->frame.slicetype = c->slice_is_intra ? KVZ_SLICE_I : KVZ_SLICE_P;
  e->ctrl.bitdepth = KVZ_BIT_DEPTH;
  e->ctrl.cfg.signhide_enable = c->signhide;
  e->ctrl.cfg.rdoq_enable = c->rdoq;
  e->ctrl.cfg.rdoq_skip = c->rdoq_skip;
  kvz_scalinglist_init(&e->ctrl.scaling_list);
  kvz_scalinglist_process(&e->ctrl.scaling_list, KVZ_BIT_DEPTH);
  if (c->rdoq) {
    kvz_init_contexts(&e->state, c->qp, e->frame.slicetype);
  }
  return e;
}

static void qenv_free(fake_qenv *e)
{
  kvz_scalinglist_destroy(&e->ctrl.scaling_list);
  free(e);
}

int ref_shim_quant(const ref_q_cfg *c, coeff_t *coef, coeff_t *q_coef, int width, int type, int scan_idx)
{
  fake_qenv *e = qenv_make(c);
  kvz_quant(&e->state, coef, q_coef, width, width, (int8_t)type, (int8_t)scan_idx,
            c->cu_is_intra ? CU_INTRA : CU_INTER);
  qenv_free(e);
  return 1;
}

int ref_shim_dequant(const ref_q_cfg *c, coeff_t *q_coef, coeff_t *coef, int width, int type)
{
  fake_qenv *e = qenv_make(c);
  kvz_dequant(&e->state, q_coef, coef, width, width, (int8_t)type,
              c->cu_is_intra ? CU_INTRA : CU_INTER);
  qenv_free(e);
  return 1;
}

int ref_shim_rdoq(const ref_q_cfg *c, coeff_t *coef, coeff_t *q_coef, int width, int type,
                  int scan_idx, int tr_depth)
{
  fake_qenv *e = qenv_make(c);
  kvz_rdoq(&e->state, coef, q_coef, width, width, (int8_t)type, (int8_t)scan_idx,
           c->cu_is_intra ? CU_INTRA : CU_INTER, (int8_t)tr_depth);
  qenv_free(e);
  return 1;
}

/* the context states kvz_rdoq reads, in the layout of kvzo_rdoq_ctx / kvz_cuda_rdoq_ctx (148 bytes) */
typedef struct {
  uint8_t cbf_luma[4], cbf_chroma[4];
  uint8_t sig_group[4];
  uint8_t sig_luma[27], sig_chroma[15];
  uint8_t last_y_luma[15], last_y_chroma[15];
  uint8_t last_x_luma[15], last_x_chroma[15];
  uint8_t one_luma[16], one_chroma[8];
  uint8_t abs_luma[4], abs_chroma[2];
  uint8_t root_cbf;
  uint8_t trskip_luma, trskip_chroma, reserved;
} ref_rdoq_ctx;

#define CTX_COPY(dir, field, member, n) \
  for (int i_ = 0; i_ < (n); i_++) { \
    if (dir) cab->ctx.member[i_].uc_state = s->field[i_]; else s->field[i_] = cab->ctx.member[i_].uc_state; }

static void ctx_transfer(int to_cabac, cabac_data_t *cab, ref_rdoq_ctx *s)
{
  CTX_COPY(to_cabac, cbf_luma, qt_cbf_model_luma, 4)
  CTX_COPY(to_cabac, cbf_chroma, qt_cbf_model_chroma, 4)
  CTX_COPY(to_cabac, sig_group, cu_sig_coeff_group_model, 4)
  CTX_COPY(to_cabac, sig_luma, cu_sig_model_luma, 27)
  CTX_COPY(to_cabac, sig_chroma, cu_sig_model_chroma, 15)
  CTX_COPY(to_cabac, last_y_luma, cu_ctx_last_y_luma, 15)
  CTX_COPY(to_cabac, last_y_chroma, cu_ctx_last_y_chroma, 15)
  CTX_COPY(to_cabac, last_x_luma, cu_ctx_last_x_luma, 15)
  CTX_COPY(to_cabac, last_x_chroma, cu_ctx_last_x_chroma, 15)
  CTX_COPY(to_cabac, one_luma, cu_one_model_luma, 16)
  CTX_COPY(to_cabac, one_chroma, cu_one_model_chroma, 8)
  CTX_COPY(to_cabac, abs_luma, cu_abs_model_luma, 4)
  CTX_COPY(to_cabac, abs_chroma, cu_abs_model_chroma, 2)
  if (to_cabac) cab->ctx.cu_qt_root_cbf_model.uc_state = s->root_cbf;
  else s->root_cbf = cab->ctx.cu_qt_root_cbf_model.uc_state;
  if (to_cabac) {
    cab->ctx.transform_skip_model_luma.uc_state = s->trskip_luma;
    cab->ctx.transform_skip_model_chroma.uc_state = s->trskip_chroma;
  } else {
    s->trskip_luma = cab->ctx.transform_skip_model_luma.uc_state;
    s->trskip_chroma = cab->ctx.transform_skip_model_chroma.uc_state;
  }
}

/* the states kvz_init_contexts (context.c:209-300) gives a slice of this QP and type */
int ref_shim_init_rdoq_ctx(int qp, int slice_is_intra, ref_rdoq_ctx *out)
{
  ref_q_cfg c; memset(&c, 0, sizeof(c));
  c.qp = qp; c.slice_is_intra = slice_is_intra; c.rdoq = 1;
  fake_qenv *e = qenv_make(&c);
  memset(out, 0, sizeof(*out));
  ctx_transfer(0, &e->state.cabac, out);
  qenv_free(e);
  return 1;
}

/* kvz_rdoq with caller-chosen context states */
int ref_shim_rdoq_ctx(const ref_q_cfg *c, const ref_rdoq_ctx *ctx, coeff_t *coef, coeff_t *q_coef, int width,
                      int type, int scan_idx, int tr_depth)
{
  fake_qenv *e = qenv_make(c);
  ref_rdoq_ctx tmp = *ctx;
  ctx_transfer(1, &e->state.cabac, &tmp);
  kvz_rdoq(&e->state, coef, q_coef, width, width, (int8_t)type, (int8_t)scan_idx,
           c->cu_is_intra ? CU_INTRA : CU_INTER, (int8_t)tr_depth);
  qenv_free(e);
  return 1;
}

int ref_shim_quantize_residual(const ref_q_cfg *c, int width, int color, int scan_idx, int use_trskip,
                               int in_stride, int out_stride, const kvz_pixel *ref_in,
                               const kvz_pixel *pred_in, kvz_pixel *rec_out, coeff_t *coeff_out)
{
  fake_qenv *e = qenv_make(c);
  cu_info_t cu; memset(&cu, 0, sizeof(cu));
  cu.type = c->cu_is_intra ? CU_INTRA : CU_INTER;
  cu.part_size = SIZE_2Nx2N;
  int r = kvz_quantize_residual(&e->state, &cu, width, (color_t)color,
                                (coeff_scan_order_t)scan_idx, use_trskip, in_stride, out_stride,
                                ref_in, pred_in, rec_out, coeff_out);
  qenv_free(e);
  return r;
}

/* dct selector exactly as transform.c:189-215 */
int ref_shim_transform2d(int16_t *block, int16_t *coeff, int size, int color, int cu_is_intra, int inverse)
{
  encoder_control_t *ctrl = calloc(1, sizeof(*ctrl));
  ctrl->bitdepth = KVZ_BIT_DEPTH;
  if (inverse) kvz_itransform2d(ctrl, block, coeff, size, (color_t)color, cu_is_intra ? CU_INTRA : CU_INTER);
  else         kvz_transform2d(ctrl, block, coeff, size, (color_t)color, cu_is_intra ? CU_INTRA : CU_INTER);
  free(ctrl);
  return 1;
}

/* ---- intra prediction: kvz_intra_predict (intra.c:281) on caller refs ---- */
int ref_shim_intra_predict(const kvz_pixel *ref_top, const kvz_pixel *ref_left, /* 2*w+1 each, [0]=corner */
                           int log2_width, int mode, int color, int filter_boundary, kvz_pixel *dst)
{
  kvz_intra_references refs;
  memset(&refs, 0, sizeof(refs));
  int n = 2 * (1 << log2_width) + 1;
  memcpy(refs.ref.top, ref_top, n * sizeof(kvz_pixel));
  memcpy(refs.ref.left, ref_left, n * sizeof(kvz_pixel));
  refs.filtered_initialized = false;
  kvz_intra_predict(&refs, log2_width, mode, (color_t)color, dst, filter_boundary);
  return 1;
}

/* ---- SAO: the strategies that read encoder state, and kvz_sao_reconstruct (sao.c:277-337) on a fake frame ---- */
#include "sao.h"
#include "strategies/strategies-sao.h"

typedef struct {
  int32_t type, eo_class, band_position[2], offsets[10];
} ref_sao_info;

static void sao_from_plain(sao_info_t *s, const ref_sao_info *p)
{
  memset(s, 0, sizeof(*s));
  s->type = (sao_type)p->type;
  s->eo_class = (sao_eo_class)p->eo_class;
  s->band_position[0] = p->band_position[0]; s->band_position[1] = p->band_position[1];
  for (int i = 0; i < 10; i++) s->offsets[i] = p->offsets[i];
}

int ref_shim_sao_band_ddistortion(const kvz_pixel *orig, const kvz_pixel *rec, int bw, int bh, int band_pos, int *bands)
{
  fake_env *e = calloc(1, sizeof(*e));
  e->state.encoder_control = &e->ctrl;
  e->ctrl.bitdepth = KVZ_BIT_DEPTH;
  int r = kvz_sao_band_ddistortion(&e->state, orig, rec, bw, bh, band_pos, bands);
  free(e);
  return r;
}

int ref_shim_sao_reconstruct_color(const kvz_pixel *rec, kvz_pixel *new_rec, const ref_sao_info *p, int stride,
                                   int new_stride, int bw, int bh, int color)
{
  encoder_control_t *ctrl = calloc(1, sizeof(*ctrl));
  ctrl->bitdepth = KVZ_BIT_DEPTH;
  sao_info_t s;
  sao_from_plain(&s, p);
  kvz_sao_reconstruct_color(ctrl, rec, new_rec, &s, stride, new_stride, bw, bh, (color_t)color);
  free(ctrl);
  return 1;
}

/* pre: deblocked plane of this colour (stride == its width); out: the frame's rec plane of this colour, which
 * kvz_sao_reconstruct writes at (x, y); luma_w / luma_h: frame size in luma samples */
int ref_shim_sao_reconstruct(const kvz_pixel *pre, kvz_pixel *out, int luma_w, int luma_h, int x, int y, int w, int h,
                             const ref_sao_info *p, int color)
{
  fake_env *e = calloc(1, sizeof(*e));
  e->state.encoder_control = &e->ctrl;
  e->state.tile = &e->tile;
  e->tile.frame = &e->vframe;
  e->ctrl.bitdepth = KVZ_BIT_DEPTH;
  e->vframe.width = luma_w; e->vframe.height = luma_h;
  e->vframe.rec = &e->pic;
  e->pic.width = luma_w; e->pic.height = luma_h; e->pic.stride = luma_w;
  e->pic.data[color] = out;
  const int plane_w = color ? luma_w >> 1 : luma_w;
  sao_info_t s;
  sao_from_plain(&s, p);
  kvz_sao_reconstruct(&e->state, pre + (size_t)y * plane_w + x, plane_w, x, y, w, h, &s, (color_t)color);
  free(e);
  return 1;
}

/* ---- deblocking: kvz_filter_deblock_lcu (filter.c:770-779) over a whole picture on a fake frame ---- */
#include "filter.h"
#include "cu.h"

typedef struct {
  uint8_t type, depth, tr_depth, part_size, cbf_y, mv_dir, mv_ref[2];
  int16_t mv[2][2];
  int8_t qp;
  uint8_t reserved[3];
} ref_scu;

typedef struct {
  int32_t width, height, qp, cu_qp, beta_offset_div2, tc_offset_div2, slice_b, chroma;
  uint8_t ref_LX[2][16];
} ref_deblock_cfg;

int ref_shim_deblock_picture(const ref_deblock_cfg *c, const ref_scu *scus, kvz_pixel *y, kvz_pixel *u, kvz_pixel *v)
{
  fake_env *e = calloc(1, sizeof(*e));
  e->state.encoder_control = &e->ctrl;
  e->state.tile = &e->tile;
  e->state.frame = &e->frame;
  e->tile.frame = &e->vframe;
  e->ctrl.bitdepth = KVZ_BIT_DEPTH;
  e->ctrl.chroma_format = c->chroma ? KVZ_CSP_420 : KVZ_CSP_400;
  e->ctrl.max_qp_delta_depth = c->cu_qp ? 0 : -1;
  e->ctrl.cfg.deblock_beta = c->beta_offset_div2;
  e->ctrl.cfg.deblock_tc = c->tc_offset_div2;
  e->state.qp = (int8_t)c->qp;
  e->frame.QP = (int8_t)c->qp;
  e->frame.slicetype = c->slice_b ? KVZ_SLICE_B : KVZ_SLICE_P;
  memcpy(e->frame.ref_LX, c->ref_LX, sizeof(c->ref_LX));
  e->vframe.width = c->width; e->vframe.height = c->height;
  e->vframe.width_in_lcu = (c->width + 63) / 64; e->vframe.height_in_lcu = (c->height + 63) / 64;
  e->vframe.rec = &e->pic;
  e->pic.width = c->width; e->pic.height = c->height; e->pic.stride = c->width;
  e->pic.y = e->pic.data[0] = y; e->pic.u = e->pic.data[1] = u; e->pic.v = e->pic.data[2] = v;
  e->pic.chroma_format = e->ctrl.chroma_format;
  e->vframe.cu_array = kvz_cu_array_alloc(c->width, c->height);
  for (int sy = 0; sy < c->height / 4; sy++)
    for (int sx = 0; sx < c->width / 4; sx++) {
      const ref_scu *s = &scus[sy * (c->width / 4) + sx];
      cu_info_t *cu = kvz_cu_array_at(e->vframe.cu_array, sx * 4, sy * 4);
      memset(cu, 0, sizeof(*cu));
      cu->type = s->type; cu->depth = s->depth; cu->tr_depth = s->tr_depth; cu->part_size = s->part_size;
      cu->qp = (uint8_t)s->qp;
      if (s->cbf_y) cbf_set(&cu->cbf, s->tr_depth, COLOR_Y);
      if (s->type == CU_INTER) {
        cu->inter.mv_dir = s->mv_dir;
        cu->inter.mv_ref[0] = s->mv_ref[0]; cu->inter.mv_ref[1] = s->mv_ref[1];
        memcpy(cu->inter.mv, s->mv, sizeof(s->mv));
      }
    }
  for (int ly = 0; ly < c->height; ly += 64)
    for (int lx = 0; lx < c->width; lx += 64) kvz_filter_deblock_lcu(&e->state, lx, ly);
  kvz_cu_array_free(&e->vframe.cu_array);
  free(e);
  return 1;
}

/* kvz_get_coeff_cost (rdo.c:208-222) on its CABAC path with caller-chosen context states and coder range */
uint32_t ref_shim_coeff_cost(const ref_rdoq_ctx *ctx, uint32_t range, const coeff_t *coeff, int width, int type,
                             int scan_mode, int signhide, int trskip_enable)
{
  ref_q_cfg c; memset(&c, 0, sizeof(c));
  c.qp = 30; c.rdoq = 1; c.signhide = signhide;
  fake_qenv *e = qenv_make(&c);
  ref_rdoq_ctx tmp = *ctx;
  ctx_transfer(1, &e->state.cabac, &tmp);
  e->state.cabac.range = range;
  e->ctrl.cfg.fast_residual_cost_limit = 0;
  e->ctrl.cfg.trskip_enable = trskip_enable;
  uint32_t bits = kvz_get_coeff_cost(&e->state, coeff, width, type, (int8_t)scan_mode);
  qenv_free(e);
  return bits;
}

/* ---- inter candidate derivation (A18): kvz_inter_get_merge_cand / kvz_inter_get_mv_cand on a fake frame ---- */
#include "inter.h"
#include "imagelist.h"

typedef struct {
  int32_t width, height, in_width, in_height, tile_offset_x, tile_offset_y, poc, slice_b, tmvp_enable, n_refs;
  int32_t ref_pocs[16];
  int32_t col_ref_pocs[2][16];
  uint8_t ref_LX[2][16];
  uint8_t ref_LX_size[2];
  uint8_t reserved[2];
  int32_t ref_scu_pitch;
  const ref_scu *scus;
  const ref_scu *ref_scus[16];
} ref_cand_frame;

typedef struct {
  fake_env env;
  image_list_t list;
  kvz_picture *images[16];
  kvz_picture pics[16];
  cu_array_t *cuas[16];
  int32_t pocs[16];
  uint8_t ref_LXs[16][2][16];
  lcu_t lcu;
} cand_env;

static void cu_from_scu(cu_info_t *cu, const ref_scu *s)
{
  memset(cu, 0, sizeof(*cu));
  cu->type = s->type; cu->depth = s->depth; cu->tr_depth = s->tr_depth; cu->part_size = s->part_size;
  if (s->type == CU_INTER) {
    cu->inter.mv_dir = s->mv_dir;
    cu->inter.mv_ref[0] = s->mv_ref[0]; cu->inter.mv_ref[1] = s->mv_ref[1];
    memcpy(cu->inter.mv, s->mv, sizeof(s->mv));
  }
}

static cand_env *cand_env_make(const ref_cand_frame *f, int x, int y)
{
  cand_env *e = calloc(1, sizeof(*e));
  e->env.state.encoder_control = &e->env.ctrl;
  e->env.state.tile = &e->env.tile;
  e->env.state.frame = &e->env.frame;
  e->env.tile.frame = &e->env.vframe;
  e->env.tile.offset_x = f->tile_offset_x; e->env.tile.offset_y = f->tile_offset_y;
  e->env.ctrl.in.width = f->in_width; e->env.ctrl.in.height = f->in_height;
  e->env.ctrl.cfg.tmvp_enable = f->tmvp_enable;
  e->env.vframe.width = f->width; e->env.vframe.height = f->height;
  e->env.frame.poc = f->poc;
  e->env.frame.slicetype = f->slice_b ? KVZ_SLICE_B : KVZ_SLICE_P;
  memcpy(e->env.frame.ref_LX, f->ref_LX, sizeof(f->ref_LX));
  e->env.frame.ref_LX_size[0] = f->ref_LX_size[0]; e->env.frame.ref_LX_size[1] = f->ref_LX_size[1];
  e->env.frame.ref = &e->list;
  e->list.images = e->images; e->list.cu_arrays = e->cuas; e->list.pocs = e->pocs; e->list.ref_LXs = e->ref_LXs;
  e->list.size = 16; e->list.used_size = f->n_refs;
  const int rows = (f->in_height + 3) / 4;
  for (int r = 0; r < f->n_refs; r++) {
    e->images[r] = &e->pics[r];
    e->pocs[r] = f->ref_pocs[r];
    e->cuas[r] = kvz_cu_array_alloc(f->ref_scu_pitch * 4, rows * 4);
    for (int sy = 0; sy < rows; sy++)
      for (int sx = 0; sx < f->ref_scu_pitch; sx++)
        cu_from_scu(kvz_cu_array_at(e->cuas[r], sx * 4, sy * 4), &f->ref_scus[r][sy * f->ref_scu_pitch + sx]);
    /* the colocated picture's own lists: identity index lists over its ref_pocs, which carry col_ref_pocs */
    for (int l = 0; l < 2; l++)
      for (int i = 0; i < 16; i++) e->ref_LXs[r][l][i] = (uint8_t)i;
  }
  if (f->n_refs && f->ref_LX_size[0]) {
    /* ref_pocs is indexed through ref_LXs[c][list][idx]; give list 1 the indices 8..15 so both tables fit */
    const int c = f->ref_LX[0][0];
    for (int i = 0; i < 8; i++) {
      e->ref_LXs[c][0][i] = (uint8_t)i;       e->pics[c].ref_pocs[i] = f->col_ref_pocs[0][i];
      e->ref_LXs[c][1][i] = (uint8_t)(8 + i); e->pics[c].ref_pocs[8 + i] = f->col_ref_pocs[1][i];
    }
  }
  /* the working LCU around (x, y): its own SCUs, the column to the left, the row above, the top-right SCU */
  const int lx = x & ~63, ly = y & ~63, per_row = f->width / 4;
  for (int sy = -1; sy < 16; sy++)
    for (int sx = -1; sx < 16; sx++) {
      const int px = lx + sx * 4, py = ly + sy * 4;
      if (px < 0 || py < 0 || px >= f->width || py >= f->height) continue;
      cu_from_scu(LCU_GET_CU_AT_PX(&e->lcu, sx * 4, sy * 4), &f->scus[(py / 4) * per_row + px / 4]);
    }
  if (ly > 0 && lx + 64 < f->width) cu_from_scu(LCU_GET_TOP_RIGHT_CU(&e->lcu), &f->scus[((ly - 4) / 4) * per_row + (lx + 64) / 4]);
  return e;
}

static void cand_env_free(cand_env *e, const ref_cand_frame *f)
{
  for (int r = 0; r < f->n_refs; r++) kvz_cu_array_free(&e->cuas[r]);
  free(e);
}

typedef struct { uint8_t dir, ref[2], reserved; int16_t mv[2][2]; } ref_merge_out;

int ref_shim_merge_cand(const ref_cand_frame *f, int x, int y, int w, int h, int use_a1, int use_b1, ref_merge_out out[5])
{
  cand_env *e = cand_env_make(f, x, y);
  inter_merge_cand_t mc[MRG_MAX_NUM_CANDS];
  memset(mc, 0, sizeof(mc));
  int n = kvz_inter_get_merge_cand(&e->env.state, x, y, w, h, use_a1, use_b1, mc, &e->lcu);
  for (int i = 0; i < 5; i++) {
    out[i].dir = mc[i].dir; out[i].ref[0] = mc[i].ref[0]; out[i].ref[1] = mc[i].ref[1]; out[i].reserved = 0;
    memcpy(out[i].mv, mc[i].mv, sizeof(out[i].mv));
  }
  cand_env_free(e, f);
  return n;
}

int ref_shim_mv_cand(const ref_cand_frame *f, int x, int y, int w, int h, int reflist, int lx_idx, int16_t mv_cand[2][2])
{
  cand_env *e = cand_env_make(f, x, y);
  cu_info_t cur;
  memset(&cur, 0, sizeof(cur));
  cur.type = CU_INTER;
  cur.inter.mv_ref[reflist] = (uint8_t)lx_idx;
  kvz_inter_get_mv_cand(&e->env.state, x, y, w, h, mv_cand, &cur, &e->lcu, (int8_t)reflist);
  cand_env_free(e, f);
  return 1;
}

/* ---- config 5: speed loops over the selected strategies, in the shape of tests/speed_tests.c:144-303 (one thread, a
 * buffer of `blocks` contiguous n x n blocks walked `reps` times).  Returns the seconds the walks took; 0 on a bad
 * request.  what: 0 sad_NxN, 1 satd_NxN, 2 dct, 3 idct, 4 dct + quant (sign hiding on), 5 quarter-pel luma
 * interpolation of an n x n block (both fractions non-zero) ---- */
#include <time.h>
#include "strategies/strategies-picture.h"
#include "strategies/strategies-ipol.h"

double ref_shim_speed(int what, int n, int blocks, int reps)
{
  const size_t nn = (size_t)n * n;
  if (blocks <= 0 || reps <= 0) return 0.0;
  struct timespec t0, t1;
  unsigned sink = 0;
  if (what == 0 || what == 1) {
    cost_pixel_nxn_func *fn = what == 0 ? kvz_pixels_get_sad_func(n) : kvz_pixels_get_satd_func(n);
    if (!fn) return 0.0;
    kvz_pixel *a = NULL, *b = NULL;
    if (posix_memalign((void **)&a, 64, nn * blocks) || posix_memalign((void **)&b, 64, nn * blocks)) return 0.0;
    for (size_t i = 0; i < nn * blocks; i++) { a[i] = (kvz_pixel)(i * 7 + (i >> 5)); b[i] = (kvz_pixel)(i * 13 + (i >> 3)); }
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int r = 0; r < reps; r++)
      for (int k = 0; k < blocks; k++) sink += fn(a + k * nn, b + k * nn);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(a); free(b);
  } else if (what >= 2 && what <= 4) {
    if (n > 32) return 0.0;
    dct_func *fn = what == 3 ? kvz_get_idct_func((int8_t)n, COLOR_Y, CU_INTER) : kvz_get_dct_func((int8_t)n, COLOR_Y, CU_INTER);
    if (!fn) return 0.0;
    int16_t *in = NULL, *out = NULL, *q = NULL;
    if (posix_memalign((void **)&in, 64, 2 * nn * blocks) || posix_memalign((void **)&out, 64, 2 * nn * blocks) ||
        posix_memalign((void **)&q, 64, 2 * nn * blocks)) return 0.0;
    for (size_t i = 0; i < nn * blocks; i++) in[i] = (int16_t)((int)((i * 31 + (i >> 4)) % 121) - 60);
    ref_q_cfg c; memset(&c, 0, sizeof(c));
    c.qp = 32; c.signhide = 1;
    fake_qenv *e = qenv_make(&c);
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int r = 0; r < reps; r++)
      for (int k = 0; k < blocks; k++) {
        fn(KVZ_BIT_DEPTH, in + k * nn, out + k * nn);
        if (what == 4) kvz_quant(&e->state, out + k * nn, q + k * nn, n, n, 0, 0, CU_INTER);
      }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    sink += (unsigned)out[0] + (unsigned)q[0];
    qenv_free(e);
    free(in); free(out); free(q);
  } else if (what == 5) {
    /* blocks cut from one plane `stride` wide with the 8-tap margins round them */
    const int stride = 4096, rows = n + 8;
    const int per_row = stride / (n + 8);
    const int plane_rows = (blocks / per_row + 1) * rows;
    kvz_pixel *src = NULL, *dst = NULL;
    if (posix_memalign((void **)&src, 64, (size_t)stride * plane_rows) || posix_memalign((void **)&dst, 64, nn * blocks)) return 0.0;
    for (size_t i = 0; i < (size_t)stride * plane_rows; i++) src[i] = (kvz_pixel)(i * 11 + (i >> 7));
    encoder_control_t *ctrl = calloc(1, sizeof(*ctrl));
    ctrl->bitdepth = KVZ_BIT_DEPTH;
    const int16_t mv[2] = { 1, 3 };
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int r = 0; r < reps; r++)
      for (int k = 0; k < blocks; k++) {
        kvz_pixel *p = src + (size_t)(k / per_row) * rows * stride + (size_t)(k % per_row) * (n + 8) + 4 * stride + 4;
        kvz_sample_quarterpel_luma(ctrl, p, (int16_t)stride, n, n, dst + k * nn, (int16_t)n, 1, 1, mv);
      }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    sink += dst[0];
    free(ctrl); free(src); free(dst);
  } else {
    return 0.0;
  }
  volatile unsigned keep = sink; (void)keep;
  return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
