/*
 * kvazaar_b200/host/kvz_glue_strategies.c -- registers the per-call "cuda" strategies with the reference's strategy
 * selector.  Compiled once per strategy family INSTEAD of src/strategies/strategies-<family>.c, with
 *   -DGLUE_SRC='"strategies/strategies-<family>.c"' -DGLUE_FAMILY=<family> -DGLUE_FAMILY_<family>
 * It includes that reference file in place (from the reference tree, at build time) with its registration function
 * renamed to kvz_strategy_register_<family>_host, and defines kvz_strategy_register_<family> as: the host's own
 * strategies first, then -- in KVZ_CUDA_MODE=percall|both -- kvz_strategy_register_<family>_cuda(opaque, bitdepth)
 * (include/kvz_cuda_strategies.h; strategy name "cuda", priority 50 > avx2's 40).  This is the one line per family
 * INTEGRATION.md section 1 describes; kvz_strategyselector_init (src/strategyselector.c:43-213) and
 * KVAZAAR_OVERRIDE_<type> (:258-324) work unchanged.
 *
 * The strategies whose signatures carry encoder structures (quant, dequant, quantize_residual: strategies-quant.h:36-46;
 * sao_band_ddistortion, sao_reconstruct_color: strategies-sao.h:44-60) get their wrappers here, where those types
 * are visible, and are registered from here.  Nothing falls back to the CPU: what the device path does not implement
 * (scaling lists) aborts with a message.
 */
#define GLUE_CAT3_(a, b, c) a##b##c
#define GLUE_CAT3(a, b, c) GLUE_CAT3_(a, b, c)
#define GLUE_REGISTER GLUE_CAT3(kvz_strategy_register_, GLUE_FAMILY, )
#define GLUE_REGISTER_HOST GLUE_CAT3(kvz_strategy_register_, GLUE_FAMILY, _host)
#define GLUE_REGISTER_CUDA GLUE_CAT3(kvz_strategy_register_, GLUE_FAMILY, _cuda)

#if defined(GLUE_FAMILY_picture)
#define kvz_strategy_register_picture kvz_strategy_register_picture_host
#elif defined(GLUE_FAMILY_dct)
#define kvz_strategy_register_dct kvz_strategy_register_dct_host
#elif defined(GLUE_FAMILY_quant)
#define kvz_strategy_register_quant kvz_strategy_register_quant_host
#elif defined(GLUE_FAMILY_intra)
#define kvz_strategy_register_intra kvz_strategy_register_intra_host
#elif defined(GLUE_FAMILY_ipol)
#define kvz_strategy_register_ipol kvz_strategy_register_ipol_host
#elif defined(GLUE_FAMILY_sao)
#define kvz_strategy_register_sao kvz_strategy_register_sao_host
#elif defined(GLUE_FAMILY_nal)
#define kvz_strategy_register_nal kvz_strategy_register_nal_host
#else
#error "GLUE_FAMILY_<family> must be defined"
#endif
#include GLUE_SRC
#undef kvz_strategy_register_picture
#undef kvz_strategy_register_dct
#undef kvz_strategy_register_quant
#undef kvz_strategy_register_intra
#undef kvz_strategy_register_ipol
#undef kvz_strategy_register_sao
#undef kvz_strategy_register_nal

#include <string.h>

#include "kvz_cuda_strategies.h"
#include "kvz_glue.h"
#include "strategyselector.h"

#if defined(GLUE_FAMILY_quant)
#include "cabac.h"
#include "encoderstate.h"

static kvz_cuda_quant_cfg glue_qcfg(const encoder_state_t *s, int8_t block_type)
{
  if (s->encoder_control->scaling_list.enable)
    kvz_glue_die("scaling lists (--scaling-list, quant-generic.c:46-47) are not implemented by the cuda quant strategies");
  kvz_cuda_quant_cfg c = { s->qp, s->encoder_control->bitdepth, s->frame->slicetype == KVZ_SLICE_I,
                           s->encoder_control->cfg.signhide_enable, block_type == CU_INTRA };
  return c;
}

/* the context states kvz_rdoq reads (rdo.c:548-885), from the calling thread's own coder (cabac.h:53-90) */
#define GLUE_CTX(field, member, n) for (int i_ = 0; i_ < (n); i_++) out->field[i_] = ctx->member[i_].uc_state
static void glue_rdoq_ctx(const encoder_state_t *s, kvz_cuda_rdoq_ctx *out)
{
  const __typeof__(s->cabac.ctx) *ctx = &s->cabac.ctx;
  memset(out, 0, sizeof(*out));
  GLUE_CTX(cbf_luma, qt_cbf_model_luma, 4);
  GLUE_CTX(cbf_chroma, qt_cbf_model_chroma, 4);
  GLUE_CTX(sig_group, cu_sig_coeff_group_model, 4);
  GLUE_CTX(sig_luma, cu_sig_model_luma, 27);
  GLUE_CTX(sig_chroma, cu_sig_model_chroma, 15);
  GLUE_CTX(last_y_luma, cu_ctx_last_y_luma, 15);
  GLUE_CTX(last_y_chroma, cu_ctx_last_y_chroma, 15);
  GLUE_CTX(last_x_luma, cu_ctx_last_x_luma, 15);
  GLUE_CTX(last_x_chroma, cu_ctx_last_x_chroma, 15);
  GLUE_CTX(one_luma, cu_one_model_luma, 16);
  GLUE_CTX(one_chroma, cu_one_model_chroma, 8);
  GLUE_CTX(abs_luma, cu_abs_model_luma, 4);
  GLUE_CTX(abs_chroma, cu_abs_model_chroma, 2);
  out->root_cbf = ctx->cu_qt_root_cbf_model.uc_state;
  out->trskip_luma = ctx->transform_skip_model_luma.uc_state;
  out->trskip_chroma = ctx->transform_skip_model_chroma.uc_state;
}

/* quant_func (strategies-quant.h:36) */
static void glue_quant_cuda(const encoder_state_t *const s, coeff_t *coef, coeff_t *q_coef, int32_t w, int32_t h,
                            int8_t type, int8_t scan_idx, int8_t block_type)
{
  kvz_cuda_quant_cfg c = glue_qcfg(s, block_type);
  if (!kvz_cuda_quant_host(&c, coef, q_coef, w, h, type, scan_idx)) kvz_glue_die("kvz_cuda_quant_host");
  KVZ_GLUE_COUNT(KVZ_GLUE_CNT_QUANT);
}

/* dequant_func (strategies-quant.h:44) */
static void glue_dequant_cuda(const encoder_state_t *const s, coeff_t *q_coef, coeff_t *coef, int32_t w, int32_t h,
                              int8_t type, int8_t block_type)
{
  kvz_cuda_quant_cfg c = glue_qcfg(s, block_type);
  if (!kvz_cuda_dequant_host(&c, q_coef, coef, w, h, type)) kvz_glue_die("kvz_cuda_dequant_host");
  KVZ_GLUE_COUNT(KVZ_GLUE_CNT_QUANT);
}

/* quant_residual_func (strategies-quant.h:38-42): with cfg.rdoq_enable the device runs kvz_rdoq too (quant-generic.c:211-221) */
static int glue_quantize_residual_cuda(encoder_state_t *const s, const cu_info_t *const cu, const int width,
                                       const color_t color, const coeff_scan_order_t scan, const int use_trskip,
                                       const int in_stride, const int out_stride, const kvz_pixel *const ref_in,
                                       const kvz_pixel *const pred_in, kvz_pixel *rec_out, coeff_t *coeff_out)
{
  kvz_cuda_quant_cfg c = glue_qcfg(s, cu->type);
  const kvz_config *cfg = &s->encoder_control->cfg;
  KVZ_GLUE_COUNT(KVZ_GLUE_CNT_QUANT);
  if (cfg->rdoq_enable && (width > 4 || !cfg->rdoq_skip)) {
    kvz_cuda_rdoq_ctx ctx;
    glue_rdoq_ctx(s, &ctx);
    const int tr_depth = cu->tr_depth - cu->depth + (cu->part_size == SIZE_NxN ? 1 : 0);
    return kvz_cuda_quantize_residual_rdoq_host(&c, &ctx, s->lambda, tr_depth, cfg->rdoq_skip, width, color, scan, use_trskip,
                                                in_stride, out_stride, ref_in, pred_in, rec_out, coeff_out);
  }
  return kvz_cuda_quantize_residual_host(&c, width, color, scan, use_trskip, in_stride, out_stride, ref_in, pred_in, rec_out,
                                         coeff_out);
}
#endif  /* quant */

#if defined(GLUE_FAMILY_sao)
#include "encoderstate.h"
#include "sao.h"

static kvz_cuda_sao_params glue_sao_params(const sao_info_t *s)
{
  kvz_cuda_sao_params p;
  memset(&p, 0, sizeof(p));
  p.type = (int8_t)s->type; p.eo_class = (int8_t)s->eo_class;
  p.band_position[0] = (int8_t)s->band_position[0]; p.band_position[1] = (int8_t)s->band_position[1];
  for (int i = 0; i < 10; i++) p.offsets[i] = (int8_t)s->offsets[i];
  return p;
}

/* sao_reconstruct_color_func (strategies-sao.h:52-60) */
static void glue_sao_reconstruct_color_cuda(const encoder_control_t *const enc, const kvz_pixel *rec, kvz_pixel *new_rec,
                                            const sao_info_t *sao, int stride, int new_stride, int w, int h, color_t color)
{
  kvz_cuda_sao_params p = glue_sao_params(sao);
  kvz_cuda_sao_reconstruct_color_host(enc->bitdepth, rec, new_rec, &p, stride, new_stride, w, h, color);
  KVZ_GLUE_COUNT(KVZ_GLUE_CNT_SAO);
}

/* sao_band_ddistortion_func (strategies-sao.h:44-50) */
static int glue_sao_band_ddistortion_cuda(const encoder_state_t *const state, const kvz_pixel *orig, const kvz_pixel *rec,
                                          int w, int h, int band_pos, int sao_bands[4])
{
  KVZ_GLUE_COUNT(KVZ_GLUE_CNT_SAO);
  return kvz_cuda_sao_band_ddistortion_host(state->encoder_control->bitdepth, orig, rec, w, h, band_pos, sao_bands);
}
#endif  /* sao */

int GLUE_REGISTER(void *opaque, uint8_t bitdepth)
{
  if (!GLUE_REGISTER_HOST(opaque, bitdepth)) return 0;
  if (!(kvz_glue_mode() & KVZ_GLUE_PERCALL)) return 1;
  int ok = GLUE_REGISTER_CUDA(opaque, bitdepth);
#if defined(GLUE_FAMILY_quant)
  if (ok && bitdepth == 8) {
    ok &= kvz_strategyselector_register(opaque, "quant", "cuda", 50, &glue_quant_cuda);
    ok &= kvz_strategyselector_register(opaque, "quantize_residual", "cuda", 50, &glue_quantize_residual_cuda);
    ok &= kvz_strategyselector_register(opaque, "dequant", "cuda", 50, &glue_dequant_cuda);
  }
#endif
#if defined(GLUE_FAMILY_sao)
  if (ok && bitdepth == 8) {
    ok &= kvz_strategyselector_register(opaque, "sao_reconstruct_color", "cuda", 50, &glue_sao_reconstruct_color_cuda);
    ok &= kvz_strategyselector_register(opaque, "sao_band_ddistortion", "cuda", 50, &glue_sao_band_ddistortion_cuda);
  }
#endif
  if (!ok) kvz_glue_die("registration of the cuda strategies");
  return ok;
}
