/*
 * oracle/kvz_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the Kvazaar inner-encode-loop arithmetic (SURVEY.md 8a
 * rows A1-A17), written from the algorithm description and checked against
 * (a) the reference's own known-answer tests (tests/sad_tests.c,
 * tests/satd_tests.c, tests/dct_tests.c, tests/coeff_sum_tests.c) and
 * (b) the unmodified reference compiled into oracle/_ref/libkvazaar_ref.so.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * load this library.  The product (libkvz_cuda.so) never links or calls it.
 *
 * Pixel-typed functions exist twice: suffix _8 (uint8_t pixels, 8-bit build
 * of the reference) and _10 (uint16_t pixels, KVZ_BIT_DEPTH=10 build).
 */
#ifndef KVZ_ORACLE_H_
#define KVZ_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- motion-search side structures (bit-depth independent) ---- */

typedef struct {
  uint8_t dir;        /* 1 = L0, 2 = L1, 3 = bi (inter.h:36-41) */
  uint8_t ref[2];
  int16_t mv[2][2];   /* quarter-pel */
} kvzo_merge_cand;

typedef struct {
  /* frame geometry */
  int32_t width, height;          /* luma plane, stride == width */
  /* PU */
  int32_t origin_x, origin_y, pu_w, pu_h;
  int32_t tile_off_x, tile_off_y;
  /* candidates */
  int16_t mv_cand[2][2];
  int32_t num_merge_cand;
  kvzo_merge_cand merge_cand[5];
  uint8_t ref_LX[2][16];
  int32_t ref_idx;
  int32_t extra_mv[2];            /* quarter-pel */
  /* config */
  int32_t lambda_sqrt_i;          /* (int32_t)(lambda_sqrt + 0.5), search_inter.c:411 */
  int32_t owf, wpp, sao_type, deblock_enable, mv_constraint;
  int32_t max_ref_lcu_right, max_ref_lcu_down;
  int32_t fme_level, me_early_termination, me_max_steps;
  int32_t ime_algorithm;          /* enum kvz_ime_algorithm (kvazaar.h) */
  /* in/out */
  int32_t best_mv[2];             /* quarter-pel */
  uint32_t best_cost, best_bitcost;
} kvzo_me_job;

enum { KVZO_IME_HEXBS = 0, KVZO_IME_TZ = 1, KVZO_IME_FULL = 2, KVZO_IME_FULL8 = 3,
       KVZO_IME_FULL16 = 4, KVZO_IME_FULL32 = 5, KVZO_IME_FULL64 = 6, KVZO_IME_DIA = 7 };

/* search_inter.c:373-412 */
uint32_t kvzo_mvd_cost(const kvzo_me_job *j, int x, int y, int mv_shift, uint32_t *bitcost);
/* search_inter.c:87-172 (x,y quarter-pel) */
int kvzo_fracmv_within_tile(const kvzo_me_job *j, int x, int y);
/* search_inter.c:326-370 with cost_out == NULL */
int kvzo_select_mv_cand(const int16_t mv_cand[2][2], int mv_x, int mv_y);

/* ---- transforms / quant (bit depth is a run-time argument) ---- */

/* strategies/generic/dct-generic.c:206-617; n in {4,8,16,32}; dst4 != 0
 * selects the 4x4 DST (strategies-dct.c:59-113) */
void kvzo_fwd_transform(int n, int dst4, int bitdepth, const int16_t *in, int16_t *out);
void kvzo_inv_transform(int n, int dst4, int bitdepth, const int16_t *in, int16_t *out);

/* the fields of cu_info_t (cu.h:134-175) the deblocking filter reads, one record per 4x4 SCU, raster order */
typedef struct {
  uint8_t type;             /* cu_type_t: 1 intra, 2 inter */
  uint8_t depth, tr_depth;  /* CU and TU quadtree depths (CU width = 64 >> depth) */
  uint8_t part_size;        /* part_mode_t 0..7 */
  uint8_t cbf_y;            /* cbf_is_set(cbf, tr_depth, COLOR_Y) */
  uint8_t mv_dir;           /* inter.mv_dir: 1 L0, 2 L1, 3 both */
  uint8_t mv_ref[2];        /* inter.mv_ref */
  int16_t mv[2][2];         /* inter.mv, quarter-pel */
  int8_t qp;
  uint8_t reserved[3];
} kvzo_scu;                 /* 20 bytes */

typedef struct {
  int32_t width, height;    /* frame (tile) size in luma samples, multiples of 8 */
  int32_t qp;               /* state->qp */
  int32_t cu_qp;            /* encoder_control->max_qp_delta_depth >= 0: QP averaged over the two CUs */
  int32_t beta_offset_div2, tc_offset_div2;   /* cfg.deblock_beta, cfg.deblock_tc */
  int32_t slice_b;          /* frame->slicetype == KVZ_SLICE_B */
  int32_t chroma;           /* chroma_format != KVZ_CSP_400 */
  uint8_t ref_LX[2][16];    /* frame->ref_LX */
} kvzo_deblock_cfg;

/* what the inter candidate derivation reads of the frame (A18) */
typedef struct {
  int32_t width, height;          /* state->tile->frame size, luma samples */
  int32_t in_width, in_height;    /* encoder_control->in.width / height */
  int32_t tile_offset_x, tile_offset_y;
  int32_t poc;                    /* state->frame->poc */
  int32_t slice_b, tmvp_enable;
  int32_t n_refs;                 /* state->frame->ref->used_size */
  int32_t ref_pocs[16];           /* state->frame->ref->pocs */
  int32_t col_ref_pocs[2][16];    /* images[c]->ref_pocs[ref_LXs[c][list][idx]] for c = ref_LX[0][0] */
  uint8_t ref_LX[2][16];
  uint8_t ref_LX_size[2];
  uint8_t reserved[2];
  int32_t ref_scu_pitch;          /* SCU records per row of the reference pictures' arrays */
  const kvzo_scu *scus;           /* this frame (tile), width / 4 records per row */
  const kvzo_scu *ref_scus[16];   /* state->frame->ref->cu_arrays */
} kvzo_cand_frame;

void kvzo_inter_mv_cand(const kvzo_cand_frame *f, int x, int y, int w, int h, int reflist, int lx_idx, int16_t mv_cand[2][2]);
int kvzo_inter_merge_cand(const kvzo_cand_frame *f, int x, int y, int w, int h, int use_a1, int use_b1, kvzo_merge_cand out[5]);
void kvzo_colocated_start_mv(const kvzo_cand_frame *f, int x, int y, int w, int h, int ref_idx, int16_t mv[2]);

/* the fields of sao_info_t (sao.h:43-51) the SAO filters and distortion functions read */
typedef struct {
  int32_t type;             /* 0 SAO_TYPE_NONE, 1 SAO_TYPE_BAND, 2 SAO_TYPE_EDGE */
  int32_t eo_class;         /* 0 horizontal, 1 vertical, 2 diagonal 135, 3 diagonal 45 (sao.h:59-64) */
  int32_t band_position[2]; /* [0] Y or U, [1] V */
  int32_t offsets[10];      /* [cat] for Y / U, [5 + cat] for V */
} kvzo_sao_info;

typedef struct {
  int32_t qp;             /* state->qp (luma QP of the CTU) */
  int32_t bitdepth;
  int32_t slice_is_intra; /* frame->slicetype == KVZ_SLICE_I */
  int32_t signhide;       /* cfg.signhide_enable */
  int32_t cu_is_intra;    /* block_type == CU_INTRA (scaling-list class; flat lists ignore it) */
} kvzo_q_cfg;

int32_t kvzo_scaled_qp(int type, int qp, int qp_offset);     /* transform.c:129-143 */
/* quant-generic.c:37-163, flat scaling list; type 0 luma / 2 chroma; scan_idx 0 diag 1 hor 2 ver */
void kvzo_quant(const kvzo_q_cfg *c, const int16_t *coef, int16_t *q_coef, int width, int type, int scan_idx);
/* quant-generic.c:279-321, flat scaling list */
void kvzo_dequant(const kvzo_q_cfg *c, const int16_t *q_coef, int16_t *coef, int width, int type);
uint32_t kvzo_coeff_abs_sum(const int16_t *c, uint32_t n);    /* quant-generic.c:323-330 */
/* kvz_g_sig_last_scan[scan_idx][log2_size-1] (tables.c): position list in coding order */
void kvzo_scan_order(int scan_idx, int log2_size, uint32_t *out);
void kvzo_transformskip(int bitdepth, const int16_t *block, int16_t *coeff, int n);   /* transform.c:151-161 */
void kvzo_itransformskip(int bitdepth, int16_t *block, const int16_t *coeff, int n);  /* transform.c:169-181 */

/* ---- A15: rate-distortion optimised quantisation (rdo.c:548-885) ---- */

/* the CABAC context states (cabac_ctx_t.uc_state, cabac.h:36-39) that kvz_rdoq reads from
 * state->cabac.ctx (cabac.h:53-90), in the order of that struct */
typedef struct {
  uint8_t cbf_luma[4], cbf_chroma[4];            /* qt_cbf_model_luma / _chroma */
  uint8_t sig_group[4];                          /* cu_sig_coeff_group_model */
  uint8_t sig_luma[27], sig_chroma[15];          /* cu_sig_model_* */
  uint8_t last_y_luma[15], last_y_chroma[15];    /* cu_ctx_last_y_* */
  uint8_t last_x_luma[15], last_x_chroma[15];    /* cu_ctx_last_x_* */
  uint8_t one_luma[16], one_chroma[8];           /* cu_one_model_* */
  uint8_t abs_luma[4], abs_chroma[2];            /* cu_abs_model_* */
  uint8_t root_cbf;                              /* cu_qt_root_cbf_model */
  uint8_t trskip_luma, trskip_chroma;            /* transform_skip_model_luma / _chroma */
  uint8_t reserved;
} kvzo_rdoq_ctx;                                 /* 148 bytes */

/* get_coeff_cabac_cost (rdo.c:158-197): bits kvz_encode_coeff_nxn (encode_coding_tree.c:103-345) spends on the
 * block in counting mode, starting from the coder range `range` (state->cabac.range) and the context states of
 * `ctx` (not modified).  0 for an all-zero block. */
uint32_t kvzo_coeff_cabac_bits(const kvzo_rdoq_ctx *ctx, uint32_t range, const int16_t *coeff, int width, int type,
                               int scan_mode, int signhide, int trskip_enable);

/* kvz_rdoq (rdo.c:548-885) + kvz_rdoq_sign_hiding (rdo.c:405-540), flat scaling list.
 * type 0 luma / 2 chroma, scan_idx 0 diag / 1 hor / 2 ver (hor/ver only for width <= 8).
 * Every entry of dest is written. */
void kvzo_rdoq(const kvzo_q_cfg *c, double lambda, const kvzo_rdoq_ctx *ctx, const int16_t *coef,
               int16_t *dest, int width, int type, int scan_idx, int tr_depth);

/* ---- pixel-typed API, generated for SFX in {8, 10} ---- */
#define KVZO_DECLARE(PIX, SFX) \
  /* A1 picture-generic.c:86-99 */ \
  unsigned kvzo_reg_sad_##SFX(const PIX *a, const PIX *b, int w, int h, unsigned sa, unsigned sb); \
  /* A2 image.c:320-479: SAD with per-pixel coordinate clamp of the reference block */ \
  unsigned kvzo_image_sad_##SFX(const PIX *pic, const PIX *ref, int W, int H, \
                                int pic_x, int pic_y, int ref_x, int ref_y, int bw, int bh); \
  /* A6 picture-generic.c:189-196,343-346,460-486 (contiguous NxN) */ \
  unsigned kvzo_sad_nxn_##SFX(int n, const PIX *a, const PIX *b); \
  unsigned kvzo_satd_nxn_##SFX(int n, const PIX *a, const PIX *b); \
  /* A7 strategies-picture.h:62-100 */ \
  unsigned kvzo_satd_any_size_##SFX(int w, int h, const PIX *a, int sa, const PIX *b, int sb); \
  /* A7 picture-generic.c:392-456 incl. the dropped 4-wide strips */ \
  void kvzo_satd_any_size_quad_##SFX(int w, int h, const PIX *const preds[4], int stride, \
                                     const PIX *orig, int orig_stride, unsigned costs[4]); \
  /* A8 image.c:488-545 */ \
  unsigned kvzo_image_satd_##SFX(const PIX *pic, const PIX *ref, int W, int H, \
                                 int pic_x, int pic_y, int ref_x, int ref_y, int bw, int bh); \
  /* A16 picture-generic.c:521-536 */ \
  unsigned kvzo_ssd_##SFX(const PIX *a, const PIX *b, int sa, int sb, int width); \
  /* A10/A11: HEVC separable interpolation of a w x h block whose top-left integer \
   * sample is (x0,y0) in the (clamped) reference plane, displaced by the \
   * quarter-pel offset (qx,qy) (any integer; >>2 = whole samples). \
   * Follows ipol-generic.c:122-157 (two passes, int16 intermediates). */ \
  void kvzo_interp_luma_##SFX(const PIX *ref, int W, int H, int x0, int y0, int qx, int qy, \
                              int w, int h, PIX *dst, int dst_stride); \
  void kvzo_interp_luma14_##SFX(const PIX *ref, int W, int H, int x0, int y0, int qx, int qy, \
                                int w, int h, int16_t *dst, int dst_stride); \
  /* chroma: eighth-pel offset, ipol-generic.c:660-728 */ \
  void kvzo_interp_chroma_##SFX(const PIX *ref, int W, int H, int x0, int y0, int ox, int oy, \
                                int w, int h, PIX *dst, int dst_stride); \
  void kvzo_interp_chroma14_##SFX(const PIX *ref, int W, int H, int x0, int y0, int ox, int oy, \
                                  int w, int h, int16_t *dst, int dst_stride); \
  /* the four candidate blocks one FME step emits (ipol-generic.c:192-658): \
   * step 0 hpel hor/ver, 1 hpel diag, 2 qpel hor/ver, 3 qpel diag around the \
   * half-pel offset (hx,hy) in {-1,0,1}; out[4][64*64], stride 64 */ \
  void kvzo_fme_blocks_##SFX(const PIX *ref, int W, int H, int x0, int y0, int w, int h, \
                             int step, int hx, int hy, PIX *out); \
  /* A12 picture-generic.c:538-588 on flat arrays */ \
  void kvzo_bipred_blend_##SFX(const int16_t *s0, const int16_t *s1, int n, PIX *dst); \
  /* A3 integer search (search_inter.c:282-956) and search_frac (:965-1128) */ \
  void kvzo_ime_##SFX(kvzo_me_job *j, const PIX *pic, const PIX *ref); \
  void kvzo_fme_##SFX(kvzo_me_job *j, const PIX *pic, const PIX *ref); \
  void kvzo_int_satd_##SFX(kvzo_me_job *j, const PIX *pic, const PIX *ref); \
  /* A14 quant-generic.c:180-273 without RDOQ, flat lists; returns has_coeffs */ \
  int kvzo_quantize_residual_##SFX(const kvzo_q_cfg *c, int width, int color, int scan_idx, \
                                   int use_trskip, int in_stride, int out_stride, \
                                   const PIX *ref_in, const PIX *pred_in, PIX *rec_out, \
                                   int16_t *coeff_out); \
  /* the same with cfg.rdoq_enable (ctx != NULL): kvz_rdoq replaces kvz_quant, except for 4x4 under rdoq_skip */ \
  int kvzo_quantize_residual_rdoq_##SFX(const kvzo_q_cfg *c, double lambda, const kvzo_rdoq_ctx *ctx, \
                                        int rdoq_skip, int tr_depth, int width, int color, int scan_idx, \
                                        int use_trskip, int in_stride, int out_stride, \
                                        const PIX *ref_in, const PIX *pred_in, PIX *rec_out, \
                                        int16_t *coeff_out); \
  /* picture hash checksum (strategies/generic/nal-generic.c:58-82), the 32-bit sum before it is unpacked \
   * big-endian; stride in samples.  At KVZ_BIT_DEPTH > 8 both bytes of a sample are hashed. */ \
  uint32_t kvzo_array_checksum_##SFX(const PIX *data, int height, int width, int stride); \
  /* the per-plane sum of compute_psnr (encmain.c:101-131) */ \
  uint64_t kvzo_plane_sse_##SFX(const PIX *src, const PIX *rec, int width, int height, int src_stride, int rec_stride); \
  /* SAO strategies (strategies/generic/sao-generic.c:50-182) and calc_sao_bands (sao.c:245-260) on contiguous \
   * block_width x block_height blocks; kvzo_sao_reconstruct = kvz_sao_reconstruct (sao.c:277-337) on whole planes \
   * of stride == frame width: `pre` holds the deblocked samples, `out` the filtered ones */ \
  int kvzo_sao_edge_ddistortion_##SFX(const PIX *orig, const PIX *rec, int bw, int bh, int eo_class, const int *offsets); \
  void kvzo_calc_sao_edge_dir_##SFX(const PIX *orig, const PIX *rec, int eo_class, int bw, int bh, int *cat_sum_cnt); \
  int kvzo_sao_band_ddistortion_##SFX(const PIX *orig, const PIX *rec, int bw, int bh, int band_pos, const int *sao_bands); \
  void kvzo_calc_sao_bands_##SFX(const PIX *orig, const PIX *rec, int bw, int bh, int *sao_bands); \
  void kvzo_sao_reconstruct_color_##SFX(const PIX *rec, PIX *new_rec, const kvzo_sao_info *sao, int stride, \
                                        int new_stride, int bw, int bh, int color); \
  void kvzo_sao_reconstruct_##SFX(const PIX *pre, PIX *out, int frame_w, int frame_h, int x, int y, int w, int h, \
                                  const kvzo_sao_info *sao, int color); \
  /* kvz_filter_deblock_lcu (filter.c:770-779) for every LCU of the picture in raster order, in place; \
   * luma stride == width, chroma stride == width / 2 */ \
  void kvzo_deblock_picture_##SFX(const kvzo_deblock_cfg *c, const kvzo_scu *scus, PIX *y, PIX *u, PIX *v); \
  /* test helper: SAD of one PU over the whole zero-centred window */ \
  void kvzo_dense_sad_map_##SFX(const PIX *pic, const PIX *ref, int W, int H, int x, int y, \
                                int bw, int bh, int r, uint32_t *out); \
  /* A17 intra.c:281-332 + intra-generic.c:37-189 + intra.c:195-279. \
   * ref_top/ref_left: 2*w+1 samples, [0] = top-left corner */ \
  void kvzo_intra_predict_##SFX(const PIX *ref_top, const PIX *ref_left, int log2_width, \
                                int mode, int is_chroma, int filter_boundary, PIX *dst);

KVZO_DECLARE(uint8_t, 8)
KVZO_DECLARE(uint16_t, 10)

#ifdef __cplusplus
}
#endif
#endif
