/*
 * include/kvz_cuda_strategies.h -- PER-CALL face of libkvz_cuda.so: one function per strategy
 * pointer of the reference's hot path, with exactly the reference's typedef signature, plus the
 * registration entry points the reference's strategy tables call.
 *
 * 8-bit build of the reference only (kvz_pixel == uint8_t, KVZ_BIT_DEPTH 8, src/kvazaar.h:72-77):
 * at another bit depth kvz_strategy_register_*_cuda() registers nothing and returns 1, the way the
 * AVX2 strategies do (src/strategies/avx2/picture-avx2.c:1230).
 *
 * Every function takes and returns HOST buffers; see INTEGRATION.md for the glue a maintainer adds
 * to src/strategies/strategies-*.c.  Cost functions cannot report errors (the reference's
 * signatures have no error channel): a CUDA failure aborts the process -- there is no CPU fallback.
 */
#ifndef KVZ_CUDA_STRATEGIES_H_
#define KVZ_CUDA_STRATEGIES_H_

#include <stddef.h>
#include <stdint.h>
#include "kvz_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef uint8_t kvz_cuda_px;      /* kvz_pixel of the 8-bit build */
typedef int16_t kvz_cuda_coeff;   /* coeff_t (src/global.h:99) */

/* ---- registration (replaces nothing; added next to kvz_strategy_register_*_avx2, e.g.
 * src/strategies/strategies-picture.c:66-88).  Same form as the reference's per-ISA registration functions:
 * int kvz_strategy_register_<family>_cuda(void *opaque, uint8_t bitdepth).  Strategy name "cuda", priority 50.
 * The host's kvz_strategyselector_register (src/strategyselector.c:216-256) is looked up in the running process
 * with dlsym(RTLD_DEFAULT, ...), which keeps this library free of link-time dependencies on the encoder; the
 * *_with forms take it as an argument instead (hosts that hide the symbol, tests).
 * Returns 1 on success; at bitdepth != 8 registers nothing and returns 1 (as avx2/picture-avx2.c:1230); without a
 * usable B200 returns 0, which fails kvz_strategyselector_init and encoder_open -- there is no CPU fallback.
 * The device is KVZ_CUDA_DEVICE, else LOCAL_RANK, else 0, unless kvz_cuda_init() already chose one. ---- */
typedef int (*kvz_cuda_register_fn)(void *opaque, const char *type, const char *strategy_name, int priority, void *fptr);
KVZ_CUDA_API int kvz_strategy_register_picture_cuda(void *opaque, uint8_t bitdepth);
KVZ_CUDA_API int kvz_strategy_register_dct_cuda(void *opaque, uint8_t bitdepth);
KVZ_CUDA_API int kvz_strategy_register_quant_cuda(void *opaque, uint8_t bitdepth);
KVZ_CUDA_API int kvz_strategy_register_intra_cuda(void *opaque, uint8_t bitdepth);
KVZ_CUDA_API int kvz_strategy_register_ipol_cuda(void *opaque, uint8_t bitdepth);
KVZ_CUDA_API int kvz_strategy_register_nal_cuda(void *opaque, uint8_t bitdepth);
KVZ_CUDA_API int kvz_strategy_register_sao_cuda(void *opaque, uint8_t bitdepth);
KVZ_CUDA_API int kvz_strategy_register_picture_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn);
KVZ_CUDA_API int kvz_strategy_register_dct_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn);
KVZ_CUDA_API int kvz_strategy_register_quant_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn);
KVZ_CUDA_API int kvz_strategy_register_intra_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn);
KVZ_CUDA_API int kvz_strategy_register_ipol_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn);
KVZ_CUDA_API int kvz_strategy_register_nal_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn);
KVZ_CUDA_API int kvz_strategy_register_sao_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn);

/* calls this face has served since the library was loaded, per strategy family: [0] picture, [1] dct, [2] quant,
 * [3] intra, [4] ipol, [5] sao, [6] nal -- parity runs read them to prove which implementation was the one called */
KVZ_CUDA_API void kvz_cuda_percall_counts(unsigned long long counts[7]);

/* ---- picture: strategies-picture.h:102-136 ---- */
/* reg_sad_func */
KVZ_CUDA_API unsigned kvz_cuda_reg_sad(const kvz_cuda_px *data1, const kvz_cuda_px *data2, int width, int height,
                                       unsigned stride1, unsigned stride2);
/* cost_pixel_nxn_func x 10 */
KVZ_CUDA_API unsigned kvz_cuda_sad_4x4(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
KVZ_CUDA_API unsigned kvz_cuda_sad_8x8(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
KVZ_CUDA_API unsigned kvz_cuda_sad_16x16(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
KVZ_CUDA_API unsigned kvz_cuda_sad_32x32(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
KVZ_CUDA_API unsigned kvz_cuda_sad_64x64(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
KVZ_CUDA_API unsigned kvz_cuda_satd_4x4(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
KVZ_CUDA_API unsigned kvz_cuda_satd_8x8(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
KVZ_CUDA_API unsigned kvz_cuda_satd_16x16(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
KVZ_CUDA_API unsigned kvz_cuda_satd_32x32(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
KVZ_CUDA_API unsigned kvz_cuda_satd_64x64(const kvz_cuda_px *block1, const kvz_cuda_px *block2);
/* cost_pixel_any_size_func */
KVZ_CUDA_API unsigned kvz_cuda_satd_any_size(int width, int height, const kvz_cuda_px *block1, int stride1,
                                             const kvz_cuda_px *block2, int stride2);
/* cost_pixel_nxn_multi_func x 10: pred_buffer is kvz_pixel (*)[32 * 32] (strategies-picture.h:30) */
typedef kvz_cuda_px (*kvz_cuda_pred_buffer)[32 * 32];
KVZ_CUDA_API void kvz_cuda_sad_4x4_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
KVZ_CUDA_API void kvz_cuda_sad_8x8_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
KVZ_CUDA_API void kvz_cuda_sad_16x16_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
KVZ_CUDA_API void kvz_cuda_sad_32x32_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
KVZ_CUDA_API void kvz_cuda_sad_64x64_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
KVZ_CUDA_API void kvz_cuda_satd_4x4_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
KVZ_CUDA_API void kvz_cuda_satd_8x8_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
KVZ_CUDA_API void kvz_cuda_satd_16x16_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
KVZ_CUDA_API void kvz_cuda_satd_32x32_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
KVZ_CUDA_API void kvz_cuda_satd_64x64_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig, unsigned num_modes, unsigned *costs_out);
/* cost_pixel_any_size_multi_func */
KVZ_CUDA_API void kvz_cuda_satd_any_size_quad(int width, int height, const kvz_cuda_px **preds, const int stride,
                                              const kvz_cuda_px *orig, const int orig_stride, unsigned num_modes,
                                              unsigned *costs_out, int8_t *valid);
/* pixels_calc_ssd_func */
KVZ_CUDA_API unsigned kvz_cuda_pixels_calc_ssd(const kvz_cuda_px *ref, const kvz_cuda_px *rec, const int ref_stride,
                                               const int rec_stride, const int width);

/* ---- dct: dct_func (strategies-dct.h:32) x 10 ---- */
KVZ_CUDA_API unsigned kvz_cuda_fast_forward_dst_4x4(int8_t bitdepth, const int16_t *input, int16_t *output);
KVZ_CUDA_API unsigned kvz_cuda_dct_4x4(int8_t bitdepth, const int16_t *input, int16_t *output);
KVZ_CUDA_API unsigned kvz_cuda_dct_8x8(int8_t bitdepth, const int16_t *input, int16_t *output);
KVZ_CUDA_API unsigned kvz_cuda_dct_16x16(int8_t bitdepth, const int16_t *input, int16_t *output);
KVZ_CUDA_API unsigned kvz_cuda_dct_32x32(int8_t bitdepth, const int16_t *input, int16_t *output);
KVZ_CUDA_API unsigned kvz_cuda_fast_inverse_dst_4x4(int8_t bitdepth, const int16_t *input, int16_t *output);
KVZ_CUDA_API unsigned kvz_cuda_idct_4x4(int8_t bitdepth, const int16_t *input, int16_t *output);
KVZ_CUDA_API unsigned kvz_cuda_idct_8x8(int8_t bitdepth, const int16_t *input, int16_t *output);
KVZ_CUDA_API unsigned kvz_cuda_idct_16x16(int8_t bitdepth, const int16_t *input, int16_t *output);
KVZ_CUDA_API unsigned kvz_cuda_idct_32x32(int8_t bitdepth, const int16_t *input, int16_t *output);

/* ---- quant: coeff_abs_sum_func (strategies-quant.h:48).  quant / dequant / quantize_residual take
 * encoder_state_t*: their glue extracts kvz_cuda_quant_cfg and calls the *_host functions below. ---- */
KVZ_CUDA_API uint32_t kvz_cuda_coeff_abs_sum(const kvz_cuda_coeff *coeffs, size_t length);
KVZ_CUDA_API int kvz_cuda_quant_host(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_coeff *coef, kvz_cuda_coeff *q_coef,
                                     int32_t width, int32_t height, int8_t type, int8_t scan_idx);
KVZ_CUDA_API int kvz_cuda_dequant_host(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_coeff *q_coef, kvz_cuda_coeff *coef,
                                       int32_t width, int32_t height, int8_t type);
KVZ_CUDA_API int kvz_cuda_quantize_residual_host(const kvz_cuda_quant_cfg *cfg, int width, int color, int scan_order,
                                                 int use_trskip, int in_stride, int out_stride,
                                                 const kvz_cuda_px *ref_in, const kvz_cuda_px *pred_in,
                                                 kvz_cuda_px *rec_out, kvz_cuda_coeff *coeff_out);

/* the same with cfg.rdoq_enable (quant-generic.c:211-221: kvz_rdoq instead of kvz_quant unless width == 4 and
 * cfg.rdoq_skip): ctx = the CABAC context states kvz_rdoq reads from state->cabac.ctx, lambda = state->lambda,
 * tr_depth = cur_cu->tr_depth - cur_cu->depth + (part_size == SIZE_NxN) */
KVZ_CUDA_API int kvz_cuda_quantize_residual_rdoq_host(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_rdoq_ctx *ctx, double lambda,
                                                      int tr_depth, int rdoq_skip, int width, int color, int scan_order,
                                                      int use_trskip, int in_stride, int out_stride,
                                                      const kvz_cuda_px *ref_in, const kvz_cuda_px *pred_in,
                                                      kvz_cuda_px *rec_out, kvz_cuda_coeff *coeff_out);
/* kvz_rdoq (rdo.c:548-885) for one block of HOST coefficients; type 0 luma / 2 chroma */
KVZ_CUDA_API int kvz_cuda_rdoq_host(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_rdoq_ctx *ctx, double lambda,
                                    const kvz_cuda_coeff *coef, kvz_cuda_coeff *dest_coeff, int width, int type,
                                    int scan_idx, int tr_depth);

/* ---- intra: angular_pred_func, intra_pred_planar_func (strategies-intra.h:32-43) ---- */
KVZ_CUDA_API void kvz_cuda_angular_pred(const int_fast8_t log2_width, const int_fast8_t intra_mode,
                                        const kvz_cuda_px *const in_ref_above, const kvz_cuda_px *const in_ref_left,
                                        kvz_cuda_px *const dst);
KVZ_CUDA_API void kvz_cuda_intra_pred_planar(const int_fast8_t log2_width, const kvz_cuda_px *const ref_top,
                                             const kvz_cuda_px *const ref_left, kvz_cuda_px *const dst);

/* ---- ipol: kvz_sample_*_func and ipol_blocks_func (strategies-ipol.h:36-48); `encoder` is not read ---- */
KVZ_CUDA_API void kvz_cuda_sample_quarterpel_luma(const void *encoder, kvz_cuda_px *src, int16_t src_stride, int width,
                                                  int height, kvz_cuda_px *dst, int16_t dst_stride, int8_t hor_flag,
                                                  int8_t ver_flag, const int16_t mv[2]);
KVZ_CUDA_API void kvz_cuda_sample_octpel_chroma(const void *encoder, kvz_cuda_px *src, int16_t src_stride, int width,
                                                int height, kvz_cuda_px *dst, int16_t dst_stride, int8_t hor_flag,
                                                int8_t ver_flag, const int16_t mv[2]);
KVZ_CUDA_API void kvz_cuda_sample_14bit_quarterpel_luma(const void *encoder, kvz_cuda_px *src, int16_t src_stride,
                                                        int width, int height, int16_t *dst, int16_t dst_stride,
                                                        int8_t hor_flag, int8_t ver_flag, const int16_t mv[2]);
KVZ_CUDA_API void kvz_cuda_sample_14bit_octpel_chroma(const void *encoder, kvz_cuda_px *src, int16_t src_stride,
                                                      int width, int height, int16_t *dst, int16_t dst_stride,
                                                      int8_t hor_flag, int8_t ver_flag, const int16_t mv[2]);
/* the four FME block filters; filtered[4][64 * 64] receives the candidate blocks, the intermediate
 * buffers are left untouched (every call filters from src) */
#define KVZ_CUDA_IPOL_BLOCKS_ARGS const void *encoder, kvz_cuda_px *src, int16_t src_stride, int width, int height, \
  kvz_cuda_px filtered[4][64 * 64], int16_t hor_intermediate[5][72 * 64], int8_t fme_level, \
  int16_t hor_first_cols[5][72], int8_t sample_off_x, int8_t sample_off_y
KVZ_CUDA_API void kvz_cuda_filter_hpel_blocks_hor_ver_luma(KVZ_CUDA_IPOL_BLOCKS_ARGS);
KVZ_CUDA_API void kvz_cuda_filter_hpel_blocks_diag_luma(KVZ_CUDA_IPOL_BLOCKS_ARGS);
KVZ_CUDA_API void kvz_cuda_filter_qpel_blocks_hor_ver_luma(KVZ_CUDA_IPOL_BLOCKS_ARGS);
KVZ_CUDA_API void kvz_cuda_filter_qpel_blocks_diag_luma(KVZ_CUDA_IPOL_BLOCKS_ARGS);

/* ---- nal: strategies-nal.h:42-47 ---- */
/* array_checksum_func kvz_array_checksum (checksum_out is SEI_HASH_MAX_LENGTH = 16 bytes, the first four are
 * written, nal-generic.c:76-81); the plane is uploaded in strips of rows.  kvz_array_md5 is not replaced. */
KVZ_CUDA_API void kvz_cuda_array_checksum(const kvz_cuda_px *data, const int height, const int width, const int stride,
                                          unsigned char checksum_out[16], const uint8_t bitdepth);

/* ---- sao: strategies-sao.h:37-60 ---- */
/* sao_edge_ddistortion_func, calc_sao_edge_dir_func: contiguous block_width x block_height blocks, at most 64x64
 * (NUM_SAO_EDGE_CATEGORIES = 5; cat_sum_cnt is accumulated into, as sao-generic.c:110-111 does) */
KVZ_CUDA_API int kvz_cuda_sao_edge_ddistortion(const kvz_cuda_px *orig_data, const kvz_cuda_px *rec_data, int block_width,
                                               int block_height, int eo_class, int offsets[5]);
KVZ_CUDA_API void kvz_cuda_calc_sao_edge_dir(const kvz_cuda_px *orig_data, const kvz_cuda_px *rec_data, int eo_class,
                                             int block_width, int block_height, int cat_sum_cnt[2][5]);
/* sao_band_ddistortion_func / sao_reconstruct_color_func take encoder_state_t / encoder_control_t / sao_info_t:
 * the glue's wrappers pass encoder_control->bitdepth and the sao_info_t fields packed as kvz_cuda_sao_params */
KVZ_CUDA_API int kvz_cuda_sao_band_ddistortion_host(int bitdepth, const kvz_cuda_px *orig_data, const kvz_cuda_px *rec_data,
                                                    int block_width, int block_height, int band_pos, const int sao_bands[4]);
KVZ_CUDA_API void kvz_cuda_sao_reconstruct_color_host(int bitdepth, const kvz_cuda_px *rec_data, kvz_cuda_px *new_rec_data,
                                                      const kvz_cuda_sao_params *sao, int stride, int new_stride,
                                                      int block_width, int block_height, int color_i);

#ifdef __cplusplus
}
#endif
#endif /* KVZ_CUDA_STRATEGIES_H_ */
