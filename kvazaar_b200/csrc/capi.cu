// kvazaar_b200/csrc/capi.cu -- extern "C" entry points of libkvz_cuda.so
// (declared in include/kvz_cuda.h; plumbing + the batch integer-ME face).
#include "../../include/kvz_cuda.h"
#include "common.cuh"
#include "ime_sad.cuh"
#include "me_search.cuh"

namespace kvz {
thread_local char g_last_error[512] = "";
unsigned long long g_launch_count = 0;
int g_device = -1;          // the device kvz_cuda_init() selected; read by every per-call function
}  // namespace kvz

namespace kvz {
int launch_block_cost(const void *base_a, const void *base_b, const kvz_cuda_block_job *jobs, int n_jobs,
                      int bitdepth, uint32_t *out, cudaStream_t stream);
int launch_image_cost(const void *pic, int pic_pitch, const void *ref, int ref_pitch, int width, int height,
                      const kvz_cuda_image_cost_job *jobs, int n_jobs, int bitdepth, uint32_t *out, cudaStream_t stream);
}
namespace kvz {
int launch_picture_checksum(const kvz_cuda_plane *planes, int n_planes, int bitdepth, uint32_t *d_out, cudaStream_t stream, int y0);
int launch_picture_sse(const kvz_cuda_plane *src, const kvz_cuda_plane *rec, int n_planes, int bitdepth,
                       unsigned long long *d_out, cudaStream_t stream);
}

namespace kvz {
int launch_sao_stats(const kvz_cuda_plane *orig, const kvz_cuda_plane *rec, int n_planes, int bitdepth,
                     kvz_cuda_sao_stats *d_stats, cudaStream_t stream);
int launch_sao_apply(const kvz_cuda_plane *rec_in, const kvz_cuda_plane *rec_out, int n_planes, int bitdepth,
                     const kvz_cuda_sao_params *d_luma, const kvz_cuda_sao_params *d_chroma, cudaStream_t stream);
}

namespace kvz {
int launch_deblock_picture(const kvz_cuda_deblock_cfg &cfg, const kvz_cuda_scu *d_scus, const kvz_cuda_plane *planes,
                           int n_planes, int bitdepth, cudaStream_t stream);
}

namespace kvz {
int launch_coeff_bits(const kvz_cuda_rdoq_ctx *d_ctx, const int16_t *d_coeffs, const kvz_cuda_coeff_job *d_jobs,
                      int n_jobs, int max_width, int signhide, int trskip_enable, uint32_t *d_bits, cudaStream_t stream);
}

namespace kvz {
int launch_inter_candidates(const kvz_cuda_cand_frame &f, kvz_cuda_me_pu *d_pus, int n_pus, const uint8_t *d_pu_flags,
                            kvz_cuda_me_job *d_jobs, cudaStream_t stream);
}

namespace kvz {
int launch_transform_batch(int n, int dst4, int inverse, int bitdepth, const int16_t *in, int16_t *out, int count,
                           cudaStream_t stream);
int launch_quant_batch(const kvz_cuda_quant_cfg &c, int n, int type, int scan_idx, int dequant, const int16_t *in,
                       int16_t *out, int count, cudaStream_t stream);
int launch_coeff_abs_sum(const int16_t *coeffs, int per_block, int count, uint32_t *out, cudaStream_t stream);
int launch_quantize_residual(const kvz_cuda_quant_cfg &c, const void *ref, const void *pred, void *rec, int16_t *coeff,
                             const kvz_cuda_residual_job *jobs, int n_jobs, uint8_t *has_coeffs, cudaStream_t stream);
int launch_rdoq_batch(const kvz_cuda_quant_cfg &cfg, const kvz_cuda_rdoq_ctx *d_ctx, const int16_t *d_coef,
                      int16_t *d_q_coef, const kvz_cuda_rdoq_job *d_jobs, int n_jobs, int max_width, int skip4,
                      cudaStream_t stream);
int launch_quantize_residual_rdoq(const kvz_cuda_quant_cfg &c, const void *ref, const void *pred, void *rec, int16_t *coeff,
                                  const kvz_cuda_residual_job *jobs, int n_jobs, uint8_t *has_coeffs,
                                  const kvz_cuda_rdoq_ctx *d_ctx, const kvz_cuda_rdoq_job *rdoq_jobs, int max_width,
                                  int16_t *work, int rdoq_skip, cudaStream_t stream);
int launch_quantize_residual_rdoq_noalloc(const kvz_cuda_quant_cfg &c, const void *ref, const void *pred, void *rec,
                                          int16_t *coeff, size_t coeff_elems, const kvz_cuda_residual_job *jobs, int n_jobs,
                                          uint8_t *has_coeffs, const kvz_cuda_rdoq_ctx *d_ctx,
                                          const kvz_cuda_rdoq_job *rdoq_jobs, int max_width, int rdoq_skip,
                                          cudaStream_t stream);
}
namespace kvz {
int launch_interp(const void *plane, int pitch, int W, int H, int bitdepth, const kvz_cuda_interp_job *jobs, int n_jobs,
                  void *dst8, int16_t *dst16, cudaStream_t stream);
int launch_bipred_blend(const int16_t *s0, const int16_t *s1, size_t n, int bitdepth, void *dst, cudaStream_t stream);
}
namespace kvz {
int launch_intra_costs(const void *refs, const void *orig, const kvz_cuda_intra_job *jobs, int n_jobs,
                       uint32_t *costs, void *preds, cudaStream_t stream, int bitdepth);
}
namespace kvz { int measure_vabsdiff4_peak(double *abs_diffs_per_s, cudaStream_t stream); }

using namespace kvz;

extern "C" {

int kvz_cuda_device_count(void)
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int kvz_cuda_init(int device)
{
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail_msg("kvz_cuda_init: no CUDA device (this library has no CPU fallback)");
  }
  if (device < 0 || device >= n) return fail_msg("kvz_cuda_init: bad device index");
  KVZ_CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  KVZ_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    snprintf(g_last_error, sizeof(g_last_error),
             "kvz_cuda_init: device is sm_%d%d, library is built for sm_100a only", prop.major, prop.minor);
    return 0;
  }
  KVZ_CUDA_TRY(cudaFree(0));
  g_device = device;
  return 1;
}

void kvz_cuda_shutdown(void) { g_device = -1; }

const char *kvz_cuda_last_error(void) { return g_last_error; }

unsigned long long kvz_cuda_launch_count(void) { return g_launch_count; }

int kvz_cuda_measure_int_pipe_peak(double *abs_diffs_per_s)
{
  if (!abs_diffs_per_s) return fail_msg("kvz_cuda_measure_int_pipe_peak: null argument");
  return measure_vabsdiff4_peak(abs_diffs_per_s, 0);
}

void *kvz_cuda_malloc(size_t bytes)
{
  void *p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) { fail("cudaMalloc", cudaGetLastError(), __FILE__, __LINE__); return nullptr; }
  return p;
}

void kvz_cuda_free(void *dptr) { if (dptr) cudaFree(dptr); }

void *kvz_cuda_malloc_host(size_t bytes)
{
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) { fail("cudaMallocHost", cudaGetLastError(), __FILE__, __LINE__); return nullptr; }
  return p;
}

void kvz_cuda_free_host(void *hptr) { if (hptr) cudaFreeHost(hptr); }

int kvz_cuda_memcpy_h2d(void *dst, const void *src, size_t bytes, void *stream)
{
  KVZ_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return 1;
}

int kvz_cuda_memcpy_d2h(void *dst, const void *src, size_t bytes, void *stream)
{
  KVZ_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return 1;
}

int kvz_cuda_memset(void *dst, int value, size_t bytes, void *stream)
{
  KVZ_CUDA_TRY(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream));
  return 1;
}

int kvz_cuda_stream_sync(void *stream)
{
  KVZ_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  return 1;
}

int kvz_cuda_pad_plane(const void *d_src, int w, int h, int src_pitch, void *d_dst, int pad,
                       int dst_pitch, int bitdepth, void *stream)
{
  if (!d_src || !d_dst || w <= 0 || h <= 0 || pad < 0) return fail_msg("kvz_cuda_pad_plane: bad argument");
  return launch_pad_plane(d_src, w, h, src_pitch, d_dst, pad, dst_pitch, bitdepth, (cudaStream_t)stream);
}

int kvz_cuda_ime_required_pad(int range) { return (range + 96 + 31) & ~31; }

int kvz_cuda_ime_map_pitch(int range) { return ime_map_pitch(range); }

size_t kvz_cuda_ime_map_bytes(const kvz_cuda_ime_desc *d, int level)
{
  if (!d || level < 0 || level > 3) return 0;
  const size_t ctus = (size_t)((d->tile_w + 63) / 64) * ((d->tile_h + 63) / 64);
  const size_t n = 8 >> level;
  const size_t k = 2 * d->range + 1;
  return ctus * n * n * (k + 1) * ime_map_pitch(d->range) * (level < 2 ? 2 : 4);   // row K = the column dx = +R
}

int kvz_cuda_ime_sad_maps(const kvz_cuda_ime_desc *d, const void *d_cur, const void *d_ref_padded,
                          uint16_t *d_sad8, uint16_t *d_sad16, uint32_t *d_sad32, uint32_t *d_sad64,
                          void *stream)
{
  if (!d || !d_cur || !d_ref_padded || !d_sad8 || !d_sad16 || !d_sad32 || !d_sad64)
    return fail_msg("kvz_cuda_ime_sad_maps: null argument");
  if (d->bitdepth != 8) return fail_msg("kvz_cuda_ime_sad_maps: only bitdepth 8 is implemented");
  {
    const int ac = abs(d->center_x) > abs(d->center_y) ? abs(d->center_x) : abs(d->center_y);
    if (d->ref_pad < kvz_cuda_ime_required_pad(d->range) + ac) return fail_msg("kvz_cuda_ime_sad_maps: ref_pad too small for this range and centre");
  }
  if ((d->width & 7) || (d->cur_pitch & 7) || (d->ref_pitch & 3) || (d->ref_pad & 3) ||
      (d->tile_x & 63) || (d->tile_y & 63) || (d->tile_w & 7) || (d->tile_h & 7))
    return fail_msg("kvz_cuda_ime_sad_maps: width/pitches/pad must be multiples of 8, tile origin of 64");
  ImeSadParams p;
  p.cur = (const uint8_t *)d_cur;
  p.ref = (const uint8_t *)d_ref_padded;
  p.width = d->width; p.height = d->height;
  p.cur_pitch = d->cur_pitch; p.ref_pitch = d->ref_pitch; p.pad = d->ref_pad;
  p.tile_x = d->tile_x; p.tile_y = d->tile_y; p.tile_w = d->tile_w; p.tile_h = d->tile_h;
  p.ctus_x = (d->tile_w + 63) / 64; p.ctus_y = (d->tile_h + 63) / 64;
  p.map_pitch = ime_map_pitch(d->range);
  p.sad8 = d_sad8; p.sad16 = d_sad16; p.sad32 = d_sad32; p.sad64 = d_sad64;
  return launch_ime_sad_dense(p, d->range, d->center_x, d->center_y, (cudaStream_t)stream);
}

size_t kvz_cuda_ime_box_map_bytes(const kvz_cuda_ime_box_desc *d, int level)
{
  if (!d || level < 0 || level > 3) return 0;
  const size_t ctus = (size_t)((d->tile_w + 63) / 64) * ((d->tile_h + 63) / 64);
  const size_t n = 8 >> level;
  return ctus * n * n * (size_t)d->n_dy * (size_t)d->n_dx * (level < 2 ? 2 : 4);
}

// validates one box descriptor and fills the kernel parameters
static int box_params(const kvz_cuda_ime_box_desc *d, const void *d_cur, const void *d_ref_padded, uint16_t *d_sad8,
                      uint16_t *d_sad16, uint32_t *d_sad32, uint32_t *d_sad64, ImeSadParams &p)
{
  if (!d || !d_cur || !d_ref_padded || !d_sad8 || !d_sad16 || !d_sad32 || !d_sad64)
    return fail_msg("kvz_cuda_ime_sad_box_maps: null argument");
  if (d->bitdepth != 8 && d->bitdepth != 10) return fail_msg("kvz_cuda_ime_sad_box_maps: bitdepth must be 8 or 10");
  if (d->n_dx <= 0 || (d->n_dx & 7) || d->n_dy <= 0 || d->n_dx > kMeMaxBox || d->n_dy > kMeMaxBox)
    return fail_msg("kvz_cuda_ime_sad_box_maps: n_dx must be a multiple of 8; both sides 1..272");
  {
    int far = abs(d->x0);
    if (abs(d->x0 + d->n_dx) > far) far = abs(d->x0 + d->n_dx);
    if (abs(d->y0) > far) far = abs(d->y0);
    if (abs(d->y0 + d->n_dy) > far) far = abs(d->y0 + d->n_dy);
    if (d->ref_pad < 104 + far) return fail_msg("kvz_cuda_ime_sad_box_maps: ref_pad too small for this box");
  }
  if ((d->width & 7) || (d->cur_pitch & 7) || (d->ref_pitch & 3) || (d->ref_pad & 3) ||
      (d->tile_x & 63) || (d->tile_y & 63) || (d->tile_w & 7) || (d->tile_h & 7))
    return fail_msg("kvz_cuda_ime_sad_box_maps: width/pitches/pad must be multiples of 8, tile origin of 64");
  p.cur = (const uint8_t *)d_cur;
  p.ref = (const uint8_t *)d_ref_padded;
  p.width = d->width; p.height = d->height;
  p.cur_pitch = d->cur_pitch; p.ref_pitch = d->ref_pitch; p.pad = d->ref_pad;
  p.tile_x = d->tile_x; p.tile_y = d->tile_y; p.tile_w = d->tile_w; p.tile_h = d->tile_h;
  p.ctus_x = (d->tile_w + 63) / 64; p.ctus_y = (d->tile_h + 63) / 64;
  p.map_pitch = d->n_dx;
  p.sad8 = d_sad8; p.sad16 = d_sad16; p.sad32 = d_sad32; p.sad64 = d_sad64;
  return 1;
}

int kvz_cuda_ime_sad_box_maps(const kvz_cuda_ime_box_desc *d, const void *d_cur, const void *d_ref_padded,
                              uint16_t *d_sad8, uint16_t *d_sad16, uint32_t *d_sad32, uint32_t *d_sad64,
                              void *stream)
{
  ImeSadParams p;
  if (!box_params(d, d_cur, d_ref_padded, d_sad8, d_sad16, d_sad32, d_sad64, p)) return 0;
  if (d->bitdepth != 8) return launch_ime_sad10_box(p, d->x0, d->y0, d->n_dx, d->n_dy, d->bitdepth, (cudaStream_t)stream);
  return launch_ime_sad_box(p, d->x0, d->y0, d->n_dx, d->n_dy, (cudaStream_t)stream);
}

int kvz_cuda_ime_sad_box_maps_multi(const kvz_cuda_ime_box_ref *refs, int n_refs, const void *d_cur, void *stream)
{
  if (!refs || n_refs < 1 || n_refs > KVZ_CUDA_IME_MAX_MULTI)
    return fail_msg("kvz_cuda_ime_sad_box_maps_multi: 1..8 references per call");
  ImeSadParams ps[KVZ_CUDA_IME_MAX_MULTI];
  int boxes[KVZ_CUDA_IME_MAX_MULTI][4];
  for (int r = 0; r < n_refs; r++) {
    const kvz_cuda_ime_box_ref &b = refs[r];
    if (!box_params(&b.desc, d_cur, b.d_ref_padded, b.d_sad8, b.d_sad16, b.d_sad32, b.d_sad64, ps[r])) return 0;
    boxes[r][0] = b.desc.x0; boxes[r][1] = b.desc.y0; boxes[r][2] = b.desc.n_dx; boxes[r][3] = b.desc.n_dy;
  }
  if (refs[0].desc.bitdepth != 8) {                 // 10-bit: one launch per reference (ime_sad10.cu)
    for (int r = 0; r < n_refs; r++)
      if (!launch_ime_sad10_box(ps[r], boxes[r][0], boxes[r][1], boxes[r][2], boxes[r][3], refs[r].desc.bitdepth,
                                (cudaStream_t)stream)) return 0;
    return 1;
  }
  return launch_ime_sad_box_multi(ps, boxes, n_refs, (cudaStream_t)stream);
}

static int me_search_common(const kvz_cuda_me_frame *fr, const kvz_cuda_me_pu *d_pus, int n_pus,
                            const kvz_cuda_me_job *d_jobs, kvz_cuda_me_result *d_results, void *stream, bool fused)
{
  if (!fr || !d_pus || !d_jobs || !d_results || !fr->d_cur) return fail_msg("kvz_cuda_me_search: null argument");
  if (fr->bitdepth != 8) return fail_msg("kvz_cuda_me_search: only bitdepth 8 is implemented");
  if (fr->range != 8 && fr->range != 16 && fr->range != 32 && fr->range != 64)
    return fail_msg("kvz_cuda_me_search: range must be 8, 16, 32 or 64");
  if (fr->n_refs < 1 || fr->n_refs > KVZ_CUDA_MAX_REFS) return fail_msg("kvz_cuda_me_search: bad n_refs");
  if (fr->fme_level < 0 || fr->fme_level > 4) return fail_msg("kvz_cuda_me_search: fme_level must be 0..4");
  MeFrame f;
  memset(&f, 0, sizeof(f));
  f.cur = (const uint8_t *)fr->d_cur; f.cur_pitch = fr->cur_pitch;
  f.width = fr->width; f.height = fr->height;
  f.tile_x = fr->tile_x; f.tile_y = fr->tile_y; f.tile_w = fr->tile_w; f.tile_h = fr->tile_h;
  f.ctus_x = (fr->tile_w + 63) / 64;
  f.range = fr->range; f.map_pitch = ime_map_pitch(fr->range);
  if (fr->ime_search < 0 || fr->ime_search > KVZ_CUDA_IME_TZ) return fail_msg("kvz_cuda_me_search: ime_search must be 0 (full), 1 (hexbs), 2 (dia) or 3 (tz)");
  f.ime_search = fr->ime_search; f.me_early_termination = fr->me_early_termination; f.me_max_steps = fr->me_max_steps;
  f.lambda_sqrt_i = fr->lambda_sqrt_i;
  f.fme_level = fr->fme_level; f.owf = fr->owf; f.wpp = fr->wpp; f.sao_type = fr->sao_type;
  f.deblock_enable = fr->deblock_enable; f.mv_constraint = fr->mv_constraint;
  f.max_ref_lcu_right = fr->max_ref_lcu_right; f.max_ref_lcu_down = fr->max_ref_lcu_down;
  f.n_refs = fr->n_refs;
  memcpy(f.ref_LX, fr->ref_LX, sizeof(f.ref_LX));
  for (int i = 0; i < fr->n_refs; i++) {
    const kvz_cuda_me_ref &r = fr->refs[i];
    // the refinement stages a (w + 20) x (h + 9) window by clamping its ORIGIN into the border: that equals the
    // per-sample clamp of kvz_get_extended_block (ipol-generic.c:761-781) only with a border of at least 64 + 20
    if (!r.d_ref_padded || r.ref_pad < 84 || (r.ref_pitch & 3))
      return fail_msg("kvz_cuda_me_search: bad reference plane (border must be at least 84 samples, pitch a multiple of 4)");
    f.refs[i].plane = (const uint8_t *)r.d_ref_padded; f.refs[i].pitch = r.ref_pitch; f.refs[i].pad = r.ref_pad;
    if (r.n_mapsets < 0 || r.n_mapsets > KVZ_CUDA_MAX_MAPSETS) return fail_msg("kvz_cuda_me_search: bad n_mapsets");
    f.refs[i].n_sets = r.n_mapsets;
    for (int m = 0; m < r.n_mapsets; m++) {
      f.refs[i].sets[m].cx = r.mapsets[m].center_x; f.refs[i].sets[m].cy = r.mapsets[m].center_y;
      f.refs[i].sets[m].n_dx = r.mapsets[m].n_dx; f.refs[i].sets[m].n_dy = r.mapsets[m].n_dy;
      if (r.mapsets[m].n_dx < 0 || (r.mapsets[m].n_dx & 7) || r.mapsets[m].n_dx > kMeMaxBox || r.mapsets[m].n_dy < 0 ||
          r.mapsets[m].n_dy > kMeMaxBox || (r.mapsets[m].n_dx > 0) != (r.mapsets[m].n_dy > 0))
        return fail_msg("kvz_cuda_me_search: bad box map set");
      f.refs[i].sets[m].sad8 = r.mapsets[m].d_sad8; f.refs[i].sets[m].sad16 = r.mapsets[m].d_sad16;
      f.refs[i].sets[m].sad32 = r.mapsets[m].d_sad32; f.refs[i].sets[m].sad64 = r.mapsets[m].d_sad64;
    }
  }
  if (fused) return launch_me_search_pu(f, d_pus, d_jobs, d_results, n_pus, (cudaStream_t)stream);
  return launch_me_search(f, d_pus, d_jobs, d_results, n_pus, (cudaStream_t)stream);
}

int kvz_cuda_me_search(const kvz_cuda_me_frame *fr, const kvz_cuda_me_pu *d_pus, int n_pus,
                       const kvz_cuda_me_job *d_jobs, kvz_cuda_me_result *d_results, void *stream)
{
  return me_search_common(fr, d_pus, n_pus, d_jobs, d_results, stream, false);
}

int kvz_cuda_me_search_fused(const kvz_cuda_me_frame *fr, const kvz_cuda_me_pu *d_pus, int n_pus,
                             const kvz_cuda_me_job *d_jobs, kvz_cuda_me_result *d_results, void *stream)
{
  return me_search_common(fr, d_pus, n_pus, d_jobs, d_results, stream, true);
}

int kvz_cuda_block_costs(const void *d_a, const void *d_b, const kvz_cuda_block_job *d_jobs, int n_jobs,
                         int bitdepth, uint32_t *d_out, void *stream)
{
  if (!d_a || !d_b || !d_jobs || !d_out || n_jobs < 0) return fail_msg("kvz_cuda_block_costs: bad argument");
  if (bitdepth != 8 && bitdepth != 10) return fail_msg("kvz_cuda_block_costs: bitdepth must be 8 or 10");
  return launch_block_cost(d_a, d_b, d_jobs, n_jobs, bitdepth, d_out, (cudaStream_t)stream);
}

int kvz_cuda_image_costs(const void *d_pic, int pic_pitch, const void *d_ref, int ref_pitch, int width, int height,
                         const kvz_cuda_image_cost_job *d_jobs, int n_jobs, int bitdepth, uint32_t *d_out, void *stream)
{
  if (!d_pic || !d_ref || !d_jobs || !d_out || n_jobs < 0 || width <= 0 || height <= 0)
    return fail_msg("kvz_cuda_image_costs: bad argument");
  if (bitdepth != 8 && bitdepth != 10) return fail_msg("kvz_cuda_image_costs: bitdepth must be 8 or 10");
  return launch_image_cost(d_pic, pic_pitch, d_ref, ref_pitch, width, height, d_jobs, n_jobs, bitdepth, d_out,
                           (cudaStream_t)stream);
}

int kvz_cuda_picture_checksum(const kvz_cuda_plane *planes, int n_planes, int bitdepth, uint32_t *d_checksums, void *stream)
{
  if (!planes || !d_checksums) return fail_msg("kvz_cuda_picture_checksum: bad argument");
  if (bitdepth < 8 || bitdepth > 16) return fail_msg("kvz_cuda_picture_checksum: bitdepth must be 8..16");
  return launch_picture_checksum(planes, n_planes, bitdepth, d_checksums, (cudaStream_t)stream, 0);
}

int kvz_cuda_picture_sse(const kvz_cuda_plane *src, const kvz_cuda_plane *rec, int n_planes, int bitdepth,
                         uint64_t *d_sse, void *stream)
{
  if (!src || !rec || !d_sse) return fail_msg("kvz_cuda_picture_sse: bad argument");
  if (bitdepth < 8 || bitdepth > 16) return fail_msg("kvz_cuda_picture_sse: bitdepth must be 8..16");
  return launch_picture_sse(src, rec, n_planes, bitdepth, reinterpret_cast<unsigned long long *>(d_sse), (cudaStream_t)stream);
}

int kvz_cuda_sao_ctu_stats(const kvz_cuda_plane *orig, const kvz_cuda_plane *rec, int n_planes, int bitdepth,
                           kvz_cuda_sao_stats *d_stats, void *stream)
{
  if (!orig || !rec || !d_stats) return fail_msg("kvz_cuda_sao_ctu_stats: bad argument");
  if (bitdepth < 8 || bitdepth > 16) return fail_msg("kvz_cuda_sao_ctu_stats: bitdepth must be 8..16");
  return launch_sao_stats(orig, rec, n_planes, bitdepth, d_stats, (cudaStream_t)stream);
}

int kvz_cuda_sao_reconstruct(const kvz_cuda_plane *rec_in, const kvz_cuda_plane *rec_out, int n_planes, int bitdepth,
                             const kvz_cuda_sao_params *d_luma, const kvz_cuda_sao_params *d_chroma, void *stream)
{
  if (!rec_in || !rec_out || !d_luma || (n_planes > 1 && !d_chroma)) return fail_msg("kvz_cuda_sao_reconstruct: bad argument");
  if (bitdepth < 8 || bitdepth > 16) return fail_msg("kvz_cuda_sao_reconstruct: bitdepth must be 8..16");
  return launch_sao_apply(rec_in, rec_out, n_planes, bitdepth, d_luma, d_chroma, (cudaStream_t)stream);
}

int kvz_cuda_inter_candidates(const kvz_cuda_cand_frame *frame, kvz_cuda_me_pu *d_pus, int n_pus, const uint8_t *d_pu_flags,
                              kvz_cuda_me_job *d_jobs, void *stream)
{
  if (!frame || !d_pus || !d_jobs || n_pus < 0 || !frame->d_scus) return fail_msg("kvz_cuda_inter_candidates: bad argument");
  if (frame->width <= 0 || frame->height <= 0 || (frame->width & 3) || (frame->height & 3) || frame->n_refs < 0 ||
      frame->n_refs > 16 || frame->ref_LX_size[0] > 16 || frame->ref_LX_size[1] > 16 || frame->ref_scu_pitch <= 0)
    return fail_msg("kvz_cuda_inter_candidates: bad frame description");
  for (int r = 0; r < frame->n_refs; r++)
    if (!frame->d_ref_scus[r]) return fail_msg("kvz_cuda_inter_candidates: missing reference SCU array");
  return launch_inter_candidates(*frame, d_pus, n_pus, d_pu_flags, d_jobs, (cudaStream_t)stream);
}

int kvz_cuda_deblock_picture(const kvz_cuda_deblock_cfg *cfg, const kvz_cuda_scu *d_scus, const kvz_cuda_plane *planes,
                             int n_planes, int bitdepth, void *stream)
{
  if (!cfg || !d_scus || !planes || (n_planes != 1 && n_planes != 3)) return fail_msg("kvz_cuda_deblock_picture: bad argument");
  if (bitdepth < 8 || bitdepth > 16) return fail_msg("kvz_cuda_deblock_picture: bitdepth must be 8..16");
  if (cfg->width < 8 || cfg->height < 8 || (cfg->width & 7) || (cfg->height & 7))
    return fail_msg("kvz_cuda_deblock_picture: width and height must be positive multiples of 8");
  if (cfg->qp < 0 || cfg->qp > 51) return fail_msg("kvz_cuda_deblock_picture: qp must be 0..51");
  return launch_deblock_picture(*cfg, d_scus, planes, n_planes, bitdepth, (cudaStream_t)stream);
}

int kvz_cuda_transform_blocks(int n, int dst4, int inverse, int bitdepth, const int16_t *d_in, int16_t *d_out,
                              int count, void *stream)
{
  if (!d_in || !d_out || count < 0) return fail_msg("kvz_cuda_transform_blocks: bad argument");
  if (bitdepth != 8 && bitdepth != 10) return fail_msg("kvz_cuda_transform_blocks: bitdepth must be 8 or 10");
  return launch_transform_batch(n, dst4, inverse, bitdepth, d_in, d_out, count, (cudaStream_t)stream);
}

static int check_qcfg(const kvz_cuda_quant_cfg *c, const char *who)
{
  if (!c) return fail_msg(who);
  if ((c->bitdepth != 8 && c->bitdepth != 10) || c->qp < -12 || c->qp > 51) return fail_msg(who);
  return 1;
}

int kvz_cuda_quant_blocks(const kvz_cuda_quant_cfg *cfg, int n, int type, int scan_idx, const int16_t *d_coef,
                          int16_t *d_q_coef, int count, void *stream)
{
  if (!check_qcfg(cfg, "kvz_cuda_quant_blocks: bad cfg") ) return 0;
  if (!d_coef || !d_q_coef || count < 0 || scan_idx < 0 || scan_idx > 2) return fail_msg("kvz_cuda_quant_blocks: bad argument");
  return launch_quant_batch(*cfg, n, type, scan_idx, 0, d_coef, d_q_coef, count, (cudaStream_t)stream);
}

int kvz_cuda_dequant_blocks(const kvz_cuda_quant_cfg *cfg, int n, int type, const int16_t *d_q_coef, int16_t *d_coef,
                            int count, void *stream)
{
  if (!check_qcfg(cfg, "kvz_cuda_dequant_blocks: bad cfg")) return 0;
  if (!d_coef || !d_q_coef || count < 0) return fail_msg("kvz_cuda_dequant_blocks: bad argument");
  return launch_quant_batch(*cfg, n, type, 0, 1, d_q_coef, d_coef, count, (cudaStream_t)stream);
}

int kvz_cuda_rdoq_blocks(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_rdoq_ctx *d_ctx, const int16_t *d_coef,
                         int16_t *d_q_coef, const kvz_cuda_rdoq_job *d_jobs, int n_jobs, int max_width, void *stream)
{
  if (!check_qcfg(cfg, "kvz_cuda_rdoq_blocks: bad cfg")) return 0;
  if (!d_ctx || !d_coef || !d_q_coef || !d_jobs || n_jobs < 0) return fail_msg("kvz_cuda_rdoq_blocks: bad argument");
  if (max_width != 4 && max_width != 8 && max_width != 16 && max_width != 32)
    return fail_msg("kvz_cuda_rdoq_blocks: max_width must be 4, 8, 16 or 32");
  return launch_rdoq_batch(*cfg, d_ctx, d_coef, d_q_coef, d_jobs, n_jobs, max_width, 0, (cudaStream_t)stream);
}

int kvz_cuda_coeff_bits(const kvz_cuda_rdoq_ctx *d_ctx, const int16_t *d_coeffs, const kvz_cuda_coeff_job *d_jobs,
                        int n_jobs, int max_width, int signhide, int trskip_enable, uint32_t *d_bits, void *stream)
{
  if (!d_ctx || !d_coeffs || !d_jobs || !d_bits || n_jobs < 0) return fail_msg("kvz_cuda_coeff_bits: bad argument");
  if (max_width != 4 && max_width != 8 && max_width != 16 && max_width != 32)
    return fail_msg("kvz_cuda_coeff_bits: max_width must be 4, 8, 16 or 32");
  return launch_coeff_bits(d_ctx, d_coeffs, d_jobs, n_jobs, max_width, signhide != 0, trskip_enable != 0, d_bits,
                           (cudaStream_t)stream);
}

int kvz_cuda_quantize_residuals_rdoq(const kvz_cuda_quant_cfg *cfg, const void *d_ref, const void *d_pred, void *d_rec,
                                     int16_t *d_coeff, const kvz_cuda_residual_job *d_jobs, int n_jobs,
                                     uint8_t *d_has_coeffs, const kvz_cuda_rdoq_ctx *d_ctx,
                                     const kvz_cuda_rdoq_job *d_rdoq_jobs, int max_width, int16_t *d_work, int rdoq_skip,
                                     void *stream)
{
  if (!check_qcfg(cfg, "kvz_cuda_quantize_residuals_rdoq: bad cfg")) return 0;

  if (!d_ref || !d_pred || !d_rec || !d_coeff || !d_jobs || !d_has_coeffs || !d_ctx || !d_rdoq_jobs || !d_work || n_jobs < 0)
    return fail_msg("kvz_cuda_quantize_residuals_rdoq: bad argument");
  if (max_width != 4 && max_width != 8 && max_width != 16 && max_width != 32)
    return fail_msg("kvz_cuda_quantize_residuals_rdoq: max_width must be 4, 8, 16 or 32");
  return launch_quantize_residual_rdoq(*cfg, d_ref, d_pred, d_rec, d_coeff, d_jobs, n_jobs, d_has_coeffs, d_ctx,
                                       d_rdoq_jobs, max_width, d_work, rdoq_skip, (cudaStream_t)stream);
}

int kvz_cuda_quantize_residuals_rdoq_noalloc(const kvz_cuda_quant_cfg *cfg, const void *d_ref, const void *d_pred,
                                             void *d_rec, int16_t *d_coeff, size_t coeff_elems,
                                             const kvz_cuda_residual_job *d_jobs, int n_jobs, uint8_t *d_has_coeffs,
                                             const kvz_cuda_rdoq_ctx *d_ctx, const kvz_cuda_rdoq_job *d_rdoq_jobs,
                                             int max_width, int rdoq_skip, void *stream)
{
  if (!check_qcfg(cfg, "kvz_cuda_quantize_residuals_rdoq_noalloc: bad cfg")) return 0;
  if (!d_ref || !d_pred || !d_rec || !d_coeff || !d_jobs || !d_has_coeffs || !d_ctx || !d_rdoq_jobs || n_jobs < 0)
    return fail_msg("kvz_cuda_quantize_residuals_rdoq_noalloc: bad argument");
  if (max_width != 4 && max_width != 8 && max_width != 16 && max_width != 32)
    return fail_msg("kvz_cuda_quantize_residuals_rdoq_noalloc: max_width must be 4, 8, 16 or 32");
  return launch_quantize_residual_rdoq_noalloc(*cfg, d_ref, d_pred, d_rec, d_coeff, coeff_elems, d_jobs, n_jobs, d_has_coeffs,
                                               d_ctx, d_rdoq_jobs, max_width, rdoq_skip, (cudaStream_t)stream);
}

int kvz_cuda_coeff_abs_sums(const int16_t *d_coeffs, int per_block, int count, uint32_t *d_out, void *stream)
{
  if (!d_coeffs || !d_out || per_block <= 0 || count < 0) return fail_msg("kvz_cuda_coeff_abs_sums: bad argument");
  return launch_coeff_abs_sum(d_coeffs, per_block, count, d_out, (cudaStream_t)stream);
}

int kvz_cuda_quantize_residuals(const kvz_cuda_quant_cfg *cfg, const void *d_ref, const void *d_pred, void *d_rec,
                                int16_t *d_coeff, const kvz_cuda_residual_job *d_jobs, int n_jobs,
                                uint8_t *d_has_coeffs, void *stream)
{
  if (!check_qcfg(cfg, "kvz_cuda_quantize_residuals: bad cfg")) return 0;

  if (!d_ref || !d_pred || !d_rec || !d_coeff || !d_jobs || !d_has_coeffs || n_jobs < 0)
    return fail_msg("kvz_cuda_quantize_residuals: bad argument");
  return launch_quantize_residual(*cfg, d_ref, d_pred, d_rec, d_coeff, d_jobs, n_jobs, d_has_coeffs, (cudaStream_t)stream);
}

int kvz_cuda_interp_blocks(const void *d_plane, int pitch, int width, int height, int bitdepth,
                           const kvz_cuda_interp_job *d_jobs, int n_jobs, void *d_dst8, int16_t *d_dst16, void *stream)
{
  if (!d_plane || !d_jobs || n_jobs < 0 || width <= 0 || height <= 0 || (!d_dst8 && !d_dst16))
    return fail_msg("kvz_cuda_interp_blocks: bad argument");
  if (bitdepth != 8 && bitdepth != 10) return fail_msg("kvz_cuda_interp_blocks: bitdepth must be 8 or 10");
  return launch_interp(d_plane, pitch, width, height, bitdepth, d_jobs, n_jobs, d_dst8, d_dst16, (cudaStream_t)stream);
}

int kvz_cuda_bipred_blend(const int16_t *d_s0, const int16_t *d_s1, size_t n, int bitdepth, void *d_dst, void *stream)
{
  if (!d_s0 || !d_s1 || !d_dst) return fail_msg("kvz_cuda_bipred_blend: null argument");
  if (bitdepth != 8 && bitdepth != 10) return fail_msg("kvz_cuda_bipred_blend: bitdepth must be 8 or 10");
  return launch_bipred_blend(d_s0, d_s1, n, bitdepth, d_dst, (cudaStream_t)stream);
}

int kvz_cuda_intra_mode_costs(const void *d_refs, const void *d_orig, const kvz_cuda_intra_job *d_jobs, int n_jobs,
                              int bitdepth, uint32_t *d_costs, void *d_preds, void *stream)
{
  if (!d_refs || !d_jobs || n_jobs < 0 || (!d_costs && !d_preds)) return fail_msg("kvz_cuda_intra_mode_costs: bad argument");
  if (bitdepth != 8 && bitdepth != 10) return fail_msg("kvz_cuda_intra_mode_costs: bitdepth must be 8 or 10");
  return launch_intra_costs(d_refs, d_orig, d_jobs, n_jobs, d_costs, d_preds, (cudaStream_t)stream, bitdepth);
}

}  // extern "C"
