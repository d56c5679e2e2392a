// kvazaar_b200/csrc/percall.cu
//
// PER-CALL face: functions with exactly the strategy typedefs of the reference
// (strategies-picture.h:102-136, strategies-dct.h:32, strategies-quant.h:48,
// strategies-intra.h:32-43, strategies-ipol.h:36-48) so that they can be handed to
// kvz_strategyselector_register() under strategy name "cuda".  Each call stages its host
// buffers through pinned memory, launches the same kernels the batch face uses on a
// batch of one, and waits.  This face exists for parity tests and the kernel sweep
// (BASELINE config 5); a launch per 8x8 block is not how throughput is obtained.
// Thread-safe through one lock (strategies are called from every threadqueue worker).
#include <dlfcn.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"
#include "../../include/kvz_cuda.h"
#include "../../include/kvz_cuda_strategies.h"

namespace kvz {
int launch_block_cost(const void *base_a, const void *base_b, const kvz_cuda_block_job *jobs, int n_jobs,
                      int bitdepth, uint32_t *out, cudaStream_t stream);
int launch_transform_batch(int n, int dst4, int inverse, int bitdepth, const int16_t *in, int16_t *out, int count,
                           cudaStream_t stream);
int launch_coeff_abs_sum(const int16_t *coeffs, int per_block, int count, uint32_t *out, cudaStream_t stream);
int launch_picture_checksum(const kvz_cuda_plane *planes, int n_planes, int bitdepth, uint32_t *d_out, cudaStream_t stream, int y0);
int launch_sao_stats(const kvz_cuda_plane *orig, const kvz_cuda_plane *rec, int n_planes, int bitdepth,
                     kvz_cuda_sao_stats *d_stats, cudaStream_t stream);
int launch_sao_apply(const kvz_cuda_plane *rec_in, const kvz_cuda_plane *rec_out, int n_planes, int bitdepth,
                     const kvz_cuda_sao_params *d_luma, const kvz_cuda_sao_params *d_chroma, cudaStream_t stream);
int launch_intra_costs(const void *refs, const void *orig, const kvz_cuda_intra_job *jobs, int n_jobs,
                       uint32_t *costs, void *preds, cudaStream_t stream, int bitdepth);
int launch_interp(const void *plane, int pitch, int W, int H, int bitdepth, const kvz_cuda_interp_job *jobs, int n_jobs,
                  void *dst8, int16_t *dst16, cudaStream_t stream);
int launch_rdoq_batch(const kvz_cuda_quant_cfg &cfg, const kvz_cuda_rdoq_ctx *d_ctx, const int16_t *d_coef,
                      int16_t *d_q_coef, const kvz_cuda_rdoq_job *d_jobs, int n_jobs, int max_width, int skip4,
                      cudaStream_t stream);
int launch_quantize_residual_rdoq(const kvz_cuda_quant_cfg &c, const void *ref, const void *pred, void *rec, int16_t *coeff,
                                  const kvz_cuda_residual_job *jobs, int n_jobs, uint8_t *has_coeffs,
                                  const kvz_cuda_rdoq_ctx *d_ctx, const kvz_cuda_rdoq_job *rdoq_jobs, int max_width,
                                  int16_t *work, int rdoq_skip, cudaStream_t stream);

namespace {

constexpr size_t kStage = 1 << 20;      // bytes of each staging buffer

struct PerCall {
  std::mutex lock;
  cudaStream_t stream = nullptr;
  uint8_t *h = nullptr, *d = nullptr;   // pinned host / device staging, kStage bytes each
  bool ready = false;
  bool init()
  {
    if (ready) return true;
    if (cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking) != cudaSuccess) return false;
    if (cudaMallocHost(&h, kStage) != cudaSuccess) return false;
    if (cudaMalloc(&d, kStage) != cudaSuccess) return false;
    ready = true;
    return true;
  }
};

PerCall g_pc;

// the GPU of this process when the host never called kvz_cuda_init(): KVZ_CUDA_DEVICE, else LOCAL_RANK (one process per GPU), else 0
int env_device()
{
  const char *v = getenv("KVZ_CUDA_DEVICE");
  if (!v) v = getenv("LOCAL_RANK");
  return v ? atoi(v) : 0;
}

[[noreturn]] void die(const char *what)
{
  // a cost function cannot report failure (SURVEY.md 8b "Errors"): fail loudly, never fall back
  fprintf(stderr, "libkvz_cuda: %s failed: %s\n", what, g_last_error);
  abort();
}

// calls served per strategy family: picture, dct, quant, intra, ipol, sao, nal (kvz_cuda_percall_counts)
unsigned long long g_family_calls[7];
int family_of(const char *who)
{
  if (strstr(who, "sao")) return 5;
  if (strstr(who, "checksum")) return 6;
  if (strstr(who, "quant") || strstr(who, "coeff_abs_sum") || strstr(who, "rdoq")) return 2;
  if (strstr(who, "dct") || strstr(who, "dst") || strstr(who, "transform")) return 1;
  if (strstr(who, "angular") || strstr(who, "planar") || strstr(who, "intra")) return 3;
  if (strstr(who, "sample") || strstr(who, "filter") || strstr(who, "ipol") || strstr(who, "interp")) return 4;
  return 0;
}

// layout of the staging buffers for one call: [0, in_bytes) inputs, then outputs
struct Stage {
  std::unique_lock<std::mutex> guard;
  size_t in_bytes = 0, out_bytes = 0;
  explicit Stage(const char *who) : guard(g_pc.lock)
  {
    // strategies are called from every threadqueue worker; a fresh thread's current device is 0, not the one
    // kvz_cuda_init() selected, so each call binds it (cheap when it already is)
    if (g_device < 0 && !kvz_cuda_init(env_device())) die(who);
    if (cudaSetDevice(g_device) != cudaSuccess) { fail("cudaSetDevice", cudaGetLastError(), __FILE__, __LINE__); die(who); }
    if (!g_pc.init()) { fail_msg("per-call staging buffers"); die(who); }
    g_family_calls[family_of(who)]++;                       // under the lock
  }
  template <typename T> T *in(size_t count)      // reserve input space, 16-byte aligned
  {
    in_bytes = (in_bytes + 15) & ~(size_t)15;
    T *p = reinterpret_cast<T *>(g_pc.h + in_bytes);
    in_bytes += count * sizeof(T);
    return p;
  }
  template <typename T> const T *dev(const T *host_ptr) const
  {
    return reinterpret_cast<const T *>(g_pc.d + (reinterpret_cast<const uint8_t *>(host_ptr) - g_pc.h));
  }
  template <typename T> T *dev(T *host_ptr) const
  {
    return reinterpret_cast<T *>(g_pc.d + (reinterpret_cast<uint8_t *>(host_ptr) - g_pc.h));
  }
  // outputs live after the inputs; call after all in<>() reservations
  template <typename T> T *out(size_t count)
  {
    size_t off = (in_bytes + out_bytes + 15) & ~(size_t)15;
    out_bytes = off - in_bytes + count * sizeof(T);
    return reinterpret_cast<T *>(g_pc.h + off);
  }
  void upload(const char *who)
  {
    if (cudaMemcpyAsync(g_pc.d, g_pc.h, in_bytes, cudaMemcpyHostToDevice, g_pc.stream) != cudaSuccess) die(who);
  }
  void download(const char *who)
  {
    if (cudaMemcpyAsync(g_pc.h + in_bytes, g_pc.d + in_bytes, out_bytes, cudaMemcpyDeviceToHost, g_pc.stream) != cudaSuccess) die(who);
    if (cudaStreamSynchronize(g_pc.stream) != cudaSuccess) { fail("stream sync", cudaGetLastError(), __FILE__, __LINE__); die(who); }
  }
};

void pack_rows(uint8_t *dst, const uint8_t *src, int w, int h, size_t stride)
{
  for (int y = 0; y < h; y++) memcpy(dst + (size_t)y * w, src + (size_t)y * stride, (size_t)w);
}

// one or more predictions against one original block
void cost_call(const char *who, int kind, int w, int h, const uint8_t *const *preds, size_t pred_stride, int n_preds,
               const uint8_t *orig, size_t orig_stride, unsigned *costs_out)
{
  Stage st(who);
  uint8_t *o = st.in<uint8_t>((size_t)w * h);
  pack_rows(o, orig, w, h, orig_stride);
  uint8_t *p = st.in<uint8_t>((size_t)w * h * n_preds);
  for (int k = 0; k < n_preds; k++) pack_rows(p + (size_t)k * w * h, preds[k], w, h, pred_stride);
  kvz_cuda_block_job *jobs = st.in<kvz_cuda_block_job>(n_preds);
  for (int k = 0; k < n_preds; k++) {
    jobs[k].a_offset = (uint32_t)(k * w * h); jobs[k].b_offset = 0;
    jobs[k].a_stride = w; jobs[k].b_stride = w;
    jobs[k].w = (int16_t)w; jobs[k].h = (int16_t)h; jobs[k].kind = kind;
  }
  uint32_t *out = st.out<uint32_t>(n_preds);
  st.upload(who);
  if (!launch_block_cost(st.dev(p), st.dev(o), st.dev(jobs), n_preds, 8, st.dev(out), g_pc.stream)) die(who);
  st.download(who);
  for (int k = 0; k < n_preds; k++) costs_out[k] = out[k];
}

unsigned cost_pair(const char *who, int kind, int w, int h, const uint8_t *a, size_t sa, const uint8_t *b, size_t sb)
{
  unsigned c = 0;
  cost_call(who, kind, w, h, &a, sa, 1, b, sb, &c);
  return c;
}

}  // namespace
}  // namespace kvz

using namespace kvz;

namespace kvz {
int launch_quant_batch(const kvz_cuda_quant_cfg &c, int n, int type, int scan_idx, int dequant, const int16_t *in,
                       int16_t *out, int count, cudaStream_t stream);
int launch_quantize_residual(const kvz_cuda_quant_cfg &c, const void *ref, const void *pred, void *rec, int16_t *coeff,
                             const kvz_cuda_residual_job *jobs, int n_jobs, uint8_t *has_coeffs, cudaStream_t stream);
}

extern "C" {

// ---------------- picture ----------------
unsigned kvz_cuda_reg_sad(const kvz_cuda_px *data1, const kvz_cuda_px *data2, int width, int height,
                          unsigned stride1, unsigned stride2)
{
  return cost_pair("kvz_cuda_reg_sad", KVZ_CUDA_COST_REG_SAD, width, height, data1, stride1, data2, stride2);
}

#define KVZ_NXN(N)                                                                                         \
  unsigned kvz_cuda_sad_##N##x##N(const kvz_cuda_px *b1, const kvz_cuda_px *b2)                            \
  { return cost_pair("kvz_cuda_sad_" #N "x" #N, KVZ_CUDA_COST_SAD_NXN, N, N, b1, N, b2, N); }              \
  unsigned kvz_cuda_satd_##N##x##N(const kvz_cuda_px *b1, const kvz_cuda_px *b2)                           \
  { return cost_pair("kvz_cuda_satd_" #N "x" #N, KVZ_CUDA_COST_SATD_NXN, N, N, b1, N, b2, N); }            \
  void kvz_cuda_sad_##N##x##N##_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig,            \
                                     unsigned num_modes, unsigned *costs_out)                              \
  {                                                                                                        \
    const uint8_t *p[8];                                                                                   \
    for (unsigned i = 0; i < num_modes && i < 8; i++) p[i] = preds[i];                                     \
    cost_call("kvz_cuda_sad_" #N "x" #N "_dual", KVZ_CUDA_COST_SAD_NXN, N, N, p, N, (int)num_modes, orig, N, costs_out);   \
  }                                                                                                        \
  void kvz_cuda_satd_##N##x##N##_dual(const kvz_cuda_pred_buffer preds, const kvz_cuda_px *orig,           \
                                      unsigned num_modes, unsigned *costs_out)                             \
  {                                                                                                        \
    const uint8_t *p[8];                                                                                   \
    for (unsigned i = 0; i < num_modes && i < 8; i++) p[i] = preds[i];                                     \
    cost_call("kvz_cuda_satd_" #N "x" #N "_dual", KVZ_CUDA_COST_SATD_NXN, N, N, p, N, (int)num_modes, orig, N, costs_out);  \
  }
KVZ_NXN(4)
KVZ_NXN(8)
KVZ_NXN(16)
KVZ_NXN(32)
KVZ_NXN(64)
#undef KVZ_NXN

unsigned kvz_cuda_satd_any_size(int width, int height, const kvz_cuda_px *block1, int stride1,
                                const kvz_cuda_px *block2, int stride2)
{
  return cost_pair("kvz_cuda_satd_any_size", KVZ_CUDA_COST_SATD_ANY, width, height, block1, (size_t)stride1, block2, (size_t)stride2);
}

void kvz_cuda_satd_any_size_quad(int width, int height, const kvz_cuda_px **preds, const int stride,
                                 const kvz_cuda_px *orig, const int orig_stride, unsigned num_modes,
                                 unsigned *costs_out, int8_t *valid)
{
  (void)valid;    // the generic strategy evaluates all predictions too (picture-generic.c:392-456)
  cost_call("kvz_cuda_satd_any_size_quad", KVZ_CUDA_COST_SATD_QUAD, width, height, preds, (size_t)stride, (int)num_modes,
            orig, (size_t)orig_stride, costs_out);
}

unsigned kvz_cuda_pixels_calc_ssd(const kvz_cuda_px *ref, const kvz_cuda_px *rec, const int ref_stride,
                                  const int rec_stride, const int width)
{
  return cost_pair("kvz_cuda_pixels_calc_ssd", KVZ_CUDA_COST_SSD, width, width, ref, (size_t)ref_stride, rec, (size_t)rec_stride);
}

// ---------------- dct ----------------
static unsigned transform_call(const char *who, int n, int dst4, int inverse, int8_t bitdepth, const int16_t *input,
                               int16_t *output)
{
  Stage st(who);
  int16_t *in = st.in<int16_t>((size_t)n * n);
  memcpy(in, input, sizeof(int16_t) * n * n);
  int16_t *out = st.out<int16_t>((size_t)n * n);
  st.upload(who);
  if (!launch_transform_batch(n, dst4, inverse, bitdepth, st.dev(in), st.dev(out), 1, g_pc.stream)) die(who);
  st.download(who);
  memcpy(output, out, sizeof(int16_t) * n * n);
  return 0;
}

#define KVZ_TR(name, n, dst4, inv) \
  unsigned name(int8_t bitdepth, const int16_t *input, int16_t *output) { return transform_call(#name, n, dst4, inv, bitdepth, input, output); }
KVZ_TR(kvz_cuda_fast_forward_dst_4x4, 4, 1, 0)
KVZ_TR(kvz_cuda_dct_4x4, 4, 0, 0)
KVZ_TR(kvz_cuda_dct_8x8, 8, 0, 0)
KVZ_TR(kvz_cuda_dct_16x16, 16, 0, 0)
KVZ_TR(kvz_cuda_dct_32x32, 32, 0, 0)
KVZ_TR(kvz_cuda_fast_inverse_dst_4x4, 4, 1, 1)
KVZ_TR(kvz_cuda_idct_4x4, 4, 0, 1)
KVZ_TR(kvz_cuda_idct_8x8, 8, 0, 1)
KVZ_TR(kvz_cuda_idct_16x16, 16, 0, 1)
KVZ_TR(kvz_cuda_idct_32x32, 32, 0, 1)
#undef KVZ_TR

// ---------------- quant ----------------
uint32_t kvz_cuda_coeff_abs_sum(const kvz_cuda_coeff *coeffs, size_t length)
{
  const char *who = "kvz_cuda_coeff_abs_sum";
  Stage st(who);
  int16_t *in = st.in<int16_t>(length);
  memcpy(in, coeffs, sizeof(int16_t) * length);
  uint32_t *out = st.out<uint32_t>(1);
  st.upload(who);
  if (!launch_coeff_abs_sum(st.dev(in), (int)length, 1, st.dev(out), g_pc.stream)) die(who);
  st.download(who);
  return out[0];
}

int kvz_cuda_quant_host(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_coeff *coef, kvz_cuda_coeff *q_coef,
                        int32_t width, int32_t height, int8_t type, int8_t scan_idx)
{
  const char *who = "kvz_cuda_quant_host";
  if (!cfg || width != height) return fail_msg("kvz_cuda_quant_host: bad argument");
  Stage st(who);
  int16_t *in = st.in<int16_t>((size_t)width * width);
  memcpy(in, coef, sizeof(int16_t) * width * width);
  int16_t *out = st.out<int16_t>((size_t)width * width);
  st.upload(who);
  if (!launch_quant_batch(*cfg, width, type, scan_idx, 0, st.dev(in), st.dev(out), 1, g_pc.stream)) return 0;
  st.download(who);
  memcpy(q_coef, out, sizeof(int16_t) * width * width);
  return 1;
}

int kvz_cuda_dequant_host(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_coeff *q_coef, kvz_cuda_coeff *coef,
                          int32_t width, int32_t height, int8_t type)
{
  const char *who = "kvz_cuda_dequant_host";
  if (!cfg || width != height) return fail_msg("kvz_cuda_dequant_host: bad argument");
  Stage st(who);
  int16_t *in = st.in<int16_t>((size_t)width * width);
  memcpy(in, q_coef, sizeof(int16_t) * width * width);
  int16_t *out = st.out<int16_t>((size_t)width * width);
  st.upload(who);
  if (!launch_quant_batch(*cfg, width, type, 0, 1, st.dev(in), st.dev(out), 1, g_pc.stream)) return 0;
  st.download(who);
  memcpy(coef, out, sizeof(int16_t) * width * width);
  return 1;
}

int kvz_cuda_quantize_residual_host(const kvz_cuda_quant_cfg *cfg, int width, int color, int scan_order, int use_trskip,
                                    int in_stride, int out_stride, const kvz_cuda_px *ref_in, const kvz_cuda_px *pred_in,
                                    kvz_cuda_px *rec_out, kvz_cuda_coeff *coeff_out)
{
  const char *who = "kvz_cuda_quantize_residual_host";
  if (!cfg) return fail_msg("kvz_cuda_quantize_residual_host: bad argument");
  Stage st(who);
  const size_t nn = (size_t)width * width;
  uint8_t *r = st.in<uint8_t>(nn), *p = st.in<uint8_t>(nn);
  pack_rows(r, ref_in, width, width, (size_t)in_stride);
  pack_rows(p, pred_in, width, width, (size_t)in_stride);
  kvz_cuda_residual_job *jb = st.in<kvz_cuda_residual_job>(1);
  memset(jb, 0, sizeof(*jb));
  jb->in_stride = width; jb->out_stride = width;
  jb->width = (uint8_t)width; jb->color = (uint8_t)color; jb->scan_idx = (uint8_t)scan_order;
  jb->use_trskip = (uint8_t)use_trskip; jb->qp = (int8_t)cfg->qp; jb->cu_is_intra = (uint8_t)cfg->cu_is_intra;
  uint8_t *rec = st.out<uint8_t>(nn);
  int16_t *co = st.out<int16_t>(nn);
  uint8_t *has = st.out<uint8_t>(1);
  st.upload(who);
  if (!launch_quantize_residual(*cfg, st.dev(r), st.dev(p), st.dev(rec), st.dev(co), st.dev(jb), 1, st.dev(has), g_pc.stream))
    die(who);
  st.download(who);
  memcpy(coeff_out, co, sizeof(int16_t) * nn);
  for (int y = 0; y < width; y++) memcpy(rec_out + (size_t)y * out_stride, rec + (size_t)y * width, (size_t)width);
  return has[0];
}

// kvz_quantize_residual with cfg.rdoq_enable (quant-generic.c:211-221): the same call with kvz_rdoq between the
// transform and the reconstruction.  ctx = the CABAC context states of the calling thread's encoder state, lambda =
// state->lambda, tr_depth as computed at quant-generic.c:216-217.
int kvz_cuda_quantize_residual_rdoq_host(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_rdoq_ctx *ctx, double lambda,
                                         int tr_depth, int rdoq_skip, int width, int color, int scan_order, int use_trskip,
                                         int in_stride, int out_stride, const kvz_cuda_px *ref_in,
                                         const kvz_cuda_px *pred_in, kvz_cuda_px *rec_out, kvz_cuda_coeff *coeff_out)
{
  const char *who = "kvz_cuda_quantize_residual_rdoq_host";
  if (!cfg || !ctx) return fail_msg("kvz_cuda_quantize_residual_rdoq_host: bad argument");
  Stage st(who);
  const size_t nn = (size_t)width * width;
  uint8_t *r = st.in<uint8_t>(nn), *p = st.in<uint8_t>(nn);
  pack_rows(r, ref_in, width, width, (size_t)in_stride);
  pack_rows(p, pred_in, width, width, (size_t)in_stride);
  kvz_cuda_residual_job *jb = st.in<kvz_cuda_residual_job>(1);
  memset(jb, 0, sizeof(*jb));
  jb->in_stride = width; jb->out_stride = width;
  jb->width = (uint8_t)width; jb->color = (uint8_t)color; jb->scan_idx = (uint8_t)scan_order;
  jb->use_trskip = (uint8_t)use_trskip; jb->qp = (int8_t)cfg->qp; jb->cu_is_intra = (uint8_t)cfg->cu_is_intra;
  kvz_cuda_rdoq_job *rj = st.in<kvz_cuda_rdoq_job>(1);
  memset(rj, 0, sizeof(*rj));
  rj->lambda = lambda; rj->width = (uint8_t)width; rj->type = color == 0 ? 0 : 2; rj->scan_idx = (uint8_t)scan_order;
  rj->tr_depth = (uint8_t)tr_depth; rj->qp = (int8_t)cfg->qp; rj->cu_is_intra = (uint8_t)cfg->cu_is_intra;
  kvz_cuda_rdoq_ctx *cx = st.in<kvz_cuda_rdoq_ctx>(1);
  *cx = *ctx;
  int16_t *work = st.in<int16_t>(nn);                  // device-only scratch: its staging copy is never read
  uint8_t *rec = st.out<uint8_t>(nn);
  int16_t *co = st.out<int16_t>(nn);
  uint8_t *has = st.out<uint8_t>(1);
  st.upload(who);
  if (!launch_quantize_residual_rdoq(*cfg, st.dev(r), st.dev(p), st.dev(rec), st.dev(co), st.dev(jb), 1, st.dev(has), st.dev(cx),
                                     st.dev(rj), width, st.dev(work), rdoq_skip, g_pc.stream)) die(who);
  st.download(who);
  memcpy(coeff_out, co, sizeof(int16_t) * nn);
  // quant-generic.c:262-271: with no coefficients and rec_out == pred_in nothing is written
  if (has[0] || rec_out != pred_in)
    for (int y = 0; y < width; y++) memcpy(rec_out + (size_t)y * out_stride, rec + (size_t)y * width, (size_t)width);
  return has[0];
}

// kvz_rdoq (rdo.c:548-885) for one block: the call kvz_quant's callers make when cfg.rdoq_enable (transform.c, intra RDO)
int kvz_cuda_rdoq_host(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_rdoq_ctx *ctx, double lambda, const kvz_cuda_coeff *coef,
                       kvz_cuda_coeff *dest_coeff, int width, int type, int scan_idx, int tr_depth)
{
  const char *who = "kvz_cuda_rdoq_host";
  if (!cfg || !ctx) return fail_msg("kvz_cuda_rdoq_host: bad argument");
  Stage st(who);
  const size_t nn = (size_t)width * width;
  int16_t *in = st.in<int16_t>(nn);
  memcpy(in, coef, sizeof(int16_t) * nn);
  kvz_cuda_rdoq_job *rj = st.in<kvz_cuda_rdoq_job>(1);
  memset(rj, 0, sizeof(*rj));
  rj->lambda = lambda; rj->width = (uint8_t)width; rj->type = (uint8_t)type; rj->scan_idx = (uint8_t)scan_idx;
  rj->tr_depth = (uint8_t)tr_depth; rj->qp = (int8_t)cfg->qp; rj->cu_is_intra = (uint8_t)cfg->cu_is_intra;
  kvz_cuda_rdoq_ctx *cx = st.in<kvz_cuda_rdoq_ctx>(1);
  *cx = *ctx;
  int16_t *out = st.out<int16_t>(nn);
  st.upload(who);
  // rdoq_job.coef_offset addresses both arrays: the output array starts where the staged input does
  if (!launch_rdoq_batch(*cfg, st.dev(cx), st.dev(in), st.dev(out), st.dev(rj), 1, width, 0, g_pc.stream)) die(who);
  st.download(who);
  memcpy(dest_coeff, out, sizeof(int16_t) * nn);
  return 1;
}

// ---------------- intra ----------------
// angular_pred_func / intra_pred_planar_func receive the references kvz_intra_predict has already
// chosen (filtered or not), so the job runs as chroma (no smoothing, no boundary filters) and
// the wanted mode is picked from the 35 predictions.
static void intra_call(const char *who, int lg, int mode, const kvz_cuda_px *top, const kvz_cuda_px *left, kvz_cuda_px *dst)
{
  const int w = 1 << lg, n = 2 * w + 1;
  Stage st(who);
  uint8_t *refs = st.in<uint8_t>((size_t)2 * n);
  memcpy(refs, top, (size_t)n);
  memcpy(refs + n, left, (size_t)n);
  kvz_cuda_intra_job *jb = st.in<kvz_cuda_intra_job>(1);
  memset(jb, 0, sizeof(*jb));
  jb->log2_width = (uint8_t)lg; jb->is_chroma = 1;
  uint8_t *preds = st.out<uint8_t>((size_t)35 * w * w);
  st.upload(who);
  if (!launch_intra_costs(st.dev(refs), nullptr, st.dev(jb), 1, nullptr, st.dev(preds), g_pc.stream, 8)) die(who);
  st.download(who);
  memcpy(dst, preds + (size_t)mode * w * w, (size_t)w * w);
}

void kvz_cuda_angular_pred(const int_fast8_t log2_width, const int_fast8_t intra_mode, const kvz_cuda_px *const in_ref_above,
                           const kvz_cuda_px *const in_ref_left, kvz_cuda_px *const dst)
{
  intra_call("kvz_cuda_angular_pred", log2_width, intra_mode, in_ref_above, in_ref_left, dst);
}

void kvz_cuda_intra_pred_planar(const int_fast8_t log2_width, const kvz_cuda_px *const ref_top,
                                const kvz_cuda_px *const ref_left, kvz_cuda_px *const dst)
{
  intra_call("kvz_cuda_intra_pred_planar", log2_width, 0, ref_top, ref_left, dst);
}

// ---------------- ipol ----------------
static void sample_call(const char *who, bool chroma, bool hi, const kvz_cuda_px *src, int src_stride, int w, int h,
                        void *dst, int dst_stride, const int16_t mv[2])
{
  const int taps = chroma ? 4 : 8, before = taps / 2 - 1, mask = chroma ? 7 : 3;
  const int ww = w + taps - 1, wh = h + taps - 1;
  Stage st(who);
  uint8_t *win = st.in<uint8_t>((size_t)ww * wh);
  pack_rows(win, src - (ptrdiff_t)before * src_stride - before, ww, wh, (size_t)src_stride);
  kvz_cuda_interp_job *jb = st.in<kvz_cuda_interp_job>(1);
  memset(jb, 0, sizeof(*jb));
  jb->x = before; jb->y = before; jb->mv_x = mv[0] & mask; jb->mv_y = mv[1] & mask;
  jb->w = (int16_t)w; jb->h = (int16_t)h; jb->chroma = chroma; jb->mode = hi ? KVZ_CUDA_INTERP_14BIT : KVZ_CUDA_INTERP_PIXEL;
  jb->dst_offset = 0; jb->dst_stride = w;
  uint8_t *o8 = hi ? nullptr : st.out<uint8_t>((size_t)w * h);
  int16_t *o16 = hi ? st.out<int16_t>((size_t)w * h) : nullptr;
  st.upload(who);
  if (!launch_interp(st.dev(win), ww, ww, wh, 8, st.dev(jb), 1, hi ? nullptr : (void *)st.dev(o8), hi ? st.dev(o16) : nullptr,
                     g_pc.stream)) die(who);
  st.download(who);
  for (int y = 0; y < h; y++) {
    if (hi) memcpy((int16_t *)dst + (size_t)y * dst_stride, o16 + (size_t)y * w, sizeof(int16_t) * w);
    else memcpy((uint8_t *)dst + (size_t)y * dst_stride, o8 + (size_t)y * w, (size_t)w);
  }
}

void kvz_cuda_sample_quarterpel_luma(const void *, kvz_cuda_px *src, int16_t src_stride, int width, int height,
                                     kvz_cuda_px *dst, int16_t dst_stride, int8_t, int8_t, const int16_t mv[2])
{ sample_call("kvz_cuda_sample_quarterpel_luma", false, false, src, src_stride, width, height, dst, dst_stride, mv); }

void kvz_cuda_sample_octpel_chroma(const void *, kvz_cuda_px *src, int16_t src_stride, int width, int height,
                                   kvz_cuda_px *dst, int16_t dst_stride, int8_t, int8_t, const int16_t mv[2])
{ sample_call("kvz_cuda_sample_octpel_chroma", true, false, src, src_stride, width, height, dst, dst_stride, mv); }

void kvz_cuda_sample_14bit_quarterpel_luma(const void *, kvz_cuda_px *src, int16_t src_stride, int width, int height,
                                           int16_t *dst, int16_t dst_stride, int8_t, int8_t, const int16_t mv[2])
{ sample_call("kvz_cuda_sample_14bit_quarterpel_luma", false, true, src, src_stride, width, height, dst, dst_stride, mv); }

void kvz_cuda_sample_14bit_octpel_chroma(const void *, kvz_cuda_px *src, int16_t src_stride, int width, int height,
                                         int16_t *dst, int16_t dst_stride, int8_t, int8_t, const int16_t mv[2])
{ sample_call("kvz_cuda_sample_14bit_octpel_chroma", true, true, src, src_stride, width, height, dst, dst_stride, mv); }

// The FME block filters: src is orig_topleft of an extended block fetched one sample up-left of the
// integer vector with (width + 1) x (height + 1) samples and 4 samples of margin (search_inter.c:1008-1013),
// so the block at the integer vector starts at src + (1, 1).
static void fme_blocks_call(const char *who, const kvz_cuda_px *src, int src_stride, int w, int h,
                            kvz_cuda_px filtered[4][64 * 64], const int q[4][2])
{
  const int ww = w + 9, wh = h + 9;
  Stage st(who);
  uint8_t *win = st.in<uint8_t>((size_t)ww * wh);
  pack_rows(win, src - (ptrdiff_t)4 * src_stride - 4, ww, wh, (size_t)src_stride);
  kvz_cuda_interp_job *jb = st.in<kvz_cuda_interp_job>(4);
  memset(jb, 0, sizeof(*jb) * 4);
  for (int k = 0; k < 4; k++) {
    jb[k].x = 5; jb[k].y = 5; jb[k].mv_x = q[k][0]; jb[k].mv_y = q[k][1];
    jb[k].w = (int16_t)w; jb[k].h = (int16_t)h; jb[k].mode = KVZ_CUDA_INTERP_FME;
    jb[k].dst_offset = (uint32_t)(k * w * h); jb[k].dst_stride = w;
  }
  uint8_t *o = st.out<uint8_t>((size_t)4 * w * h);
  st.upload(who);
  if (!launch_interp(st.dev(win), ww, ww, wh, 8, st.dev(jb), 4, st.dev(o), nullptr, g_pc.stream)) die(who);
  st.download(who);
  for (int k = 0; k < 4; k++)
    for (int y = 0; y < h; y++) memcpy(&filtered[k][y * 64], o + ((size_t)k * h + y) * w, (size_t)w);
}

void kvz_cuda_filter_hpel_blocks_hor_ver_luma(KVZ_CUDA_IPOL_BLOCKS_ARGS)
{
  (void)encoder; (void)hor_intermediate; (void)fme_level; (void)hor_first_cols; (void)sample_off_x; (void)sample_off_y;
  const int q[4][2] = { { -2, 0 }, { 2, 0 }, { 0, -2 }, { 0, 2 } };
  fme_blocks_call("kvz_cuda_filter_hpel_blocks_hor_ver_luma", src, src_stride, width, height, filtered, q);
}

void kvz_cuda_filter_hpel_blocks_diag_luma(KVZ_CUDA_IPOL_BLOCKS_ARGS)
{
  (void)encoder; (void)hor_intermediate; (void)fme_level; (void)hor_first_cols; (void)sample_off_x; (void)sample_off_y;
  const int q[4][2] = { { -2, -2 }, { 2, -2 }, { -2, 2 }, { 2, 2 } };
  fme_blocks_call("kvz_cuda_filter_hpel_blocks_diag_luma", src, src_stride, width, height, filtered, q);
}

void kvz_cuda_filter_qpel_blocks_hor_ver_luma(KVZ_CUDA_IPOL_BLOCKS_ARGS)
{
  (void)encoder; (void)hor_intermediate; (void)fme_level; (void)hor_first_cols;
  const int ox = 2 * sample_off_x, oy = 2 * sample_off_y;
  const int q[4][2] = { { ox - 1, oy }, { ox + 1, oy }, { ox, oy - 1 }, { ox, oy + 1 } };
  fme_blocks_call("kvz_cuda_filter_qpel_blocks_hor_ver_luma", src, src_stride, width, height, filtered, q);
}

void kvz_cuda_filter_qpel_blocks_diag_luma(KVZ_CUDA_IPOL_BLOCKS_ARGS)
{
  (void)encoder; (void)hor_intermediate; (void)fme_level; (void)hor_first_cols;
  const int ox = 2 * sample_off_x, oy = 2 * sample_off_y;
  const int q[4][2] = { { ox - 1, oy - 1 }, { ox + 1, oy - 1 }, { ox - 1, oy + 1 }, { ox + 1, oy + 1 } };
  fme_blocks_call("kvz_cuda_filter_qpel_blocks_diag_luma", src, src_stride, width, height, filtered, q);
}

// ---------------- registration ----------------
// Same type strings as STRATEGIES_*_EXPORTS (strategies-picture.h:173-198, -dct.h:54-68,
// -quant.h:58-62, -intra.h:54-56, -ipol.h:64-73).
#define KVZ_REG(type, fn) ok &= register_fn(opaque, type, "cuda", 50, (void *)&fn)

static int cuda_usable()
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail_msg("kvz_strategy_register_*_cuda: no CUDA device (there is no CPU fallback)");
  }
  if (g_device < 0 && !kvz_cuda_init(env_device())) return 0;
  return 1;
}

int kvz_strategy_register_picture_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn)
{
  if (!register_fn) return 0;
  if (bitdepth != 8) return 1;
  if (!cuda_usable()) return 0;
  int ok = 1;
  KVZ_REG("reg_sad", kvz_cuda_reg_sad);
  KVZ_REG("sad_4x4", kvz_cuda_sad_4x4); KVZ_REG("sad_8x8", kvz_cuda_sad_8x8); KVZ_REG("sad_16x16", kvz_cuda_sad_16x16);
  KVZ_REG("sad_32x32", kvz_cuda_sad_32x32); KVZ_REG("sad_64x64", kvz_cuda_sad_64x64);
  KVZ_REG("satd_4x4", kvz_cuda_satd_4x4); KVZ_REG("satd_8x8", kvz_cuda_satd_8x8); KVZ_REG("satd_16x16", kvz_cuda_satd_16x16);
  KVZ_REG("satd_32x32", kvz_cuda_satd_32x32); KVZ_REG("satd_64x64", kvz_cuda_satd_64x64);
  KVZ_REG("satd_any_size", kvz_cuda_satd_any_size);
  KVZ_REG("sad_4x4_dual", kvz_cuda_sad_4x4_dual); KVZ_REG("sad_8x8_dual", kvz_cuda_sad_8x8_dual);
  KVZ_REG("sad_16x16_dual", kvz_cuda_sad_16x16_dual); KVZ_REG("sad_32x32_dual", kvz_cuda_sad_32x32_dual);
  KVZ_REG("sad_64x64_dual", kvz_cuda_sad_64x64_dual);
  KVZ_REG("satd_4x4_dual", kvz_cuda_satd_4x4_dual); KVZ_REG("satd_8x8_dual", kvz_cuda_satd_8x8_dual);
  KVZ_REG("satd_16x16_dual", kvz_cuda_satd_16x16_dual); KVZ_REG("satd_32x32_dual", kvz_cuda_satd_32x32_dual);
  KVZ_REG("satd_64x64_dual", kvz_cuda_satd_64x64_dual);
  KVZ_REG("satd_any_size_quad", kvz_cuda_satd_any_size_quad);
  KVZ_REG("pixels_calc_ssd", kvz_cuda_pixels_calc_ssd);
  return ok;
}

int kvz_strategy_register_dct_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn)
{
  if (!register_fn) return 0;
  if (bitdepth != 8) return 1;
  if (!cuda_usable()) return 0;
  int ok = 1;
  KVZ_REG("fast_forward_dst_4x4", kvz_cuda_fast_forward_dst_4x4);
  KVZ_REG("dct_4x4", kvz_cuda_dct_4x4); KVZ_REG("dct_8x8", kvz_cuda_dct_8x8);
  KVZ_REG("dct_16x16", kvz_cuda_dct_16x16); KVZ_REG("dct_32x32", kvz_cuda_dct_32x32);
  KVZ_REG("fast_inverse_dst_4x4", kvz_cuda_fast_inverse_dst_4x4);
  KVZ_REG("idct_4x4", kvz_cuda_idct_4x4); KVZ_REG("idct_8x8", kvz_cuda_idct_8x8);
  KVZ_REG("idct_16x16", kvz_cuda_idct_16x16); KVZ_REG("idct_32x32", kvz_cuda_idct_32x32);
  return ok;
}

int kvz_strategy_register_quant_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn)
{
  if (!register_fn) return 0;
  if (bitdepth != 8) return 1;
  if (!cuda_usable()) return 0;
  int ok = 1;
  // quant / quantize_residual / dequant read encoder_state_t: their one-line wrappers live in the
  // glue and call kvz_cuda_{quant,quantize_residual,dequant}_host
  KVZ_REG("coeff_abs_sum", kvz_cuda_coeff_abs_sum);
  return ok;
}

int kvz_strategy_register_intra_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn)
{
  if (!register_fn) return 0;
  if (bitdepth != 8) return 1;
  if (!cuda_usable()) return 0;
  int ok = 1;
  KVZ_REG("angular_pred", kvz_cuda_angular_pred);
  KVZ_REG("intra_pred_planar", kvz_cuda_intra_pred_planar);
  return ok;
}

int kvz_strategy_register_ipol_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn)
{
  if (!register_fn) return 0;
  if (bitdepth != 8) return 1;
  if (!cuda_usable()) return 0;
  int ok = 1;
  KVZ_REG("filter_hpel_blocks_hor_ver_luma", kvz_cuda_filter_hpel_blocks_hor_ver_luma);
  KVZ_REG("filter_hpel_blocks_diag_luma", kvz_cuda_filter_hpel_blocks_diag_luma);
  KVZ_REG("filter_qpel_blocks_hor_ver_luma", kvz_cuda_filter_qpel_blocks_hor_ver_luma);
  KVZ_REG("filter_qpel_blocks_diag_luma", kvz_cuda_filter_qpel_blocks_diag_luma);
  KVZ_REG("sample_quarterpel_luma", kvz_cuda_sample_quarterpel_luma);
  KVZ_REG("sample_octpel_chroma", kvz_cuda_sample_octpel_chroma);
  KVZ_REG("sample_14bit_quarterpel_luma", kvz_cuda_sample_14bit_quarterpel_luma);
  KVZ_REG("sample_14bit_octpel_chroma", kvz_cuda_sample_14bit_octpel_chroma);
  // get_extended_block stays with the host's generic strategy: it only copies memory
  return ok;
}

// ---------------- nal ----------------
void kvz_cuda_array_checksum(const kvz_cuda_px *data, const int height, const int width, const int stride,
                             unsigned char checksum_out[16], const uint8_t bitdepth)
{
  const char *who = "kvz_cuda_array_checksum";
  if (bitdepth != 8 || width <= 0 || height <= 0 || (size_t)width + 64 > kStage) { fail_msg("kvz_cuda_array_checksum: bad argument"); die(who); }
  uint32_t sum = 0;                                           // the checksum is a sum: strips of rows add up
  const int rows_per = (int)((kStage - 64) / (size_t)width);
  for (int y0 = 0; y0 < height; y0 += rows_per) {
    const int rows = height - y0 < rows_per ? height - y0 : rows_per;
    Stage st(who);
    uint8_t *in = st.in<uint8_t>((size_t)rows * width);
    pack_rows(in, data + (size_t)y0 * stride, width, rows, (size_t)stride);
    uint32_t *out = st.out<uint32_t>(1);
    st.upload(who);
    kvz_cuda_plane pl = { st.dev(in), width, rows, width };
    if (!launch_picture_checksum(&pl, 1, 8, st.dev(out), g_pc.stream, y0)) die(who);
    st.download(who);
    sum += *out;
  }
  checksum_out[0] = (unsigned char)(sum >> 24); checksum_out[1] = (unsigned char)(sum >> 16);
  checksum_out[2] = (unsigned char)(sum >> 8); checksum_out[3] = (unsigned char)sum;
}

int kvz_strategy_register_nal_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn)
{
  if (!register_fn) return 0;
  if (bitdepth != 8) return 1;
  if (!cuda_usable()) return 0;
  int ok = 1;
  KVZ_REG("array_checksum", kvz_cuda_array_checksum);
  return ok;
}

// ---------------- sao ----------------
// the statistics of one contiguous block_width x block_height block (<= 64x64, one CTU of the statistics kernel)
static void sao_block_stats(const char *who, const kvz_cuda_px *orig, const kvz_cuda_px *rec, int bw, int bh, int bitdepth,
                            kvz_cuda_sao_stats *result)
{
  if (bw <= 0 || bh <= 0 || bw > 64 || bh > 64) { fail_msg("SAO block larger than an LCU"); die(who); }
  Stage st(who);
  uint8_t *o = st.in<uint8_t>((size_t)bw * bh), *r = st.in<uint8_t>((size_t)bw * bh);
  memcpy(o, orig, (size_t)bw * bh);
  memcpy(r, rec, (size_t)bw * bh);
  kvz_cuda_sao_stats *out = st.out<kvz_cuda_sao_stats>(1);
  st.upload(who);
  kvz_cuda_plane po = { st.dev(o), bw, bh, bw }, pr = { st.dev(r), bw, bh, bw };
  if (!launch_sao_stats(&po, &pr, 1, bitdepth, st.dev(out), g_pc.stream)) die(who);
  st.download(who);
  *result = *out;
}

int kvz_cuda_sao_edge_ddistortion(const kvz_cuda_px *orig_data, const kvz_cuda_px *rec_data, int block_width, int block_height,
                                  int eo_class, int offsets[5])
{
  kvz_cuda_sao_stats s;
  sao_block_stats("kvz_cuda_sao_edge_ddistortion", orig_data, rec_data, block_width, block_height, 8, &s);
  int sum = 0;                                                // n * h^2 - 2 * h * E per category == sao-generic.c:51-81
  for (int c = 0; c < 5; c++) sum += s.edge[eo_class & 3][1][c] * offsets[c] * offsets[c] - 2 * offsets[c] * s.edge[eo_class & 3][0][c];
  return sum;
}

void kvz_cuda_calc_sao_edge_dir(const kvz_cuda_px *orig_data, const kvz_cuda_px *rec_data, int eo_class, int block_width,
                                int block_height, int cat_sum_cnt[2][5])
{
  kvz_cuda_sao_stats s;
  sao_block_stats("kvz_cuda_calc_sao_edge_dir", orig_data, rec_data, block_width, block_height, 8, &s);
  for (int k = 0; k < 2; k++)
    for (int c = 0; c < 5; c++) cat_sum_cnt[k][c] += s.edge[eo_class & 3][k][c];      // the reference accumulates
}

int kvz_cuda_sao_band_ddistortion_host(int bitdepth, const kvz_cuda_px *orig_data, const kvz_cuda_px *rec_data, int block_width,
                                       int block_height, int band_pos, const int sao_bands[4])
{
  if (bitdepth != 8) { fail_msg("kvz_cuda_sao_band_ddistortion_host: the per-call face is 8-bit"); die("kvz_cuda_sao_band_ddistortion_host"); }
  kvz_cuda_sao_stats s;
  sao_block_stats("kvz_cuda_sao_band_ddistortion_host", orig_data, rec_data, block_width, block_height, 8, &s);
  int sum = 0;
  for (int b = 0; b < 4; b++) {
    const int band = band_pos + b;
    if (band < 0 || band > 31) continue;
    sum += s.band[1][band] * sao_bands[b] * sao_bands[b] - 2 * sao_bands[b] * s.band[0][band];
  }
  return sum;
}

void kvz_cuda_sao_reconstruct_color_host(int bitdepth, const kvz_cuda_px *rec_data, kvz_cuda_px *new_rec_data,
                                         const kvz_cuda_sao_params *sao, int stride, int new_stride, int block_width,
                                         int block_height, int color_i)
{
  const char *who = "kvz_cuda_sao_reconstruct_color_host";
  if (bitdepth != 8 || !sao || block_width <= 0 || block_height <= 0 || block_width > 66 || block_height > 66 || color_i < 0 || color_i > 2) {
    fail_msg("kvz_cuda_sao_reconstruct_color_host: bad argument"); die(who);
  }
  // The filter reads one sample beyond the block on every side for the edge type (the caller owns that border,
  // sao-generic.c:138-148); stage the block with its border, filter the bordered plane as one CTU of a 1-CTU
  // picture and return the interior.  A band offset reads no neighbours.
  const int edge = sao->type == 2, b = edge ? 1 : 0, w = block_width + 2 * b, h = block_height + 2 * b;
  if (w > 64 || h > 64) {
    // larger than one CTU of the kernel with its border: filter row strips of up to 62 x 62
    for (int y0 = 0; y0 < block_height; y0 += 62)
      for (int x0 = 0; x0 < block_width; x0 += 62)
        kvz_cuda_sao_reconstruct_color_host(bitdepth, rec_data + (size_t)y0 * stride + x0, new_rec_data + (size_t)y0 * new_stride + x0,
                                            sao, stride, new_stride, block_width - x0 < 62 ? block_width - x0 : 62,
                                            block_height - y0 < 62 ? block_height - y0 : 62, color_i);
    return;
  }
  Stage st(who);
  uint8_t *in = st.in<uint8_t>((size_t)w * h);
  pack_rows(in, rec_data - (ptrdiff_t)b * stride - b, w, h, (size_t)stride);
  kvz_cuda_sao_params *prm = st.in<kvz_cuda_sao_params>(1);
  *prm = *sao;
  if (color_i == 2) {                                         // the kernel picks the V fields for its third plane only
    prm->band_position[0] = sao->band_position[1];
    for (int c = 0; c < 5; c++) prm->offsets[c] = sao->offsets[5 + c];
  }
  uint8_t *out = st.out<uint8_t>((size_t)w * h);
  st.upload(who);
  kvz_cuda_plane pi = { st.dev(in), w, h, w }, po = { st.dev(out), w, h, w };
  if (!launch_sao_apply(&pi, &po, 1, 8, st.dev(prm), nullptr, g_pc.stream)) die(who);
  st.download(who);
  for (int y = 0; y < block_height; y++) memcpy(new_rec_data + (size_t)y * new_stride, out + (size_t)(y + b) * w + b, (size_t)block_width);
}

int kvz_strategy_register_sao_cuda_with(void *opaque, uint8_t bitdepth, kvz_cuda_register_fn register_fn)
{
  if (!register_fn) return 0;
  if (bitdepth != 8) return 1;
  if (!cuda_usable()) return 0;
  int ok = 1;
  // sao_reconstruct_color / sao_band_ddistortion read encoder structures: their one-line wrappers live in the glue
  //  and call kvz_cuda_sao_{reconstruct_color,band_ddistortion}_host
  KVZ_REG("sao_edge_ddistortion", kvz_cuda_sao_edge_ddistortion);
  KVZ_REG("calc_sao_edge_dir", kvz_cuda_calc_sao_edge_dir);
  return ok;
}

void kvz_cuda_percall_counts(unsigned long long counts[7])
{
  std::lock_guard<std::mutex> guard(g_pc.lock);
  for (int i = 0; i < 7; i++) counts[i] = g_family_calls[i];
}

// ---------------- the two-argument forms of SURVEY.md 8(b) ----------------
// int kvz_strategy_register_<family>_cuda(void *opaque, uint8_t bitdepth), exactly like the reference's per-ISA
// registration functions (e.g. kvz_strategy_register_picture_avx2, strategies/avx2/picture-avx2.h).  The host's
// kvz_strategyselector_register (src/strategyselector.c:216-256) is looked up in the running process, so this
// library keeps no link-time dependency on the encoder; a host that hides the symbol uses the *_with forms.
static kvz_cuda_register_fn host_register_fn()
{
  static kvz_cuda_register_fn fn = nullptr;
  if (!fn) fn = reinterpret_cast<kvz_cuda_register_fn>(dlsym(RTLD_DEFAULT, "kvz_strategyselector_register"));
  if (!fn) fail_msg("kvz_strategy_register_*_cuda: kvz_strategyselector_register is not visible in this process "
                    "(use kvz_strategy_register_*_cuda_with)");
  return fn;
}
#define KVZ_REG2(family)                                                                      \
  int kvz_strategy_register_##family##_cuda(void *opaque, uint8_t bitdepth)                   \
  {                                                                                           \
    kvz_cuda_register_fn fn = host_register_fn();                                             \
    return fn ? kvz_strategy_register_##family##_cuda_with(opaque, bitdepth, fn) : 0;         \
  }
KVZ_REG2(picture) KVZ_REG2(dct) KVZ_REG2(quant) KVZ_REG2(intra) KVZ_REG2(ipol) KVZ_REG2(nal) KVZ_REG2(sao)

}  // extern "C"
