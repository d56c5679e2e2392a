// kvazaar_b200/csrc/enc_session.cu -- the encoder-session face (include/kvz_cuda_enc.h).
//
// A Kvazaar host keeps its pictures in host memory and searches one PU at a time, in CTU wavefront order, from
// all of its threadqueue workers (src/encoderstate.c:777-828, src/search_inter.c:1451-1524).  The session turns
// that call pattern into the batch face's: per frame the source and the new reference are uploaded once, the
// references are bordered and ONE launch produces the dense SAD maps of every PU of the frame against every
// reference (ime_sad.cu); each per-PU call then only prices / scans those maps and refines the winner
// (me_search.cu) on the calling thread's own stream, so the calls of the wavefront's CTU rows overlap on the GPU.
#include <sched.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <vector>

#include "../../include/kvz_cuda_enc.h"
#include "common.cuh"
#include "ime_sad.cuh"
#include "me_search.cuh"

namespace kvz {
namespace {

constexpr int kPad = 240;             // border of the resident reference planes: 104 + kFar (ime_sad box kernels)
constexpr int kFar = 136;             // largest |displacement| + 1 a box may reach
constexpr int kMaxSlots = 24;         // resident reference planes (the DPB holds at most 16 + the one being uploaded)

struct PlaneSlot {
  uint64_t key = 0;
  int poc = 0;
  bool live = false;
  uint64_t last_listed = 0;           // number of the newest frame whose reference list holds this picture
  uint8_t *d = nullptr;               // (w + 2 kPad) x (h + 2 kPad), pitch bytes per row
  // a picture that was still being reconstructed when a frame first listed it (--owf > 0): its 64x64 blocks are
  // uploaded one by one as the host's CTU dependencies guarantee them (ensure_blocks)
  bool partial = false;
  const uint8_t *host = nullptr;      // its luma in host memory
  int host_stride = 0;
  std::vector<uint8_t> have;          // per 64x64 block of the picture: uploaded (and complete on the device)
};

struct CurPlane {
  uint64_t key = 0, frame_num = ~0ull;
  uint8_t *d = nullptr;
  bool preloaded = false;             // d belongs to a Preload record (kvz_cuda_enc_preload_source), not to this slot
};

struct Preload {
  int64_t pts;
  uint8_t *d;
};

struct FrameCtx {
  uint64_t id = ~0ull;
  MeFrame f;
  uint8_t *maps = nullptr;
  size_t map_bytes = 0;               // allocated
  cudaEvent_t ready = nullptr;
  cudaEvent_t dense0 = nullptr, dense1 = nullptr;    // round the dense SAD launch of this frame
  bool dense_pending = false;
  unsigned long long dense_absdiffs = 0, dense_map_bytes = 0;
  std::atomic<int> inflight{0};
  int box[KVZ_CUDA_MAX_REFS][4];      // x0, y0, n_dx, n_dy per reference
  // references that were in flight when the frame began: their maps are made CTU by CTU (ctu_begin)
  int n_partial = 0;
  PlaneSlot *partial_slot[KVZ_CUDA_MAX_REFS];
  ImeSadParams partial_ps[KVZ_CUDA_MAX_REFS];
  int partial_box[KVZ_CUDA_MAX_REFS][4];
};

// per calling thread: a stream and the staging of one PU, its jobs and results
// Per calling thread: kStreams streams and a ring of kSlots launch slots in pinned, mapped host memory.  A slot holds the
// results of the up to kMePuArgsMax PUs of one launch and one completion word per PU; slot i always uses stream
// i % kStreams, so a slot is not written by two launches at once.  A ticket = slot | generation << 8.
constexpr int kStreams = 4, kSlots = 64;
constexpr size_t kSlotResults = (size_t)kMePuArgsMax * KVZ_CUDA_MAX_REFS * sizeof(kvz_cuda_me_result);
constexpr size_t kSlotBytes = kSlotResults + 64;                      // + the completion words on a line of their own
constexpr size_t kStageBytes = kSlots * kSlotBytes;

struct Worker {
  int device = -1;
  cudaStream_t streams[kStreams] = {};
  cudaEvent_t ctu_ready = nullptr;    // --owf: the blocks and maps ctu_begin prepared on streams[0]
  uint8_t *h = nullptr, *d = nullptr;
  const void *seen_enc = nullptr;
  uint64_t seen_serial = 0;           // a later session may get the address of a closed one
  uint64_t seen_ctx[kStreams];        // the frame context whose `ready` event the stream has waited for
  int seen_ctu[kStreams];             // the CTU of that context whose ctu_ready the stream has waited for
  int cur_ctu = -1;                   // the CTU this thread last prepared (ctu_begin)
  uint64_t cur_ctx = ~0ull;
  FrameCtx *ctx_of[64] = {};          // frame handle -> context, as last looked up under the session's lock
  uint32_t next_slot = 0;
  uint32_t gen[kSlots] = {};          // generation of the slot's last launch (the value its completion words receive)
  uint8_t n_pus[kSlots] = {};
  uint8_t n_jobs[kSlots][kMePuArgsMax] = {};
  uint8_t size_class[kSlots][kMePuArgsMax] = {};
  // a host creates its worker threads per encoder: what a thread holds goes when the thread ends (launches made ahead
  // and never collected may still be running: the streams are drained first).  After the runtime has shut down the
  // calls fail and are ignored.
  ~Worker()
  {
    if (device < 0) return;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return; }
    for (auto &s : streams)
      if (s) { cudaStreamSynchronize(s); cudaStreamDestroy(s); s = nullptr; }
    if (ctu_ready) cudaEventDestroy(ctu_ready);
    if (h) cudaFreeHost(h);
    cudaGetLastError();
  }
};
thread_local Worker t_worker;

// How a thread waits for a search it needs (KVZ_CUDA_WAIT): "spin" (default) = poll the completion word the kernel
// writes into mapped memory after its results; "yield" = the same with sched_yield() between polls, for hosts that run
// more worker threads than cores.
enum WaitMode { kWaitSpin = 1, kWaitYield = 2 };
WaitMode wait_mode()
{
  static const WaitMode m = [] {
    const char *v = getenv("KVZ_CUDA_WAIT");
    return v && !strcmp(v, "yield") ? kWaitYield : kWaitSpin;
  }();
  return m;
}

}  // namespace
}  // namespace kvz

using namespace kvz;

struct kvz_cuda_enc {
  kvz_cuda_enc_cfg cfg;
  uint64_t serial = 0;                // unique per open
  int pitch = 0;                      // bytes per row of the bordered planes
  size_t plane_bytes = 0;
  std::mutex lock;
  cudaStream_t stream = nullptr;      // frame-level work: uploads, borders, dense maps
  uint8_t *d_tmp = nullptr;           // un-bordered upload of a reference
  PlaneSlot slots[kMaxSlots];
  std::vector<CurPlane> cur;
  std::vector<FrameCtx *> ctxs;
  std::vector<Preload> preloads;
  std::vector<uint8_t *> spare_planes;                 // source planes handed back by consumed preloads
  std::atomic<unsigned long long> stats[8];
  std::atomic<unsigned long long> meas[8];
  // KVZ_CUDA_STATS: wall time of kvz_cuda_enc_search_pu by PU size (8, 16, 32, 64) and by whether every window of
  // every job lay inside its reference's box
  bool timing = false;
  std::atomic<unsigned long long> t_ns[4][2], t_calls[4][2];
  std::atomic<unsigned long long> t_host[4][2];      // KVZ_CUDA_STATS: {ns, calls} of ctu_begin, the PU launch call, frame_begin, stream waits
};

namespace {

int fail_enc(const char *what) { return fail_msg(what); }

// One box per reference: the zero window plus the windows round the vectors the hints predict.  Window centres are
// the whole-sample parts of the frame's merge candidates and colocated vectors (search_inter.c:898-955) -- vectors of
// ANY reference, so one box serves all references.  The hints are trimmed to their 0.5th..99.5th percentile per axis
// (a few wild vectors must not blow the box up) and widened by kBoxMargin, because the vectors this frame finds
// scatter round the ones its references found.  Caps: the dense kernel stages n_dy + 63 reference rows in shared
// memory (ime_box_max_rows(), 177) and the box maps are at most 272 displacements wide.
constexpr int kBoxMargin = 24;
void choose_box(int R, const int16_t *hints, int n_hints, int shrink_pct, int box[4])
{
  static const int margin = getenv("KVZ_CUDA_BOX_MARGIN") ? atoi(getenv("KVZ_CUDA_BOX_MARGIN")) : kBoxMargin;   // experiments
  int lox = 0, hix = 0, loy = 0, hiy = 0;
  if (hints && n_hints > 0) {
    std::vector<int> xs(n_hints), ys(n_hints);
    for (int i = 0; i < n_hints; i++) { xs[i] = hints[2 * i] >> 2; ys[i] = hints[2 * i + 1] >> 2; }
    std::sort(xs.begin(), xs.end());
    std::sort(ys.begin(), ys.end());
    const int a = n_hints / 200, b = n_hints - 1 - n_hints / 200;
    lox = std::min(lox, xs[a]); hix = std::max(hix, xs[b]);
    loy = std::min(loy, ys[a]); hiy = std::max(hiy, ys[b]);
  }
  lox -= margin; loy -= margin; hix += margin; hiy += margin;
  const int cap_x = std::max(0, std::min((272 - 8 - (2 * R + 1)) / 2, kFar - 8 - R)) * shrink_pct / 100;
  static const int max_rows = ime_box_max_rows();
  const int cap_y = std::max(0, (max_rows - (2 * R + 1)) / 2) * shrink_pct / 100;
  lox = std::max(lox, -cap_x); hix = std::min(hix, cap_x);
  loy = std::max(loy, -cap_y); hiy = std::min(hiy, cap_y);
  box[0] = lox - R; box[1] = loy - R;
  box[2] = ((hix + R) - (lox - R) + 1 + 7) & ~7;
  box[3] = (hiy + R) - (loy - R) + 1;
}

size_t level_entries(int tile_w, int tile_h, int level)
{
  const size_t ctus = (size_t)((tile_w + 63) / 64) * ((tile_h + 63) / 64);
  const size_t n = 8 >> level;
  return ctus * n * n;
}

size_t box_bytes(int tile_w, int tile_h, const int box[4])
{
  size_t b = 0;
  for (int l = 0; l < 4; l++) {
    size_t lb = level_entries(tile_w, tile_h, l) * (size_t)box[2] * box[3] * (l < 2 ? 2 : 4);
    b += (lb + 255) & ~(size_t)255;
  }
  return b;
}

// the window centres search_mv_full visits for one job (search_inter.c:891-955; the list scan_job builds on the device):
// true when all of them lie inside the reference's box, i.e. the job is served from the dense maps alone
bool job_in_box(const FrameCtx &c, int R, const kvz_cuda_me_pu &pu, const kvz_cuda_me_job &job)
{
  const int *b = c.box[job.ref_idx];
  int lox = 0, hix = 0, loy = 0, hiy = 0;
  auto add = [&](int x, int y) { lox = std::min(lox, x); hix = std::max(hix, x); loy = std::min(loy, y); hiy = std::max(hiy, y); };
  const int ex = job.extra_mv[0] >> 2, ey = job.extra_mv[1] >> 2;
  bool in_merge = false;
  for (int i = 0; i < pu.num_merge_cand; i++) {
    const kvz_cuda_merge_cand &m = pu.merge_cand[i];
    if (m.dir == 3) continue;
    in_merge |= ((m.mv[m.dir - 1][0] + 2) >> 2) == ex && ((m.mv[m.dir - 1][1] + 2) >> 2) == ey;
    add(m.mv[m.dir - 1][0] >> 2, m.mv[m.dir - 1][1] >> 2);
  }
  if (!in_merge) add(ex, ey);
  return lox - R >= b[0] && hix + R < b[0] + b[2] && loy - R >= b[1] && hiy + R < b[1] + b[3];
}

// the duration of a context's dense SAD launch, once it has finished (CUDA events on the launching stream)
void harvest_dense(kvz_cuda_enc *e, FrameCtx *c, bool wait)
{
  if (!c->dense_pending) return;
  if (wait) cudaEventSynchronize(c->dense1);
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, c->dense0, c->dense1) != cudaSuccess) { cudaGetLastError(); return; }
  c->dense_pending = false;
  e->meas[3]++;
  e->meas[4] += (unsigned long long)((double)ms * 1e6);
  e->meas[5] += c->dense_absdiffs;
  e->meas[6] += c->dense_map_bytes;
}

// --owf > 0.  Before a thread searches its first PU of CTU (cx, cy) of a frame that lists pictures still in flight:
// (1) every 64x64 block of those pictures that the motion vector restriction lets this CTU reach (fracmv_within_tile,
// search_inter.c:87-140: at most max_ref_lcu_down rows below, right + down at most max_ref_lcu_right + max_ref_lcu_down
// blocks away) is brought to the device if it is not there yet.  The host's CTU job dependency (encoderstate.c:786-801)
// guarantees those blocks are reconstructed; the last samples of a block wait for its right / lower neighbour's in-loop
// filters (the restriction's margins keep vectors off them until then), so every block is uploaded with 16 samples of
// its left / upper neighbours: a block's late strips arrive with the neighbour that finishes them.  (2) the dense SAD maps
// of this CTU against those pictures, on the thread's own stream, in front of its PU searches.  Samples that are not
// final yet only ever feed displacements the restriction excludes.
int ctu_begin(kvz_cuda_enc *e, FrameCtx *ctx, int cx, int cy, cudaStream_t stream)
{
  const kvz_cuda_enc_cfg &cfg = e->cfg;
  const int W = cfg.width, H = cfg.height;
  const int ctus_x = (W + 63) / 64, ctus_y = (H + 63) / 64;
  const int down = cfg.max_ref_lcu_down, reach = cfg.max_ref_lcu_down + cfg.max_ref_lcu_right;
  // picture-level CTU position (tiles with --owf are refused by the glue; tile offsets are multiples of 64 anyway)
  const int pcx = cx + ctx->f.tile_x / 64, pcy = cy + ctx->f.tile_y / 64;
  bool uploaded = false;
  int marks[64][2], n_marks = 0;
  PlaneSlot *mark_slot[64];
  for (int i = 0; i < ctx->n_partial; i++) {
    PlaneSlot *s = ctx->partial_slot[i];
    if (!s->partial) continue;                       // completed by a later frame's begin in the meantime
    for (int ry = 0; ry <= std::min(pcy + down, ctus_y - 1); ry++) {
      const int rx_max = std::min(pcx + reach - (ry - pcy), ctus_x - 1);
      for (int rx = 0; rx <= rx_max; rx++) {
        if (s->have[(size_t)ry * ctus_x + rx]) continue;
        const int x0 = std::max(0, rx * 64 - 16), y0 = std::max(0, ry * 64 - 16);
        const int x1 = std::min(W, (rx + 1) * 64), y1 = std::min(H, (ry + 1) * 64);
        KVZ_CUDA_TRY(cudaMemcpy2DAsync(s->d + (size_t)(y0 + kPad) * e->pitch + x0 + kPad, e->pitch,
                                       s->host + (size_t)y0 * s->host_stride + x0, s->host_stride, x1 - x0, y1 - y0,
                                       cudaMemcpyHostToDevice, stream));
        if (!launch_pad_patch(s->d, W, H, kPad, e->pitch, x0, y0, x1 - x0, y1 - y0, 8, stream)) return 0;
        e->meas[0] += (unsigned long long)(x1 - x0) * (y1 - y0);
        uploaded = true;
        if (n_marks == 64) {                         // flush: complete on the device before anybody is told
          KVZ_CUDA_TRY(cudaStreamSynchronize(stream));
          for (int k = 0; k < n_marks; k++) mark_slot[k]->have[(size_t)marks[k][1] * ctus_x + marks[k][0]] = 1;
          n_marks = 0;
        }
        marks[n_marks][0] = rx; marks[n_marks][1] = ry; mark_slot[n_marks] = s; n_marks++;
      }
    }
  }
  if (uploaded) {
    // other frames in flight may list the same picture: a block counts as present only once it is complete on the device
    KVZ_CUDA_TRY(cudaStreamSynchronize(stream));
    for (int k = 0; k < n_marks; k++) mark_slot[k]->have[(size_t)marks[k][1] * ctus_x + marks[k][0]] = 1;
  }
  if (ctx->f.ime_search == KVZ_CUDA_IME_FULL && ctx->n_partial > 0) {
    const int ctu = cy * ctx->f.ctus_x + cx;
    for (int r0 = 0; r0 < ctx->n_partial; r0 += kImeMaxMulti) {
      const int n = std::min(kImeMaxMulti, ctx->n_partial - r0);
      if (!launch_ime_sad_box_multi(ctx->partial_ps + r0, ctx->partial_box + r0, n, stream, ctu)) return 0;
    }
  }
  return 1;
}

bool worker_ready(kvz_cuda_enc *e)
{
  Worker &w = t_worker;
  if (w.device == e->cfg.device && w.streams[0]) return true;
  if (cudaSetDevice(e->cfg.device) != cudaSuccess) return false;      // every calling thread binds the session's device
  for (int i = 0; i < kStreams; i++) {
    if (cudaStreamCreateWithFlags(&w.streams[i], cudaStreamNonBlocking) != cudaSuccess) return false;
    w.seen_ctx[i] = ~0ull; w.seen_ctu[i] = -1;
  }
  if (cudaEventCreateWithFlags(&w.ctu_ready, cudaEventDisableTiming) != cudaSuccess) return false;
  // pinned AND mapped: the kernel writes its results and completion words straight into this host buffer (one small
  // PCIe write per PU instead of a copy command); the PUs and their jobs travel as kernel arguments
  if (cudaHostAlloc(&w.h, kStageBytes, cudaHostAllocMapped) != cudaSuccess) return false;
  if (cudaHostGetDevicePointer(reinterpret_cast<void **>(&w.d), w.h, 0) != cudaSuccess) return false;
  memset(w.h, 0, kStageBytes);
  w.device = e->cfg.device;
  return true;
}

}  // namespace

extern "C" {

kvz_cuda_enc *kvz_cuda_enc_open(const kvz_cuda_enc_cfg *cfg)
{
  if (!cfg) { fail_enc("kvz_cuda_enc_open: null cfg"); return nullptr; }
  if (cfg->bitdepth != 8) { fail_enc("kvz_cuda_enc_open: only bitdepth 8 is implemented on this face"); return nullptr; }
  if (cfg->width <= 0 || cfg->height <= 0 || (cfg->width & 7) || (cfg->height & 7)) {
    fail_enc("kvz_cuda_enc_open: width and height must be positive multiples of 8 (the CLI pads its input, encmain.c:80)");
    return nullptr;
  }
  if (cfg->ime_search < KVZ_CUDA_IME_FULL || cfg->ime_search > KVZ_CUDA_IME_TZ) { fail_enc("kvz_cuda_enc_open: bad ime_search"); return nullptr; }
  if (cfg->range != 8 && cfg->range != 16 && cfg->range != 32 && cfg->range != 64) { fail_enc("kvz_cuda_enc_open: range must be 8, 16, 32 or 64"); return nullptr; }
  if (cfg->fme_level < 0 || cfg->fme_level > 4) { fail_enc("kvz_cuda_enc_open: fme_level must be 0..4"); return nullptr; }
  if (!kvz_cuda_init(cfg->device)) return nullptr;
  kvz_cuda_enc *e = new kvz_cuda_enc();
  static std::atomic<uint64_t> serials{0};
  e->serial = ++serials;
  e->cfg = *cfg;
  if (e->cfg.max_map_mbytes <= 0) e->cfg.max_map_mbytes = 98304;       // of 180 GB; shared by the owf + 1 frames in flight
  for (auto &s : e->stats) s = 0;
  for (auto &s : e->meas) s = 0;
  for (int a = 0; a < 4; a++) for (int b = 0; b < 2; b++) { e->t_ns[a][b] = 0; e->t_calls[a][b] = 0; e->t_host[a][b] = 0; }
  e->timing = getenv("KVZ_CUDA_STATS") != nullptr;
  e->pitch = (cfg->width + 2 * kPad + 63) & ~63;
  e->plane_bytes = (size_t)e->pitch * (cfg->height + 2 * kPad);
  if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaMalloc(&e->d_tmp, (size_t)cfg->width * cfg->height) != cudaSuccess) {
    fail("kvz_cuda_enc_open", cudaGetLastError(), __FILE__, __LINE__);
    delete e;
    return nullptr;
  }
  return e;
}

void kvz_cuda_enc_close(kvz_cuda_enc *e)
{
  if (!e) return;
  cudaSetDevice(e->cfg.device);
  cudaDeviceSynchronize();
  for (auto &s : e->slots) if (s.d) cudaFree(s.d);
  for (auto &c : e->cur) if (c.d && !c.preloaded) cudaFree(c.d);
  for (auto &p : e->preloads) cudaFree(p.d);
  for (auto *p : e->spare_planes) cudaFree(p);
  for (FrameCtx *c : e->ctxs) {
    if (c->maps) cudaFree(c->maps);
    if (c->ready) cudaEventDestroy(c->ready);
    if (c->dense0) cudaEventDestroy(c->dense0);
    if (c->dense1) cudaEventDestroy(c->dense1);
    delete c;
  }
  if (e->d_tmp) cudaFree(e->d_tmp);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

int kvz_cuda_enc_frame_begin(kvz_cuda_enc *e, const kvz_cuda_enc_frame *fr)
{
  if (!e || !fr || !fr->src_luma) { fail_enc("kvz_cuda_enc_frame_begin: null argument"); return -1; }
  if (fr->n_refs < 1 || fr->n_refs > KVZ_CUDA_MAX_REFS) { fail_enc("kvz_cuda_enc_frame_begin: bad n_refs"); return -1; }
  if ((fr->tile_x & 63) || (fr->tile_y & 63) || (fr->tile_w & 7) || (fr->tile_h & 7) || fr->tile_w <= 0 || fr->tile_h <= 0 ||
      fr->tile_x + fr->tile_w > e->cfg.width || fr->tile_y + fr->tile_h > e->cfg.height) {
    fail_enc("kvz_cuda_enc_frame_begin: bad tile rectangle");
    return -1;
  }
  std::lock_guard<std::mutex> guard(e->lock);
  for (size_t i = 0; i < e->ctxs.size(); i++)
    if (e->ctxs[i]->id == fr->frame_id) return (int)i;              // another thread began this (frame, tile)
  struct FbTimer {
    kvz_cuda_enc *e; struct timespec t0;
    explicit FbTimer(kvz_cuda_enc *e_) : e(e_) { if (e->timing) clock_gettime(CLOCK_MONOTONIC, &t0); }
    ~FbTimer() {
      if (!e->timing) return;
      struct timespec t1; clock_gettime(CLOCK_MONOTONIC, &t1);
      e->t_host[2][0] += (unsigned long long)((t1.tv_sec - t0.tv_sec) * 1000000000ll + (t1.tv_nsec - t0.tv_nsec));
      e->t_host[2][1]++;
    }
  } fb_timer(e);
  if (cudaSetDevice(e->cfg.device) != cudaSuccess) { fail("cudaSetDevice", cudaGetLastError(), __FILE__, __LINE__); return -1; }
  const kvz_cuda_enc_cfg &cfg = e->cfg;
  const uint64_t frame_num = fr->frame_id >> 16;

  // ---- source plane: one upload per frame, shared by its tiles ----
  // planes, contexts and slots of frame n may be reused once frame n + owf + 1 begins: the encoder has owf + 1 states in
  // a ring (kvazaar.c:96-126), so that frame's state is being reused, i.e. the frame is complete
  const uint64_t window = (uint64_t)(cfg.owf > 0 ? cfg.owf : 0) + 1;
  CurPlane *cur = nullptr;
  for (auto &c : e->cur) if (c.d && c.key == fr->src_key && c.frame_num == frame_num) cur = &c;
  if (!cur) {
    for (auto &c : e->cur) if (c.frame_num == ~0ull || c.frame_num + window <= frame_num) { cur = &c; break; }
    if (!cur) { e->cur.emplace_back(); cur = &e->cur.back(); }
    if (cur->preloaded) { e->spare_planes.push_back(cur->d); cur->d = nullptr; cur->preloaded = false; }
    cur->key = fr->src_key; cur->frame_num = frame_num;
    bool found = false;
    for (size_t i = 0; i < e->preloads.size(); i++) {
      if (e->preloads[i].pts != fr->src_pts) continue;
      if (cur->d) e->spare_planes.push_back(cur->d);
      cur->d = e->preloads[i].d; cur->preloaded = true;
      e->preloads.erase(e->preloads.begin() + i);
      e->meas[7]++;
      found = true;
      break;
    }
    if (!found) {
      if (!cur->d && cudaMalloc(&cur->d, (size_t)cfg.width * cfg.height) != cudaSuccess) {
        fail("cudaMalloc(source plane)", cudaGetLastError(), __FILE__, __LINE__); return -1;
      }
      if (cudaMemcpy2DAsync(cur->d, cfg.width, fr->src_luma, fr->src_stride, cfg.width, cfg.height, cudaMemcpyHostToDevice,
                            e->stream) != cudaSuccess) { fail("upload of the source plane", cudaGetLastError(), __FILE__, __LINE__); return -1; }
      e->meas[0] += (unsigned long long)cfg.width * cfg.height;
    }
  }

  // ---- reference planes: keep what some frame in flight lists, upload and border what is new ----
  for (auto &s : e->slots) {
    if (!s.live) continue;
    for (int r = 0; r < fr->n_refs; r++)
      if (fr->refs[r].key == s.key && fr->refs[r].poc == s.poc && frame_num > s.last_listed) s.last_listed = frame_num;
    if (s.last_listed + window <= frame_num) s.live = false;       // no frame that can still be searching lists it
  }
  const int pic_ctus_x = (cfg.width + 63) / 64, pic_ctus_y = (cfg.height + 63) / 64;
  PlaneSlot *ref_slot[KVZ_CUDA_MAX_REFS];
  for (int r = 0; r < fr->n_refs; r++) {
    const kvz_cuda_enc_ref &rr = fr->refs[r];
    if (!rr.luma) { fail_enc("kvz_cuda_enc_frame_begin: null reference plane"); return -1; }
    PlaneSlot *s = nullptr;
    for (auto &q : e->slots) if (q.live && q.key == rr.key && q.poc == rr.poc) s = &q;
    const bool fresh = s == nullptr;
    if (fresh) {
      for (auto &q : e->slots) if (!q.live) { s = &q; break; }
      if (!s) { fail_enc("kvz_cuda_enc_frame_begin: out of reference plane slots"); return -1; }
      if (!s->d && cudaMalloc(&s->d, e->plane_bytes) != cudaSuccess) {
        fail("cudaMalloc(reference plane)", cudaGetLastError(), __FILE__, __LINE__); return -1;
      }
      s->key = rr.key; s->poc = rr.poc; s->live = true; s->last_listed = frame_num;
      s->partial = rr.in_flight != 0;
      s->host = static_cast<const uint8_t *>(rr.luma); s->host_stride = rr.stride;
      if (s->partial) s->have.assign((size_t)pic_ctus_x * pic_ctus_y, 0);
    }
    // whole picture at once: a new complete picture, or one that was in flight when it was first listed and is done now
    if ((fresh && !s->partial) || (!fresh && s->partial && !rr.in_flight)) {
      if (cudaMemcpy2DAsync(e->d_tmp, cfg.width, rr.luma, rr.stride, cfg.width, cfg.height, cudaMemcpyHostToDevice,
                            e->stream) != cudaSuccess) { fail("upload of a reference plane", cudaGetLastError(), __FILE__, __LINE__); return -1; }
      if (!launch_pad_plane(e->d_tmp, cfg.width, cfg.height, cfg.width, s->d, kPad, e->pitch, 8, e->stream)) return -1;
      if (s->partial) {
        // threads of an older frame may be bringing this picture's blocks in one by one: they stop doing so when the
        // flag falls, so the whole picture has to be on the device before it does
        if (cudaStreamSynchronize(e->stream) != cudaSuccess) { fail("stream sync", cudaGetLastError(), __FILE__, __LINE__); return -1; }
        s->partial = false;
      }
      e->stats[2]++;
      e->meas[0] += (unsigned long long)cfg.width * cfg.height;
    }
    ref_slot[r] = s;
  }

  // ---- a context: one of an older frame whose searches are over, or a new one ----
  FrameCtx *ctx = nullptr;
  int handle = -1;
  for (size_t i = 0; i < e->ctxs.size(); i++) {
    FrameCtx *c = e->ctxs[i];
    if (c->id != ~0ull && (c->id >> 16) + window <= frame_num && c->inflight.load() == 0) { ctx = c; handle = (int)i; break; }
  }
  if (!ctx) {
    ctx = new FrameCtx();
    if (cudaEventCreateWithFlags(&ctx->ready, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreate(&ctx->dense0) != cudaSuccess || cudaEventCreate(&ctx->dense1) != cudaSuccess) {
      fail("cudaEventCreate", cudaGetLastError(), __FILE__, __LINE__); delete ctx; return -1;
    }
    e->ctxs.push_back(ctx);
    handle = (int)e->ctxs.size() - 1;
  }
  ctx->id = ~0ull;                                     // not findable until it is complete
  harvest_dense(e, ctx, true);

  // ---- boxes, within the map budget ----
  const bool full = cfg.ime_search == KVZ_CUDA_IME_FULL;
  size_t need = 0;
  if (full) {
    int shrink = 100;                                  // percent of the caps: boxes shrink until the maps fit the budget
    for (;;) {
      need = 0;
      for (int r = 0; r < fr->n_refs; r++) {
        choose_box(cfg.range, fr->refs[r].mv_hints, fr->refs[r].n_mv_hints, shrink, ctx->box[r]);
        need += box_bytes(fr->tile_w, fr->tile_h, ctx->box[r]);
      }
      if (need <= ((size_t)cfg.max_map_mbytes << 20) / window || shrink == 0) break;     // owf + 1 frames hold maps at once
      shrink = shrink * 3 / 4;
    }
    if (need > ctx->map_bytes) {
      if (ctx->maps) { cudaStreamSynchronize(e->stream); cudaFree(ctx->maps); ctx->maps = nullptr; ctx->map_bytes = 0; }
      if (cudaMalloc(&ctx->maps, need) != cudaSuccess) { fail("cudaMalloc(SAD maps)", cudaGetLastError(), __FILE__, __LINE__); return -1; }
      ctx->map_bytes = need;
    }
  }

  // ---- frame description of the decision kernels ----
  MeFrame &f = ctx->f;
  memset(&f, 0, sizeof(f));
  f.cur = cur->d; f.cur_pitch = cfg.width;
  f.width = cfg.width; f.height = cfg.height;
  f.tile_x = fr->tile_x; f.tile_y = fr->tile_y; f.tile_w = fr->tile_w; f.tile_h = fr->tile_h;
  f.ctus_x = (fr->tile_w + 63) / 64;
  f.range = cfg.range; f.map_pitch = ime_map_pitch(cfg.range);
  f.ime_search = cfg.ime_search; f.me_early_termination = cfg.me_early_termination; f.me_max_steps = cfg.me_max_steps;
  f.fme_level = cfg.fme_level; f.owf = cfg.owf; f.wpp = cfg.wpp; f.sao_type = cfg.sao_type;
  f.deblock_enable = cfg.deblock_enable; f.mv_constraint = cfg.mv_constraint;
  f.max_ref_lcu_right = cfg.max_ref_lcu_right; f.max_ref_lcu_down = cfg.max_ref_lcu_down;
  f.n_refs = fr->n_refs;
  memcpy(f.ref_LX, fr->ref_LX, sizeof(f.ref_LX));
  ImeSadParams ps[KVZ_CUDA_MAX_REFS];
  int ps_box[KVZ_CUDA_MAX_REFS][4];
  int n_full = 0;
  uint8_t *at = ctx->maps;
  unsigned long long disp = 0, disp_full = 0;
  ctx->n_partial = 0;
  for (int r = 0; r < fr->n_refs; r++) {
    f.refs[r].plane = ref_slot[r]->d; f.refs[r].pitch = e->pitch; f.refs[r].pad = kPad;
    f.refs[r].n_sets = 0;
    if (!full) {
      // the heuristic searches read the reference plane directly: a picture in flight needs its blocks too
      if (ref_slot[r]->partial) { ctx->partial_slot[ctx->n_partial] = ref_slot[r]; ctx->n_partial++; }
      continue;
    }
    const int *b = ctx->box[r];
    void *lv[4];
    for (int l = 0; l < 4; l++) {
      lv[l] = at;
      const size_t lb = level_entries(fr->tile_w, fr->tile_h, l) * (size_t)b[2] * b[3] * (l < 2 ? 2 : 4);
      at += (lb + 255) & ~(size_t)255;
    }
    MeMapSet &ms = f.refs[r].sets[0];
    f.refs[r].n_sets = 1;
    ms.cx = b[0]; ms.cy = b[1]; ms.n_dx = b[2]; ms.n_dy = b[3];
    ms.sad8 = (const uint16_t *)lv[0]; ms.sad16 = (const uint16_t *)lv[1];
    ms.sad32 = (const uint32_t *)lv[2]; ms.sad64 = (const uint32_t *)lv[3];
    const bool part = ref_slot[r]->partial;
    ImeSadParams &p = part ? ctx->partial_ps[ctx->n_partial] : ps[n_full];
    memset(&p, 0, sizeof(p));
    p.cur = cur->d; p.ref = ref_slot[r]->d;
    p.width = cfg.width; p.height = cfg.height;
    p.cur_pitch = cfg.width; p.ref_pitch = e->pitch; p.pad = kPad;
    p.tile_x = fr->tile_x; p.tile_y = fr->tile_y; p.tile_w = fr->tile_w; p.tile_h = fr->tile_h;
    p.ctus_x = (fr->tile_w + 63) / 64; p.ctus_y = (fr->tile_h + 63) / 64;
    p.map_pitch = b[2];
    p.sad8 = (uint16_t *)lv[0]; p.sad16 = (uint16_t *)lv[1]; p.sad32 = (uint32_t *)lv[2]; p.sad64 = (uint32_t *)lv[3];
    disp += (unsigned long long)b[2] * b[3];
    if (part) {
      memcpy(ctx->partial_box[ctx->n_partial], b, sizeof(int) * 4);
      ctx->partial_slot[ctx->n_partial] = ref_slot[r];
      ctx->n_partial++;
    } else {
      memcpy(ps_box[n_full], b, sizeof(int) * 4);
      disp_full += (unsigned long long)b[2] * b[3];
      n_full++;
    }
  }
  if (full && n_full > 0) {
    cudaEventRecord(ctx->dense0, e->stream);
    for (int r0 = 0; r0 < n_full; r0 += kImeMaxMulti) {
      const int n = std::min(kImeMaxMulti, n_full - r0);
      if (!launch_ime_sad_box_multi(ps + r0, ps_box + r0, n, e->stream)) return -1;
    }
    cudaEventRecord(ctx->dense1, e->stream);
    ctx->dense_pending = true;
    ctx->dense_absdiffs = level_entries(fr->tile_w, fr->tile_h, 0) * 64ull * disp_full;
    ctx->dense_map_bytes = need;
  }
  if (cudaEventRecord(ctx->ready, e->stream) != cudaSuccess) { fail("cudaEventRecord", cudaGetLastError(), __FILE__, __LINE__); return -1; }
  e->stats[0]++;
  e->stats[3] = need;
  e->stats[4] = disp;
  ctx->id = fr->frame_id;
  return handle;
}

int kvz_cuda_enc_search_begin(kvz_cuda_enc *e, int handle, int lambda_sqrt_i, int n_pus, const kvz_cuda_me_pu *pus,
                              const kvz_cuda_me_job *jobs, int *ticket)
{
  if (!e || !pus || !jobs || !ticket) return fail_enc("kvz_cuda_enc_search_begin: null argument");
  if (n_pus < 1 || n_pus > kMePuArgsMax) return fail_enc("kvz_cuda_enc_search_begin: 1..8 PUs per launch");
  // contexts are created once and live until the session closes, so a thread may keep the pointer it looked up under
  // the lock (the vector that holds them may grow); a context is recycled only for a frame owf + 1 later, when the
  // host has long stopped searching this one
  FrameCtx *ctx = nullptr;
  Worker &w0 = t_worker;
  if (w0.seen_enc == e && w0.seen_serial == e->serial && handle >= 0 && handle < 64) ctx = w0.ctx_of[handle];
  if (!ctx) {
    std::lock_guard<std::mutex> guard(e->lock);
    if (handle >= 0 && handle < (int)e->ctxs.size()) ctx = e->ctxs[handle];
  }
  if (!ctx || ctx->id == ~0ull) return fail_enc("kvz_cuda_enc_search_begin: bad frame handle");
  ctx->inflight++;
  struct Release { FrameCtx *c; ~Release() { c->inflight--; } } release{ctx};
  const int ctu_x = pus[0].x >> 6, ctu_y = pus[0].y >> 6;
  unsigned long long n_jobs_total = 0;
  for (int p = 0; p < n_pus; p++) {
    const kvz_cuda_me_pu &pu = pus[p];
    const int nj = pu.num_jobs;
    if (nj < 1 || nj > KVZ_CUDA_MAX_REFS) return fail_enc("kvz_cuda_enc_search_begin: bad num_jobs");
    if ((pu.w & 7) || (pu.h & 7) || pu.w == 0 || pu.h == 0 || pu.w > 64 || pu.h > 64)
      return fail_enc("kvz_cuda_enc_search_begin: PU sides must be multiples of 8, at most 64 (SMP/AMP shapes 4, 12 are not searched here)");
    if ((pu.x >> 6) != ctu_x || (pu.y >> 6) != ctu_y) return fail_enc("kvz_cuda_enc_search_begin: the PUs of one launch must lie in one CTU");
    for (int i = 0; i < nj; i++)
      if (jobs[(size_t)p * KVZ_CUDA_MAX_REFS + i].ref_idx >= ctx->f.n_refs)
        return fail_enc("kvz_cuda_enc_search_begin: ref_idx outside the frame's reference list");
    n_jobs_total += (unsigned long long)nj;
  }
  if (!worker_ready(e)) return fail("per-thread streams / staging", cudaGetLastError(), __FILE__, __LINE__);
  Worker &w = t_worker;
  if (ctx->f.ime_search == KVZ_CUDA_IME_FULL)
    for (int p = 0; p < n_pus; p++)
      for (int i = 0; i < pus[p].num_jobs; i++)
        if (!job_in_box(*ctx, e->cfg.range, pus[p], jobs[(size_t)p * KVZ_CUDA_MAX_REFS + i])) e->stats[6]++;

  const uint64_t tag = ((uint64_t)handle << 48) ^ ctx->id;
  const bool same_session = w.seen_enc == e && w.seen_serial == e->serial;
  if (!same_session) for (auto &c : w.ctx_of) c = nullptr;
  if (handle < 64) w.ctx_of[handle] = ctx;
  if (!same_session || w.cur_ctx != tag) {
    w.seen_enc = e; w.seen_serial = e->serial; w.cur_ctx = tag; w.cur_ctu = -1;
    for (int i = 0; i < kStreams; i++) { w.seen_ctx[i] = ~0ull; w.seen_ctu[i] = -1; }
  }
  const int ctu = ctu_y * ctx->f.ctus_x + ctu_x;
  if (ctx->n_partial > 0 && w.cur_ctu != ctu) {            // --owf: references in flight are prepared CTU by CTU
    if (w.seen_ctx[0] != tag) { KVZ_CUDA_TRY(cudaStreamWaitEvent(w.streams[0], ctx->ready, 0)); w.seen_ctx[0] = tag; }
    struct timespec c0, c1;
    if (e->timing) clock_gettime(CLOCK_MONOTONIC, &c0);
    if (!ctu_begin(e, ctx, ctu_x, ctu_y, w.streams[0])) return 0;
    KVZ_CUDA_TRY(cudaEventRecord(w.ctu_ready, w.streams[0]));
    if (e->timing) {
      clock_gettime(CLOCK_MONOTONIC, &c1);
      e->t_host[0][0] += (unsigned long long)((c1.tv_sec - c0.tv_sec) * 1000000000ll + (c1.tv_nsec - c0.tv_nsec));
      e->t_host[0][1]++;
    }
    w.cur_ctu = ctu; w.seen_ctu[0] = ctu;
  }
  const uint32_t slot = w.next_slot++ % kSlots;
  const int si = (int)(slot % kStreams);
  cudaStream_t stream = w.streams[si];
  if (w.seen_ctx[si] != tag) {                               // first launch of this stream for this frame: wait for its maps
    KVZ_CUDA_TRY(cudaStreamWaitEvent(stream, ctx->ready, 0));
    w.seen_ctx[si] = tag; w.seen_ctu[si] = -1;
  }
  if (ctx->n_partial > 0 && w.seen_ctu[si] != ctu) {
    KVZ_CUDA_TRY(cudaStreamWaitEvent(stream, w.ctu_ready, 0));
    w.seen_ctu[si] = ctu;
  }
  MeFrame f = ctx->f;
  f.lambda_sqrt_i = lambda_sqrt_i;
  const uint32_t gen = ++w.gen[slot];
  uint8_t *dslot = w.d + (size_t)slot * kSlotBytes;
  // the PUs and their jobs go in the launch itself; results and completion words come back through the mapped slot
  struct timespec l0, l1;
  if (e->timing) clock_gettime(CLOCK_MONOTONIC, &l0);
  if (!launch_me_search_pu(f, nullptr, nullptr, reinterpret_cast<kvz_cuda_me_result *>(dslot), n_pus, stream, pus, jobs,
                           reinterpret_cast<uint32_t *>(dslot + kSlotResults), gen)) return 0;
  if (e->timing) {
    clock_gettime(CLOCK_MONOTONIC, &l1);
    e->t_host[1][0] += (unsigned long long)((l1.tv_sec - l0.tv_sec) * 1000000000ll + (l1.tv_nsec - l0.tv_nsec));
    e->t_host[1][1]++;
  }
  w.n_pus[slot] = (uint8_t)n_pus;
  for (int p = 0; p < n_pus; p++) {
    w.n_jobs[slot][p] = pus[p].num_jobs;
    const int m = std::max(pus[p].w, pus[p].h);
    w.size_class[slot][p] = m > 32 ? 3 : m > 16 ? 2 : m > 8 ? 1 : 0;
  }
  *ticket = (int)(slot | (gen & 0x7FFFFFu) << 8);
  e->stats[1] += (unsigned long long)n_pus;
  e->stats[5] += n_jobs_total;
  e->meas[1] += (unsigned long long)n_pus * sizeof(kvz_cuda_me_pu) + n_jobs_total * sizeof(kvz_cuda_me_job);
  return 1;
}

int kvz_cuda_enc_search_end(kvz_cuda_enc *e, int ticket, int pu_index, kvz_cuda_me_result *results)
{
  if (!e || !results || ticket < 0) return fail_enc("kvz_cuda_enc_search_end: bad argument");
  Worker &w = t_worker;
  if (!w.h) return fail_enc("kvz_cuda_enc_search_end: this thread began no search");
  const uint32_t slot = (uint32_t)ticket & 0xFFu, gen = w.gen[slot % kSlots];
  if (slot >= (uint32_t)kSlots || ((uint32_t)ticket >> 8) != (gen & 0x7FFFFFu))
    return fail_enc("kvz_cuda_enc_search_end: the ticket has expired (its slot was used by a later launch of this thread)");
  if (pu_index < 0 || pu_index >= w.n_pus[slot]) return fail_enc("kvz_cuda_enc_search_end: bad pu_index");
  const uint8_t *hslot = w.h + (size_t)slot * kSlotBytes;
  const volatile uint32_t *flag = reinterpret_cast<const volatile uint32_t *>(hslot + kSlotResults) + pu_index;
  const WaitMode wm = wait_mode();
  struct timespec ts0;
  const bool timed = e->timing;
  if (timed) clock_gettime(CLOCK_MONOTONIC, &ts0);
  for (unsigned spins = 1; *flag != gen; spins++) {
    if (wm == kWaitYield) sched_yield(); else __builtin_ia32_pause();
    if ((spins & 0x3FFFF) == 0) {                          // now and then: did the launch die?
      const cudaError_t st = cudaStreamQuery(w.streams[slot % kStreams]);
      if (st == cudaSuccess) {
        if (*flag != gen) return fail_enc("kvz_cuda_enc_search_end: the search finished without reporting");
        break;
      }
      if (st != cudaErrorNotReady) return fail("kvz_cuda_enc_search_end", st, __FILE__, __LINE__);
    }
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  if (timed) {
    struct timespec ts1;
    clock_gettime(CLOCK_MONOTONIC, &ts1);
    const int sz = w.size_class[slot][pu_index];
    e->t_ns[sz][1] += (unsigned long long)((ts1.tv_sec - ts0.tv_sec) * 1000000000ll + (ts1.tv_nsec - ts0.tv_nsec));
    e->t_calls[sz][1]++;
  }
  const int nj = w.n_jobs[slot][pu_index];
  memcpy(results, hslot + (size_t)pu_index * KVZ_CUDA_MAX_REFS * sizeof(kvz_cuda_me_result), (size_t)nj * sizeof(kvz_cuda_me_result));
  e->meas[2] += (unsigned long long)nj * sizeof(kvz_cuda_me_result);
  return 1;
}

int kvz_cuda_enc_search_ready(kvz_cuda_enc *e, int ticket, int pu_index)
{
  Worker &w = t_worker;
  if (!e || ticket < 0 || !w.h) return -1;
  const uint32_t slot = (uint32_t)ticket & 0xFFu;
  if (slot >= (uint32_t)kSlots) return -1;
  const uint32_t gen = w.gen[slot];
  if (((uint32_t)ticket >> 8) != (gen & 0x7FFFFFu) || pu_index < 0 || pu_index >= w.n_pus[slot]) return -1;
  const volatile uint32_t *flag = reinterpret_cast<const volatile uint32_t *>(w.h + (size_t)slot * kSlotBytes + kSlotResults) + pu_index;
  return *flag == gen ? 1 : 0;
}

int kvz_cuda_enc_search_pu(kvz_cuda_enc *e, int handle, int lambda_sqrt_i, const kvz_cuda_me_pu *pu,
                           const kvz_cuda_me_job *jobs, kvz_cuda_me_result *results)
{
  if (!e || !pu || !jobs || !results) return fail_enc("kvz_cuda_enc_search_pu: null argument");
  struct timespec ts0;
  if (e->timing) clock_gettime(CLOCK_MONOTONIC, &ts0);
  kvz_cuda_me_job padded[KVZ_CUDA_MAX_REFS];
  const int nj = pu->num_jobs;
  if (nj < 1 || nj > KVZ_CUDA_MAX_REFS) return fail_enc("kvz_cuda_enc_search_pu: bad num_jobs");
  memcpy(padded, jobs, (size_t)nj * sizeof(kvz_cuda_me_job));
  int ticket = -1;
  if (!kvz_cuda_enc_search_begin(e, handle, lambda_sqrt_i, 1, pu, padded, &ticket)) return 0;
  if (!kvz_cuda_enc_search_end(e, ticket, 0, results)) return 0;
  if (e->timing) {
    struct timespec ts1;
    clock_gettime(CLOCK_MONOTONIC, &ts1);
    const int sz = std::max(pu->w, pu->h) > 32 ? 3 : std::max(pu->w, pu->h) > 16 ? 2 : std::max(pu->w, pu->h) > 8 ? 1 : 0;
    e->t_ns[sz][0] += (unsigned long long)((ts1.tv_sec - ts0.tv_sec) * 1000000000ll + (ts1.tv_nsec - ts0.tv_nsec));
    e->t_calls[sz][0]++;
  }
  return 1;
}

int kvz_cuda_enc_preload_source(kvz_cuda_enc *e, int64_t pts, const void *luma, int32_t stride)
{
  if (!e || !luma) return fail_enc("kvz_cuda_enc_preload_source: null argument");
  std::lock_guard<std::mutex> guard(e->lock);
  KVZ_CUDA_TRY(cudaSetDevice(e->cfg.device));
  Preload p;
  p.pts = pts; p.d = nullptr;
  if (!e->spare_planes.empty()) { p.d = e->spare_planes.back(); e->spare_planes.pop_back(); }
  else KVZ_CUDA_TRY(cudaMalloc(&p.d, (size_t)e->cfg.width * e->cfg.height));
  KVZ_CUDA_TRY(cudaMemcpy2DAsync(p.d, e->cfg.width, luma, stride, e->cfg.width, e->cfg.height, cudaMemcpyHostToDevice, e->stream));
  KVZ_CUDA_TRY(cudaStreamSynchronize(e->stream));
  e->preloads.push_back(p);
  return 1;
}

void kvz_cuda_enc_measurements(kvz_cuda_enc *e, unsigned long long m[8])
{
  if (e) {
    std::lock_guard<std::mutex> guard(e->lock);
    cudaSetDevice(e->cfg.device);
    for (FrameCtx *c : e->ctxs) harvest_dense(e, c, true);
  }
  for (int i = 0; i < 8; i++) m[i] = e ? e->meas[i].load() : 0ull;
}

void kvz_cuda_enc_stats(const kvz_cuda_enc *e, unsigned long long stats[8])
{
  for (int i = 0; i < 8; i++) stats[i] = e ? e->stats[i].load() : 0ull;
  if (e && e->timing) {
    unsigned long long ph[4][4];
    cudaSetDevice(e->cfg.device);
    cudaDeviceSynchronize();
    if (me_pu_phase_times(ph))
      for (int sz = 0; sz < 4; sz++)
        if (ph[sz][3])
          fprintf(stderr, "kvz_cuda_enc: PU %2d  device time per call: box scan %.1f us, CTA scan %.1f us, refinement %.1f us (%llu calls)\n",
                  8 << sz, ph[sz][0] / 1000.0 / ph[sz][3], ph[sz][1] / 1000.0 / ph[sz][3], ph[sz][2] / 1000.0 / ph[sz][3], ph[sz][3]);
  }
  if (e && e->timing) {
    static const char *names[4] = { "ctu_begin (blocks of pictures in flight + the CTU's dense launch)", "the PU launch call", "frame_begin", "-" };
    for (int i = 0; i < 3; i++)
      if (e->t_host[i][1].load())
        fprintf(stderr, "kvz_cuda_enc: host time in %s: %llu calls, %.1f us each\n", names[i], e->t_host[i][1].load(),
                e->t_host[i][0].load() / 1000.0 / e->t_host[i][1].load());
  }
  if (e && e->timing)
    for (int sz = 0; sz < 4; sz++)
      for (int k = 0; k < 2; k++)
        if (e->t_calls[sz][k].load())
          fprintf(stderr, "kvz_cuda_enc: PU %2d  %-22s %8llu calls  %8.1f us/call\n", 8 << sz,
                  k ? "waited in search_end          " : "blocking search_pu (begin + end)", e->t_calls[sz][k].load(),
                  e->t_ns[sz][k].load() / 1000.0 / e->t_calls[sz][k].load());
}

}  // extern "C"
