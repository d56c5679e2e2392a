/*
 * kvazaar_b200/host/kvz_glue_search_inter.c -- Kvazaar's inter PU search with the reference loop on the device.
 *
 * Compiled INSTEAD of src/search_inter.c: it includes that file in place (from the reference tree, at build time)
 * with its two entry points renamed, so every file-static helper of the reference (search_pu_inter_bipred,
 * fracmv_within_tile, ...) is reachable here and the reference's own search stays available under the names
 * kvz_search_cu_inter_host / kvz_search_cu_smp_host for KVZ_CUDA_MODE=off|percall.
 *
 * What this file adds is the replacement of search_pu_inter (src/search_inter.c:1451-1524): the merge and AMVP
 * candidates are still derived by the host's own kvz_inter_get_merge_cand / kvz_inter_get_mv_cand from the working
 * lcu_t -- that is what keeps the decisions in CTU wavefront order -- and the loop over the reference pictures
 * (search_pu_inter_ref, :1134-1297: integer search, fractional refinement, merge match, AMVP choice, *inter_cost
 * carry) is ONE call of kvz_cuda_enc_search_pu.
 */
#define kvz_search_cu_inter kvz_search_cu_inter_host
#define kvz_search_cu_smp kvz_search_cu_smp_host
#include "search_inter.c"
#undef kvz_search_cu_inter
#undef kvz_search_cu_smp

#include <pthread.h>
#include <time.h>
#include <stdio.h>
#include <stdlib.h>

#include "kvz_cuda_enc.h"
#include "kvz_glue.h"

/* ---- the session: one per process ("Only one encoder may be open at a time", src/kvazaar.h:619) ---- */
static kvz_cuda_enc *g_enc = NULL;
static uint64_t g_epoch = 0;                 /* bumped when a session opens: launches made under an older one never match */
static unsigned long long g_part_ns[6];          /* KVZ_CUDA_STATS: merge list, AMVP, colocated vectors, launch call, record derivations, calls of glue_build_inputs */
static unsigned long long g_glue_ns[4];          /* KVZ_CUDA_STATS: ns in deriving the real records, waiting for results, launching ahead; calls */
static unsigned long long g_origin8[3][2];         /* KVZ_CUDA_STATS, 8x8 PUs answered ahead by origin of the launch: count, ns waited */
static unsigned long long g_spec_by_size[4][2];   /* by PU size 8, 16, 32, 64: answered ahead, searched on the spot */
static unsigned long long g_spec_cnt[4];     /* PU searches answered by a launch made ahead, searched on the spot, PUs launched ahead, PUs launched */
static const encoder_control_t *g_enc_ctrl = NULL;
static pthread_mutex_t g_enc_lock = PTHREAD_MUTEX_INITIALIZER;

static void glue_close_session(void)
{
  if (g_enc) {
    if (getenv("KVZ_CUDA_STATS")) {
      unsigned long long s[8];
      kvz_cuda_enc_stats(g_enc, s);
      fprintf(stderr, "kvz_cuda_enc: frames %llu, PU searches %llu, reference jobs %llu, reference planes uploaded %llu, "
              "SAD maps of the last frame %.1f MiB (%llu displacements), reference jobs with a window outside the box %llu\n",
              s[0], s[1], s[5], s[2], s[3] / 1048576.0, s[4], s[6]);
      fprintf(stderr, "kvz_glue: PU searches the host asked for: %llu answered by a launch made ahead, %llu searched on the spot; "
              "PUs launched %llu, of them ahead of time %llu\n", g_spec_cnt[0], g_spec_cnt[1], g_spec_cnt[3], g_spec_cnt[2]);
      if (g_glue_ns[3])
        fprintf(stderr, "kvz_glue: per search_pu_inter: records %.1f us, lookup + wait for the results %.1f us, bi-prediction + launching ahead %.1f us\n",
                g_glue_ns[0] / 1e3 / g_glue_ns[3], g_glue_ns[1] / 1e3 / g_glue_ns[3], g_glue_ns[2] / 1e3 / g_glue_ns[3]);
      if (g_part_ns[5])
        fprintf(stderr, "kvz_glue: per derivation of a PU's records (%llu): merge list %.2f us, AMVP %.2f us, colocated vectors %.2f us; per launch (%llu): %.2f us\n",
                g_part_ns[5], g_part_ns[0] / 1e3 / g_part_ns[5], g_part_ns[1] / 1e3 / g_part_ns[5], g_part_ns[2] / 1e3 / g_part_ns[5],
                g_part_ns[4], g_part_ns[4] ? g_part_ns[3] / 1e3 / g_part_ns[4] : 0.0);
      {
        static const char *on[3] = { "rode along with a search on the spot", "next CU of the walk and its chain", "child under the inherit hypothesis" };
        for (int i = 0; i < 3; i++)
          if (g_origin8[i][0])
            fprintf(stderr, "kvz_glue: 8x8 searches answered by a launch that %s: %llu, waited %.1f us each\n", on[i], g_origin8[i][0],
                    g_origin8[i][1] / 1e3 / g_origin8[i][0]);
      }
      for (int i = 0; i < 4; i++)
        fprintf(stderr, "kvz_glue: PU %2d: %llu answered ahead, %llu searched on the spot\n", 8 << i, g_spec_by_size[i][0], g_spec_by_size[i][1]);
    }
    kvz_cuda_enc_close(g_enc);
    g_enc = NULL;
  }
}

void kvz_glue_session_close(void)
{
  pthread_mutex_lock(&g_enc_lock);
  glue_close_session();
  g_enc_ctrl = NULL;
  pthread_mutex_unlock(&g_enc_lock);
}

static int glue_ime_search(const kvz_config *cfg, int *range)
{
  *range = 32;
  switch (cfg->ime_algorithm) {
    case KVZ_IME_FULL64: *range = 64; return KVZ_CUDA_IME_FULL;
    case KVZ_IME_FULL32: case KVZ_IME_FULL: return KVZ_CUDA_IME_FULL;     /* search_inter.c:1208-1215 */
    case KVZ_IME_FULL16: *range = 16; return KVZ_CUDA_IME_FULL;
    case KVZ_IME_FULL8: *range = 8; return KVZ_CUDA_IME_FULL;
    case KVZ_IME_TZ: return KVZ_CUDA_IME_TZ;
    case KVZ_IME_DIA: return KVZ_CUDA_IME_DIA;
    default: return KVZ_CUDA_IME_HEXBS;
  }
}

/* per (frame, tile): a handle of the session, found again by every worker thread of the frame */
#define GLUE_FRAME_SLOTS 64
static struct { uint64_t id; int handle; } g_frames[GLUE_FRAME_SLOTS];
static int g_frames_next = 0;
static pthread_mutex_t g_frames_lock = PTHREAD_MUTEX_INITIALIZER;

static kvz_cuda_enc *glue_session(const encoder_state_t *state)
{
  const encoder_control_t *ctrl = state->encoder_control;
  if (g_enc && g_enc_ctrl == ctrl) return g_enc;       /* the session lives as long as its encoder (kvz_glue_api.c) */
  pthread_mutex_lock(&g_enc_lock);
  if (!g_enc || g_enc_ctrl != ctrl) {
    static int registered = 0;
    glue_close_session();
    pthread_mutex_lock(&g_frames_lock);
    memset(g_frames, 0, sizeof(g_frames));
    pthread_mutex_unlock(&g_frames_lock);
    const kvz_config *cfg = &ctrl->cfg;
    /* what the device search does not reproduce is refused, loudly -- never searched on the CPU instead */
    if (cfg->owf != 0 && (cfg->tiles_width_count > 1 || cfg->tiles_height_count > 1))
      kvz_glue_die("KVZ_CUDA_MODE=batch: --owf > 0 together with --tiles is not supported (pictures in flight are brought "
                   "to the device by the CTU dependencies of one wavefront, encoderstate.c:786-801)");
    if (cfg->owf != 0 && !cfg->wpp)
      kvz_glue_die("KVZ_CUDA_MODE=batch: --owf > 0 needs --wpp (without it the motion vectors into a picture in flight are not restricted)");
    if (cfg->mv_rdo) kvz_glue_die("--mv-rdo (kvz_calc_mvd_cost_cabac, rdo.c:908) is not on the device");
    if (cfg->smp_enable || cfg->amp_enable) kvz_glue_die("--smp / --amp PU shapes (4 and 12 sample sides) are not searched on the device");
    if (ctrl->bitdepth != 8) kvz_glue_die("the encoder-session face is 8-bit only");
    kvz_cuda_enc_cfg c;
    memset(&c, 0, sizeof(c));
    c.device = kvz_glue_device();
    c.width = ctrl->in.width; c.height = ctrl->in.height; c.bitdepth = ctrl->bitdepth;
    c.ime_search = glue_ime_search(cfg, &c.range);
    c.me_early_termination = cfg->me_early_termination; c.me_max_steps = cfg->me_max_steps;
    c.fme_level = cfg->fme_level;
    c.owf = cfg->owf; c.wpp = cfg->wpp; c.sao_type = cfg->sao_type; c.deblock_enable = cfg->deblock_enable;
    c.mv_constraint = cfg->mv_constraint;
    c.max_ref_lcu_right = ctrl->max_inter_ref_lcu.right; c.max_ref_lcu_down = ctrl->max_inter_ref_lcu.down;
    if (getenv("KVZ_CUDA_MAP_MBYTES")) c.max_map_mbytes = atoi(getenv("KVZ_CUDA_MAP_MBYTES"));
    /* every worker thread launches on four streams of its own: give them hardware queues (read when CUDA starts) */
    setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
    g_enc = kvz_cuda_enc_open(&c);
    if (!g_enc) kvz_glue_die("kvz_cuda_enc_open");
    __atomic_fetch_add(&g_epoch, 1, __ATOMIC_SEQ_CST);
    g_enc_ctrl = ctrl;
    if (!registered) { atexit(glue_close_session); registered = 1; }
  }
  pthread_mutex_unlock(&g_enc_lock);
  return g_enc;
}

/* ---- for drivers of the library (bench.py): measurement counters of the session and source preloading ---- */
KVZ_PUBLIC int kvz_glue_enc_measurements(unsigned long long m[8])
{
  if (!g_enc) return 0;
  kvz_cuda_enc_measurements(g_enc, m);
  return 1;
}

KVZ_PUBLIC int kvz_glue_enc_stats(unsigned long long s[8])
{
  if (!g_enc) return 0;
  kvz_cuda_enc_stats(g_enc, s);
  return 1;
}

/* c[0] searches answered by a launch made ahead, c[1] searched on the spot, c[2] PUs launched ahead, c[3] PUs launched */
KVZ_PUBLIC void kvz_glue_spec_counts(unsigned long long c[4])
{
  for (int i = 0; i < 4; i++) c[i] = __atomic_load_n(&g_spec_cnt[i], __ATOMIC_RELAXED);
}

/* the source luma of the picture the host will feed with presentation number `pts`, put on the device ahead of time */
KVZ_PUBLIC int kvz_glue_preload_source(int64_t pts, const uint8_t *luma, int stride)
{
  return g_enc ? kvz_cuda_enc_preload_source(g_enc, pts, luma, stride) : 0;
}

/* quarter-pel vectors of the motion fields of the DPB, one sample per 16x16: where extra_mv (search_inter.c:1188-1206)
 * and the merge candidates of this frame will probably point */
static int glue_mv_hints(const encoder_state_t *state, int16_t *hints, int max_hints)
{
  const image_list_t *ref = state->frame->ref;
  int n = 0;
  for (unsigned r = 0; r < ref->used_size; r++) {
    const cu_array_t *cua = ref->cu_arrays[r];
    if (!cua) continue;
    for (unsigned y = 8; y < cua->height; y += 16) {
      for (unsigned x = 8; x < cua->width; x += 16) {
        const cu_info_t *cu = kvz_cu_array_at_const(cua, x, y);
        if (cu->type != CU_INTER) continue;
        for (int l = 0; l < 2; l++) {
          if (!(cu->inter.mv_dir & (1 << l)) || n >= max_hints) continue;
          hints[2 * n] = cu->inter.mv[l][0]; hints[2 * n + 1] = cu->inter.mv[l][1];
          n++;
        }
      }
    }
  }
  return n;
}

static int glue_frame_handle(kvz_cuda_enc *enc, const encoder_state_t *state)
{
  const uint64_t id = ((uint64_t)(uint32_t)state->frame->num << 16) | (uint64_t)(state->tile->id & 0xFFFF);
  for (int i = 0; i < GLUE_FRAME_SLOTS; i++)            /* racy read of a slot is fine: ids are unique and written last */
    if (__atomic_load_n(&g_frames[i].id, __ATOMIC_ACQUIRE) == id + 1) return g_frames[i].handle;
  pthread_mutex_lock(&g_frames_lock);
  int handle = -1;
  for (int i = 0; i < GLUE_FRAME_SLOTS; i++) if (g_frames[i].id == id + 1) handle = g_frames[i].handle;
  if (handle < 0) {
    const videoframe_t *vf = state->tile->frame;
    const kvz_picture *src = vf->source->base_image ? vf->source->base_image : vf->source;
    const image_list_t *ref = state->frame->ref;
    const int max_hints = 1 << 16;
    int16_t *hints = malloc(sizeof(int16_t) * 2 * max_hints);
    const int n_hints = hints ? glue_mv_hints(state, hints, max_hints) : 0;
    kvz_cuda_enc_frame f;
    memset(&f, 0, sizeof(f));
    f.frame_id = id;
    f.src_luma = src->y; f.src_stride = src->stride; f.src_key = (uint64_t)(uintptr_t)src;
    f.src_pts = src->pts;
    f.tile_x = state->tile->offset_x; f.tile_y = state->tile->offset_y;
    f.tile_w = vf->width; f.tile_h = vf->height;
    memcpy(f.ref_LX, state->frame->ref_LX, sizeof(f.ref_LX));
    f.n_refs = (int)ref->used_size;
    for (int r = 0; r < f.n_refs; r++) {
      const kvz_picture *im = ref->images[r];
      f.refs[r].luma = im->y; f.refs[r].stride = im->stride; f.refs[r].poc = ref->pocs[r];
      f.refs[r].key = (uint64_t)(uintptr_t)im;
      f.refs[r].mv_hints = hints; f.refs[r].n_mv_hints = n_hints;
      /* still being reconstructed?  the owf previous encoder states are the frames that can be (kvazaar.c:96-126) */
      const encoder_state_t *ps = state;
      for (int k = 0; k < state->encoder_control->cfg.owf; k++) {
        ps = ps->previous_encoder_state;
        if (ps == state) break;
        /* a tile-level state holds a sub-image of the picture (encoderstate.c:869-903): compare the owners of the samples */
        const kvz_picture *prec = ps->tile->frame->rec;
        if (prec && (prec == im || prec->base_image == im->base_image)) f.refs[r].in_flight = 1;
      }
    }
    handle = kvz_cuda_enc_frame_begin(enc, &f);
    free(hints);
    if (handle < 0) kvz_glue_die("kvz_cuda_enc_frame_begin");
    KVZ_GLUE_COUNT(KVZ_GLUE_CNT_ME_FRAMES);
    const int slot = g_frames_next;
    g_frames_next = (g_frames_next + 1) % GLUE_FRAME_SLOTS;
    g_frames[slot].handle = handle;
    __atomic_store_n(&g_frames[slot].id, id + 1, __ATOMIC_RELEASE);
  }
  pthread_mutex_unlock(&g_frames_lock);
  return handle;
}

/* which list holds picture ref_idx, and where (search_inter.c:1142-1162) */
static void glue_list_of(const encoder_state_t *state, int ref_idx, int8_t *ref_list, int8_t *lx_idx)
{
  const int n = MAX(state->frame->ref_LX_size[0], state->frame->ref_LX_size[1]);
  *ref_list = -1;
  for (*lx_idx = 0; *lx_idx < n; (*lx_idx)++) {
    if (*lx_idx < state->frame->ref_LX_size[0] && state->frame->ref_LX[0][*lx_idx] == ref_idx) { *ref_list = 0; return; }
    if (*lx_idx < state->frame->ref_LX_size[1] && state->frame->ref_LX[1][*lx_idx] == ref_idx) { *ref_list = 1; return; }
  }
  kvz_glue_die("reference picture in neither list");
}

void kvz_glue_inter_get_mv_cands(const encoder_state_t *const state, int32_t x, int32_t y, int32_t width, int32_t height,
                                 lcu_t *lcu, int n, const int8_t *reflist, const int8_t *lx_idx, int16_t (*mv_cand)[2][2]);

/* ---- the inputs of one device search, and searches launched ahead of time ----
 *
 * The host walks the CU quadtree in z-order and every kvz_search_cu_inter blocks for the device (~50 us).  Most of
 * that wait is hidden by launching a PU BEFORE the walk reaches it, with the candidates the PU will get if the
 * decisions in between turn out as expected: after the search of CU c the next CU of the walk (its next sibling, or
 * an ancestor's) is launched under the hypothesis "c stays as just searched", together with the chain of top-left
 * children at the same origin (their neighbours all lie outside the CU, so their candidates are known already).
 * When the walk arrives, the REAL candidates are derived from the real lcu_t as always and compared, record by
 * record, with what was sent: equal -> the results are taken (same inputs, same frame state, a deterministic search);
 * different -> the launch is ignored and the PU is searched now.  Decisions are therefore exactly the reference's;
 * a wrong guess costs device time only.  KVZ_CUDA_SPECULATE=0 switches it off. */
typedef struct {
  kvz_cuda_me_pu pu;
  kvz_cuda_me_job jobs[KVZ_CUDA_MAX_REFS];
} glue_inputs;

#define GLUE_SPEC_SLOTS 64          /* = the ticket slots of a thread (kvz_cuda_enc.h) */
#define GLUE_SPEC_PUS 8           /* = the PUs one kvz_cuda_enc_search_begin takes */
typedef struct {
  int valid, ticket, idx, handle, lambda;
  int origin;                       /* statistics: 0 rode along with a search made on the spot, 1 next CU of the walk / its chain, 2 child under the inherit hypothesis */
  uint64_t epoch, frame_id;
  glue_inputs in;
} glue_spec;
static __thread glue_spec t_spec[GLUE_SPEC_SLOTS][GLUE_SPEC_PUS];
/* what a lookup scans: 8 bytes per entry (the records themselves are 368 bytes apart); 0 = no entry */
static __thread uint64_t t_key[GLUE_SPEC_SLOTS][GLUE_SPEC_PUS];
static uint64_t glue_key(const glue_inputs *in)
{
  return 1ull << 63 | (uint64_t)(uint16_t)in->pu.x << 32 | (uint64_t)(uint16_t)in->pu.y << 16 | (uint64_t)in->pu.w << 8 | in->pu.h;
}
static __thread lcu_t t_hyp_storage;                          /* the hypothetical CU array (only ->cu is used); freed with its thread */

/* KVZ_GLUE_TRACE=1: the searches and launches of the first CTU of frame 1 on stderr (debugging the launching ahead) */
static int kvz_glue_trace(void)
{
  static int on = -1;
  if (on < 0) on = getenv("KVZ_GLUE_TRACE") != NULL;
  return on;
}

static int glue_timing(void)
{
  static int on = -1;
  if (on < 0) on = getenv("KVZ_CUDA_STATS") != NULL;
  return on;
}
static unsigned long long glue_now(void)
{
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (unsigned long long)ts.tv_sec * 1000000000ull + (unsigned long long)ts.tv_nsec;
}

static int glue_speculate(void)
{
  static int on = -1;
  if (on < 0) on = !(getenv("KVZ_CUDA_SPECULATE") && atoi(getenv("KVZ_CUDA_SPECULATE")) == 0);
  return on;
}

/* the PU record and the per-picture job records search_pu_inter / search_pu_inter_ref derive for a PU
 * (search_inter.c:1469-1481 merge candidates, :1164-1181 AMVP candidates, :1188-1206 colocated vector), from the CU
 * array of `lcu` -- the real one, or a hypothetical copy.  Reads only lcu->cu; writes nothing but *in, *info. */
static void glue_build_inputs(encoder_state_t *state, int x, int y, int width, int height, bool merge_a1, bool merge_b1,
                              lcu_t *lcu, inter_search_info_t *info, glue_inputs *in, int8_t *ref_list, int8_t *lx_idx)
{
  const int n_refs = (int)state->frame->ref->used_size;
  const int timing = glue_timing();
  const unsigned long long p0 = timing ? glue_now() : 0;
  memset(in, 0, sizeof(*in));
  info->num_merge_cand = kvz_inter_get_merge_cand(state, x, y, width, height, merge_a1, merge_b1, info->merge_cand, lcu);
  const unsigned long long p1 = timing ? glue_now() : 0;
  kvz_cuda_me_pu *pu = &in->pu;
  pu->x = (int16_t)x; pu->y = (int16_t)y; pu->w = (uint8_t)width; pu->h = (uint8_t)height;
  pu->num_merge_cand = (uint8_t)info->num_merge_cand;
  pu->num_jobs = (uint8_t)n_refs;
  for (int i = 0; i < info->num_merge_cand; i++) {
    /* only the lists `dir` names are set by the derivation (inter.c:1375-1530); the others stay zero here, so that
     * two derivations from the same CU array give the same bytes whatever their callers' stacks held */
    const inter_merge_cand_t *m = &info->merge_cand[i];
    pu->merge_cand[i].dir = m->dir;
    for (int l = 0; l < 2; l++)
      if (m->dir & (1 << l)) {
        pu->merge_cand[i].ref[l] = m->ref[l];
        pu->merge_cand[i].mv[l][0] = m->mv[l][0]; pu->merge_cand[i].mv[l][1] = m->mv[l][1];
      }
  }
  /* the AMVP candidates of every picture from ONE derivation of the PU's neighbours (kvz_glue_inter.c) */
  int16_t cands[KVZ_CUDA_MAX_REFS][2][2];
  for (int r = 0; r < n_refs; r++) glue_list_of(state, r, &ref_list[r], &lx_idx[r]);
  kvz_glue_inter_get_mv_cands(state, x, y, width, height, lcu, n_refs, ref_list, lx_idx, cands);
  if (n_refs > 0) memcpy(info->mv_cand, cands[n_refs - 1], sizeof(info->mv_cand));      /* as the reference's loop leaves it */
  const unsigned long long p2 = timing ? glue_now() : 0;
  for (int r = 0; r < n_refs; r++) {
    memcpy(in->jobs[r].mv_cand, cands[r], sizeof(in->jobs[r].mv_cand));
    /* the colocated CU's vector in that picture (:1188-1206) */
    const int mid_x = state->tile->offset_x + x + (width >> 1), mid_y = state->tile->offset_y + y + (height >> 1);
    const cu_info_t *ref_cu = kvz_cu_array_at_const(state->frame->ref->cu_arrays[r], mid_x, mid_y);
    if (ref_cu->type == CU_INTER) {
      const int l = (ref_cu->inter.mv_dir & 1) ? 0 : 1;
      in->jobs[r].extra_mv[0] = ref_cu->inter.mv[l][0]; in->jobs[r].extra_mv[1] = ref_cu->inter.mv[l][1];
    }
    in->jobs[r].ref_idx = (uint8_t)r;
  }
  if (timing) {
    const unsigned long long p3 = glue_now();
    __atomic_fetch_add(&g_part_ns[0], p1 - p0, __ATOMIC_RELAXED);
    __atomic_fetch_add(&g_part_ns[1], p2 - p1, __ATOMIC_RELAXED);
    __atomic_fetch_add(&g_part_ns[2], p3 - p2, __ATOMIC_RELAXED);
    __atomic_fetch_add(&g_part_ns[5], 1ull, __ATOMIC_RELAXED);
  }
}

static bool glue_same_inputs(const glue_inputs *a, const glue_inputs *b)
{
  return memcmp(&a->pu, &b->pu, sizeof(a->pu)) == 0 &&
         memcmp(a->jobs, b->jobs, (size_t)a->pu.num_jobs * sizeof(kvz_cuda_me_job)) == 0;
}

static glue_spec *glue_spec_find(uint64_t frame_id, int handle, int lambda, const glue_inputs *in)
{
  const uint64_t key = glue_key(in);
  for (int s = 0; s < GLUE_SPEC_SLOTS; s++)
    for (int k = 0; k < GLUE_SPEC_PUS; k++) {
      if (t_key[s][k] != key) continue;
      glue_spec *e = &t_spec[s][k];
      if (e->valid && e->epoch == g_epoch && e->frame_id == frame_id && e->handle == handle && e->lambda == lambda &&
          glue_same_inputs(&e->in, in)) return e;
    }
  return NULL;
}

/* search_cu's test for calling kvz_search_cu_inter at `depth` (search.c:446-468) */
static bool glue_cu_is_searched(const encoder_state_t *state, int x, int y, int depth)
{
  const encoder_control_t *ctrl = state->encoder_control;
  const videoframe_t *frame = state->tile->frame;
  const int cu_width = LCU_WIDTH >> depth;
  if (x + cu_width > frame->width || y + cu_width > frame->height) return false;
  if (state->frame->slicetype == KVZ_SLICE_I || depth > MAX_DEPTH) return false;
  const int cu_width_inter_min = LCU_WIDTH >> ctrl->cfg.pu_depth_inter.max;
  return WITHIN(depth, ctrl->cfg.pu_depth_inter.min, ctrl->cfg.pu_depth_inter.max) ||
         (x & ~(cu_width_inter_min - 1)) + cu_width_inter_min > frame->width ||
         (y & ~(cu_width_inter_min - 1)) + cu_width_inter_min > frame->height;
}

/* the PUs of one launch being put together */
typedef struct {
  glue_inputs sent[GLUE_SPEC_PUS];
  uint8_t origin[GLUE_SPEC_PUS];
  int n, n_needed_now, cur_origin;
} glue_batch;

/* adds CU (x, y, depth) as search_cu would search it, with the records `lcu` gives it -- unless search_cu would not
 * search it there, or the same records are in flight already */
static void glue_batch_add_cu(encoder_state_t *state, glue_batch *b, uint64_t frame_id, int handle, int lambda,
                              int x, int y, int depth, lcu_t *lcu)
{
  if (b->n >= GLUE_SPEC_PUS || !glue_cu_is_searched(state, x, y, depth)) return;
  inter_search_info_t info;
  int8_t ref_list[KVZ_CUDA_MAX_REFS], lx_idx[KVZ_CUDA_MAX_REFS];
  const int w = LCU_WIDTH >> depth;
  glue_inputs *in = &b->sent[b->n];
  glue_build_inputs(state, x, y, w, w, true, true, lcu, &info, in, ref_list, lx_idx);
  if (in->pu.num_jobs == 0 || glue_spec_find(frame_id, handle, lambda, in)) return;
  b->origin[b->n] = (uint8_t)b->cur_origin;
  b->n++;
}

/* the top-left chain (x, y, depth), (x, y, depth + 1), ...: search_cu reaches each of them before any other CU of the
 * subtree, so their neighbours are the ones `lcu` has now */
static void glue_batch_add_chain(encoder_state_t *state, glue_batch *b, uint64_t frame_id, int handle, int lambda,
                                 int x, int y, int depth, lcu_t *lcu)
{
  for (int d = depth; d <= MAX_DEPTH; d++) glue_batch_add_cu(state, b, frame_id, handle, lambda, x, y, d, lcu);
}

/* one launch for the batch; returns its ticket, or -1 when the batch is empty */
static int glue_batch_launch(kvz_cuda_enc *enc, uint64_t frame_id, int handle, int lambda, const glue_batch *b)
{
  if (b->n == 0) return -1;
  kvz_cuda_me_pu pus[GLUE_SPEC_PUS];
  kvz_cuda_me_job jobs[GLUE_SPEC_PUS * KVZ_CUDA_MAX_REFS];
  for (int i = 0; i < b->n; i++) {
    pus[i] = b->sent[i].pu;
    memcpy(jobs + i * KVZ_CUDA_MAX_REFS, b->sent[i].jobs, sizeof(b->sent[i].jobs));
  }
  if (kvz_glue_trace() && (frame_id >> 16) == 1 && pus[0].x < 64 && pus[0].y < 64) {
    fprintf(stderr, "T launch");
    for (int i = 0; i < b->n; i++) fprintf(stderr, " (%d,%d,%d)", pus[i].x, pus[i].y, pus[i].w);
    fprintf(stderr, "\n");
  }
  int ticket = -1;
  const unsigned long long l0 = glue_timing() ? glue_now() : 0;
  if (!kvz_cuda_enc_search_begin(enc, handle, lambda, b->n, pus, jobs, &ticket)) kvz_glue_die("kvz_cuda_enc_search_begin");
  if (l0) { __atomic_fetch_add(&g_part_ns[3], glue_now() - l0, __ATOMIC_RELAXED); __atomic_fetch_add(&g_part_ns[4], 1ull, __ATOMIC_RELAXED); }
  if ((ticket & 0xFF) >= GLUE_SPEC_SLOTS) kvz_glue_die("kvz_cuda_enc_search_begin: ticket slot beyond the glue's table");
  glue_spec *row = t_spec[ticket & 0xFF];
  for (int i = 0; i < GLUE_SPEC_PUS; i++) { row[i].valid = 0; t_key[ticket & 0xFF][i] = 0; }
  for (int i = 0; i < b->n; i++) {
    glue_spec *e = &row[i];
    t_key[ticket & 0xFF][i] = glue_key(&b->sent[i]);
    e->valid = 1; e->ticket = ticket; e->idx = i; e->handle = handle; e->lambda = lambda;
    e->epoch = g_epoch; e->frame_id = frame_id; e->in = b->sent[i]; e->origin = b->origin[i];
  }
  __atomic_fetch_add(&g_spec_cnt[3], (unsigned long long)b->n, __ATOMIC_RELAXED);
  __atomic_fetch_add(&g_spec_cnt[2], (unsigned long long)(b->n - b->n_needed_now), __ATOMIC_RELAXED);
  return ticket;
}

/* the CU search_cu visits once the subtree of CU (x, y, depth) is finished (search.c:680-683: z-order; a CU that
 * lies outside the frame returns at once, :437-440) -- inside the same CTU, else 0 */
static int glue_next_cu(const videoframe_t *frame, int *x, int *y, int *depth)
{
  int cx = *x, cy = *y, d = *depth;
  while (d > 0) {
    const int s = LCU_WIDTH >> d;
    int i = ((cy / s) & 1) * 2 + ((cx / s) & 1);
    const int px = cx & ~(2 * s - 1), py = cy & ~(2 * s - 1);
    if (i == 3) { cx = px; cy = py; d--; continue; }
    i++;
    cx = px + (i & 1) * s; cy = py + (i >> 1) * s;
    if (cx >= frame->width || cy >= frame->height) continue;
    *x = cx; *y = cy; *depth = d;
    return 1;
  }
  return 0;
}

/* what search_pu_inter_ref writes into the CU when a picture lowers *inter_cost (:1281-1296), pictures in order;
 * true when some picture did */
static bool glue_apply_results(cu_info_t *cu, const kvz_cuda_me_result *res, int n_refs, const int8_t *ref_list,
                               const int8_t *lx_idx, double *inter_cost, uint32_t *inter_bitcost)
{
  bool any = false;
  for (int r = 0; r < n_refs; r++) {
    if (!res[r].took) continue;
    const int l = ref_list[r];
    cu->inter.mv_dir = l + 1;
    cu->merged = res[r].merged;
    cu->merge_idx = res[r].merge_idx;
    cu->inter.mv_ref[l] = lx_idx[r];
    cu->inter.mv[l][0] = res[r].mv[0];
    cu->inter.mv[l][1] = res[r].mv[1];
    CU_SET_MV_CAND(cu, l, res[r].mv_cand_idx);
    *inter_cost = res[r].cost;
    *inter_bitcost = res[r].bitcost + cu->inter.mv_dir - 1 + lx_idx[r];
    any = true;
  }
  return any;
}

/* After the search of CU (x, y, depth), under the hypothesis that the CU stays as `cur_cu` says now: launches the next
 * CU of the walk with its top-left chain, and -- when the CU can be split -- its children 1..3 as they will be searched
 * if the children before them come out like the CU itself (smooth motion; when they do not, the child is launched
 * again after its predecessor, by this same function). */
static void glue_launch_next(encoder_state_t *state, kvz_cuda_enc *enc, uint64_t frame_id, int handle, int lambda,
                             int x, int y, int depth, const lcu_t *lcu, const cu_info_t *cur_cu, bool found_inter)
{
  int nx = x, ny = y, nd = depth;
  const bool has_next = glue_next_cu(state->tile->frame, &nx, &ny, &nd) != 0;
  static int children = -1;                                   /* KVZ_CUDA_SPEC_CHILDREN=0: experiments */
  if (children < 0) children = !(getenv("KVZ_CUDA_SPEC_CHILDREN") && atoi(getenv("KVZ_CUDA_SPEC_CHILDREN")) == 0);
  const bool has_children = children && depth < MAX_DEPTH && depth < state->encoder_control->cfg.pu_depth_inter.max;
  if (!has_next && !has_children) return;
  lcu_t *const t_hyp = &t_hyp_storage;
  memcpy(t_hyp->cu, lcu->cu, sizeof(lcu->cu));
  cu_info_t filled = *cur_cu;
  filled.type = found_inter ? CU_INTER : CU_INTRA;
  const int w = LCU_WIDTH >> depth, xl = SUB_SCU(x), yl = SUB_SCU(y);
  for (int yy = yl; yy < yl + w; yy += 4)
    for (int xx = xl; xx < xl + w; xx += 4) *LCU_GET_CU_AT_PX(t_hyp, xx, yy) = filled;
  glue_batch b;
  b.n = b.n_needed_now = 0;
  b.cur_origin = 1;
  if (has_next) glue_batch_add_chain(state, &b, frame_id, handle, lambda, nx, ny, nd, t_hyp);
  b.cur_origin = 2;
  if (has_children && found_inter) {
    const int h = w >> 1;
    glue_batch_add_cu(state, &b, frame_id, handle, lambda, x + h, y, depth + 1, t_hyp);
    glue_batch_add_cu(state, &b, frame_id, handle, lambda, x, y + h, depth + 1, t_hyp);
    glue_batch_add_cu(state, &b, frame_id, handle, lambda, x + h, y + h, depth + 1, t_hyp);
  }
  glue_batch_launch(enc, frame_id, handle, lambda, &b);
}

/* search_pu_inter (search_inter.c:1451-1524) with the reference loop on the device */
static void glue_search_pu_inter(encoder_state_t *const state, int x_cu, int y_cu, int depth, part_mode_t part_mode,
                                 int i_pu, lcu_t *lcu, double *inter_cost, uint32_t *inter_bitcost)
{
  *inter_cost = MAX_INT;
  *inter_bitcost = MAX_INT;
  const kvz_config *cfg = &state->encoder_control->cfg;
  const videoframe_t *const frame = state->tile->frame;
  const int width_cu = LCU_WIDTH >> depth;
  const int x = PU_GET_X(part_mode, width_cu, x_cu, i_pu), y = PU_GET_Y(part_mode, width_cu, y_cu, i_pu);
  const int width = PU_GET_W(part_mode, width_cu, i_pu), height = PU_GET_H(part_mode, width_cu, i_pu);
  const bool merge_a1 = i_pu == 0 || width >= height;       /* second PUs of Nx2N.. / 2NxN.. (:1472-1477) */
  const bool merge_b1 = i_pu == 0 || width <= height;
  cu_info_t *cur_cu = LCU_GET_CU_AT_PX(lcu, SUB_SCU(x), SUB_SCU(y));

  inter_search_info_t info = {
    .state = state, .pic = frame->source, .origin = { x, y }, .width = width, .height = height,
    .mvd_cost_func = calc_mvd_cost,
  };
  glue_inputs in;
  int8_t ref_list[KVZ_CUDA_MAX_REFS], lx_idx[KVZ_CUDA_MAX_REFS];
  const int timing = glue_timing();
  unsigned long long t0 = timing ? glue_now() : 0, t1 = 0, t2 = 0;
  glue_build_inputs(state, x, y, width, height, merge_a1, merge_b1, lcu, &info, &in, ref_list, lx_idx);
  CU_SET_MV_CAND(cur_cu, 0, 0);
  CU_SET_MV_CAND(cur_cu, 1, 0);

  kvz_cuda_enc *enc = glue_session(state);
  const int handle = glue_frame_handle(enc, state);
  const uint64_t frame_id = ((uint64_t)state->frame->num << 16) | (uint64_t)(state->tile->id & 0xFFFF);
  const int lambda = (int)(state->lambda_sqrt + 0.5);
  const int n_refs = (int)state->frame->ref->used_size;
  const bool square = part_mode == SIZE_2Nx2N;
  kvz_cuda_me_result res[KVZ_CUDA_MAX_REFS];
  memset(res, 0, sizeof(res));

  if (timing) t1 = glue_now();
  if (n_refs > 0) {
    bool have = false;
    glue_spec *e = glue_speculate() ? glue_spec_find(frame_id, handle, lambda, &in) : NULL;
    if (e) {                                             /* launched ahead with exactly these records */
      const unsigned long long w0 = timing ? glue_now() : 0;
      have = kvz_cuda_enc_search_end(enc, e->ticket, e->idx, res) != 0;
      e->valid = 0;
      if (timing && have && width == 8) {
        __atomic_fetch_add(&g_origin8[e->origin][0], 1ull, __ATOMIC_RELAXED);
        __atomic_fetch_add(&g_origin8[e->origin][1], glue_now() - w0, __ATOMIC_RELAXED);
      }
    }
    __atomic_fetch_add(&g_spec_cnt[have ? 0 : 1], 1ull, __ATOMIC_RELAXED);
    if (kvz_glue_trace() && x < 64 && y < 64 && state->frame->num == 1) fprintf(stderr, "T search %d %d %d %s\n", x, y, width, have ? "hit" : "miss");
    __atomic_fetch_add(&g_spec_by_size[width > 32 ? 3 : width > 16 ? 2 : width > 8 ? 1 : 0][have ? 0 : 1], 1ull, __ATOMIC_RELAXED);
    if (!have) {                                         /* search it now; its top-left children ride along */
      glue_batch b;
      b.n = b.n_needed_now = 1;
      b.cur_origin = 0; b.origin[0] = 0;
      b.sent[0] = in;
      if (square && glue_speculate()) glue_batch_add_chain(state, &b, frame_id, handle, lambda, x, y, depth + 1, lcu);
      const int ticket = glue_batch_launch(enc, frame_id, handle, lambda, &b);
      if (!kvz_cuda_enc_search_end(enc, ticket, 0, res)) kvz_glue_die("kvz_cuda_enc_search_end");
      t_spec[ticket & 0xFF][0].valid = 0;
    }
    KVZ_GLUE_COUNT(KVZ_GLUE_CNT_ME_PU);
  }
  if (timing) t2 = glue_now();

  glue_apply_results(cur_cu, res, n_refs, ref_list, lx_idx, inter_cost, inter_bitcost);

  /* bi-prediction from pairs of merge candidates stays the reference's own function (:1304-1434): it predicts with
   * the ipol strategies and costs with kvz_satd_any_size; info.mv_cand holds the last picture's candidates as there */
  const bool can_use_bipred = state->frame->slicetype == KVZ_SLICE_B && cfg->bipred && width + height >= 16;
  if (can_use_bipred) {
    if (n_refs > 0) { info.ref_idx = n_refs - 1; info.ref = state->frame->ref->images[n_refs - 1]; }
    search_pu_inter_bipred(&info, depth, lcu, cur_cu, inter_cost, inter_bitcost);
  }

  if (n_refs > 0 && square && glue_speculate())
    glue_launch_next(state, enc, frame_id, handle, lambda, x, y, depth, lcu, cur_cu, *inter_cost < MAX_INT);
  if (timing) {
    const unsigned long long t3 = glue_now();
    __atomic_fetch_add(&g_glue_ns[0], t1 - t0, __ATOMIC_RELAXED);
    __atomic_fetch_add(&g_glue_ns[1], t2 - t1, __ATOMIC_RELAXED);
    __atomic_fetch_add(&g_glue_ns[2], t3 - t2, __ATOMIC_RELAXED);
    __atomic_fetch_add(&g_glue_ns[3], 1ull, __ATOMIC_RELAXED);
  }
}

/* kvz_search_cu_inter (search_inter.c:1588-1610) */
void kvz_search_cu_inter(encoder_state_t *const state, int x, int y, int depth, lcu_t *lcu, double *inter_cost,
                         uint32_t *inter_bitcost)
{
  if (!(kvz_glue_mode() & KVZ_GLUE_BATCH)) {
    kvz_search_cu_inter_host(state, x, y, depth, lcu, inter_cost, inter_bitcost);
    return;
  }
  glue_search_pu_inter(state, x, y, depth, SIZE_2Nx2N, 0, lcu, inter_cost, inter_bitcost);
  if (state->encoder_control->cfg.rdo >= 2) kvz_cu_cost_inter_rd2(state, x, y, depth, lcu, inter_cost, inter_bitcost);
}

/* kvz_search_cu_smp (search_inter.c:1627-1700): SMP / AMP are refused when the session opens, so in batch mode this
 * is never reached with them enabled; the reference's function serves the other modes */
void kvz_search_cu_smp(encoder_state_t *const state, int x, int y, int depth, part_mode_t part_mode, lcu_t *lcu,
                       double *inter_cost, uint32_t *inter_bitcost)
{
  if (kvz_glue_mode() & KVZ_GLUE_BATCH) kvz_glue_die("kvz_search_cu_smp: SMP / AMP partitions are not searched on the device");
  kvz_search_cu_smp_host(state, x, y, depth, part_mode, lcu, inter_cost, inter_bitcost);
}
