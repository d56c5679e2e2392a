/*
 * tests/stub/enc_stub.c -- TEST INFRASTRUCTURE ONLY.  A stand-in for the device behind the encoder-session face
 * (include/kvz_cuda_enc.h), preloaded (LD_PRELOAD) in front of libkvz_cuda.so so that the HOST glue of
 * kvazaar_b200/host/ can be exercised on a machine without a GPU: its PU walk, the records it derives, and above all
 * its launching of PUs ahead of the walk.  This is NOT a motion search: the "results" are a hash of exactly the
 * records a launch carries (PU, jobs, lambda, frame), i.e. a pure function of what the device is given -- which is
 * the property the glue relies on when it takes the results of a launch made ahead.  A glue that matched a launch
 * made with other records, or of another frame, produces a different bitstream than the same run without launching
 * ahead; tests/test_glue_speculation.py checks that they are equal.  Never linked into the product.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "kvz_cuda_enc.h"

struct kvz_cuda_enc { kvz_cuda_enc_cfg cfg; uint64_t frames[64]; int n_frames; unsigned long long stats[8]; };

#define SLOTS 64
#define PUS 8
static __thread struct {
  uint32_t gen[SLOTS], next;
  int n_pus[SLOTS];
  kvz_cuda_me_result res[SLOTS][PUS][KVZ_CUDA_MAX_REFS];
  int n_jobs[SLOTS][PUS];
} t;
static char g_err[256] = "";

const char *kvz_cuda_last_error(void) { return g_err; }
static int fail(const char *m) { strncpy(g_err, m, sizeof(g_err) - 1); return 0; }

static uint64_t fnv(uint64_t h, const void *p, size_t n)
{
  const uint8_t *b = p;
  for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
  return h;
}

kvz_cuda_enc *kvz_cuda_enc_open(const kvz_cuda_enc_cfg *cfg)
{
  kvz_cuda_enc *e = calloc(1, sizeof(*e));
  if (e) e->cfg = *cfg;
  return e;
}
void kvz_cuda_enc_close(kvz_cuda_enc *e) { free(e); }

int kvz_cuda_enc_frame_begin(kvz_cuda_enc *e, const kvz_cuda_enc_frame *f)
{
  for (int i = 0; i < 64; i++) if (e->frames[i] == f->frame_id + 1) return i;
  const int i = __atomic_fetch_add(&e->n_frames, 1, __ATOMIC_SEQ_CST) % 64;
  e->frames[i] = f->frame_id + 1;
  return i;
}

int kvz_cuda_enc_search_begin(kvz_cuda_enc *e, int handle, int lambda, int n_pus, const kvz_cuda_me_pu *pus,
                              const kvz_cuda_me_job *jobs, int *ticket)
{
  if (n_pus < 1 || n_pus > PUS) return fail("stub: 1..8 PUs per launch");
  if (handle < 0 || handle >= 64 || !e->frames[handle]) return fail("stub: bad frame handle");
  const uint32_t slot = t.next++ % SLOTS, gen = ++t.gen[slot];
  t.n_pus[slot] = n_pus;
  for (int p = 0; p < n_pus; p++) {
    kvz_cuda_me_pu pu = pus[p];
    pu.first_job = 0;
    uint32_t inter_cost = 0x7FFFFFFFu;
    t.n_jobs[slot][p] = pu.num_jobs;
    for (int r = 0; r < pu.num_jobs; r++) {
      uint64_t h = fnv(1469598103934665603ull, &pu, sizeof(pu));
      h = fnv(h, &jobs[p * KVZ_CUDA_MAX_REFS + r], sizeof(kvz_cuda_me_job));
      h = fnv(h, &lambda, sizeof(lambda));
      h = fnv(h, &e->frames[handle], sizeof(uint64_t));
      kvz_cuda_me_result *o = &t.res[slot][p][r];
      memset(o, 0, sizeof(*o));
      o->mv[0] = (int16_t)((int)(h & 63) - 32);
      o->mv[1] = (int16_t)((int)(h >> 6 & 63) - 32);
      o->cost = 2000 + (uint32_t)(h >> 12 & 0xFFF) * (uint32_t)(pu.w / 8);
      o->bitcost = 8 + (uint32_t)(h >> 24 & 15);
      o->mv_cand_idx = (uint8_t)(h >> 28 & 1);
      o->took = o->cost < inter_cost;
      if (o->took) inter_cost = o->cost;
    }
  }
  e->stats[1] += (unsigned long long)n_pus;
  *ticket = (int)(slot | (gen & 0x7FFFFFu) << 8);
  return 1;
}

int kvz_cuda_enc_search_end(kvz_cuda_enc *e, int ticket, int pu_index, kvz_cuda_me_result *results)
{
  (void)e;
  const uint32_t slot = (uint32_t)ticket & 0xFF;
  if (slot >= SLOTS || ((uint32_t)ticket >> 8) != (t.gen[slot] & 0x7FFFFFu)) return fail("stub: the ticket has expired");
  if (pu_index < 0 || pu_index >= t.n_pus[slot]) return fail("stub: bad pu_index");
  memcpy(results, t.res[slot][pu_index], (size_t)t.n_jobs[slot][pu_index] * sizeof(kvz_cuda_me_result));
  return 1;
}

int kvz_cuda_enc_search_ready(kvz_cuda_enc *e, int ticket, int pu_index)
{
  (void)e;
  const uint32_t slot = (uint32_t)ticket & 0xFF;
  if (ticket < 0 || slot >= SLOTS || ((uint32_t)ticket >> 8) != (t.gen[slot] & 0x7FFFFFu) || pu_index < 0 || pu_index >= t.n_pus[slot]) return -1;
  return 1;                                   /* the stand-in "searches" inside begin */
}

int kvz_cuda_enc_search_pu(kvz_cuda_enc *e, int handle, int lambda, const kvz_cuda_me_pu *pu, const kvz_cuda_me_job *jobs,
                           kvz_cuda_me_result *results)
{
  kvz_cuda_me_job padded[KVZ_CUDA_MAX_REFS];
  memset(padded, 0, sizeof(padded));
  memcpy(padded, jobs, (size_t)pu->num_jobs * sizeof(kvz_cuda_me_job));
  int ticket;
  return kvz_cuda_enc_search_begin(e, handle, lambda, 1, pu, padded, &ticket) && kvz_cuda_enc_search_end(e, ticket, 0, results);
}

int kvz_cuda_enc_preload_source(kvz_cuda_enc *e, int64_t pts, const void *luma, int32_t stride)
{
  (void)e; (void)pts; (void)luma; (void)stride;
  return 1;
}
void kvz_cuda_enc_stats(const kvz_cuda_enc *e, unsigned long long s[8]) { memcpy(s, e->stats, sizeof(e->stats)); }
void kvz_cuda_enc_measurements(kvz_cuda_enc *e, unsigned long long m[8]) { (void)e; memset(m, 0, 8 * sizeof(*m)); }
