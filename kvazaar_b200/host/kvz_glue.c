/* kvazaar_b200/host/kvz_glue.c -- mode, device and counters of the host glue (see kvz_glue.h) */
#include "kvz_glue.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "kvz_cuda_strategies.h"

unsigned long long kvz_glue_count[KVZ_GLUE_CNT_N];

static const char *const glue_count_names[KVZ_GLUE_CNT_N] = { "picture", "dct", "quant", "intra", "ipol", "sao", "nal",
                                                              "me_pu_searches", "me_frames" };

void kvz_glue_counters(unsigned long long out[KVZ_GLUE_CNT_N])
{
  for (int i = 0; i < KVZ_GLUE_CNT_N; i++) out[i] = __atomic_load_n(&kvz_glue_count[i], __ATOMIC_RELAXED);
}

static void glue_print_stats(void)
{
  fprintf(stderr, "kvz_glue: mode %d, kernels launched %llu;", kvz_glue_mode(), kvz_cuda_launch_count());
  for (int i = 0; i < KVZ_GLUE_CNT_N; i++) fprintf(stderr, " %s=%llu", glue_count_names[i], kvz_glue_count[i]);
  unsigned long long pc[7];
  kvz_cuda_percall_counts(pc);
  fprintf(stderr, "; per-call strategies served: picture=%llu dct=%llu quant=%llu intra=%llu ipol=%llu sao=%llu nal=%llu\n",
          pc[0], pc[1], pc[2], pc[3], pc[4], pc[5], pc[6]);
}

int kvz_glue_mode(void)
{
  static int mode = -1;
  if (mode < 0) {
    const char *m = getenv("KVZ_CUDA_MODE");
    int v = KVZ_GLUE_BATCH;
    if (m) {
      if (!strcmp(m, "off")) v = KVZ_GLUE_OFF;
      else if (!strcmp(m, "batch")) v = KVZ_GLUE_BATCH;
      else if (!strcmp(m, "percall")) v = KVZ_GLUE_PERCALL;
      else if (!strcmp(m, "both")) v = KVZ_GLUE_BOTH;
      else { fprintf(stderr, "kvz_glue: KVZ_CUDA_MODE must be off, batch, percall or both (got '%s')\n", m); abort(); }
    }
    if (getenv("KVZ_CUDA_STATS")) atexit(glue_print_stats);
    mode = v;
  }
  return mode;
}

int kvz_glue_device(void)
{
  const char *v = getenv("KVZ_CUDA_DEVICE");
  if (!v) v = getenv("LOCAL_RANK");
  return v ? atoi(v) : 0;
}

void kvz_glue_die(const char *what)
{
  const char *e = kvz_cuda_last_error();
  fprintf(stderr, "kvz_glue: %s%s%s -- there is no CPU fallback on this path\n", what, (e && *e) ? ": " : "", (e && *e) ? e : "");
  abort();
}
