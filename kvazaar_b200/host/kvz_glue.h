/*
 * kvazaar_b200/host/kvz_glue.h -- shared state of the host glue that binds libkvz_cuda.so into Kvazaar.
 *
 * The glue is the handful of translation units a Kvazaar maintainer adds: each one #includes the
 * reference source file it extends IN PLACE from the reference tree at build time (nothing of the reference is
 * copied into this repository), renames the one function it takes over, and defines the replacement.  The result is
 * libkvazaar_cuda.so / kvazaar_cuda: the unmodified reference encoder with the CUDA path behind its own
 * strategy selector (src/strategyselector.c:43-213) and behind kvz_search_cu_inter (src/search_inter.c:1588).
 *
 * Mode, from the environment variable KVZ_CUDA_MODE:
 *   batch    (default) motion search of every PU on the device through the encoder-session face
 *            (include/kvz_cuda_enc.h); the per-call "cuda" strategies are not registered, so what stays on the host
 *            (intra search, residual coding, in-loop filters, CABAC) runs the reference's own SIMD strategies.
 *   percall  the per-call "cuda" strategies (include/kvz_cuda_strategies.h) are registered at priority 50 and win
 *            every strategy type unless KVAZAAR_OVERRIDE_<type> says otherwise (src/strategyselector.c:258-324);
 *            motion search runs the reference's host code over those strategies.  One launch per call: for parity.
 *   both     percall + batch.
 *   off      the reference as it is.
 * KVZ_CUDA_DEVICE (else LOCAL_RANK, else 0) selects the GPU of this process.
 */
#ifndef KVZ_GLUE_H_
#define KVZ_GLUE_H_

#include <stdint.h>

enum { KVZ_GLUE_OFF = 0, KVZ_GLUE_BATCH = 1, KVZ_GLUE_PERCALL = 2, KVZ_GLUE_BOTH = 3 };

int kvz_glue_mode(void);          /* parsed once from KVZ_CUDA_MODE */
int kvz_glue_device(void);

/* per-family counters of calls served by the CUDA library, printed at exit when KVZ_CUDA_STATS is set and readable
 * through kvz_glue_counters() (tests check them to prove which path ran) */
enum { KVZ_GLUE_CNT_PICTURE, KVZ_GLUE_CNT_DCT, KVZ_GLUE_CNT_QUANT, KVZ_GLUE_CNT_INTRA, KVZ_GLUE_CNT_IPOL,
       KVZ_GLUE_CNT_SAO, KVZ_GLUE_CNT_NAL, KVZ_GLUE_CNT_ME_PU, KVZ_GLUE_CNT_ME_FRAMES, KVZ_GLUE_CNT_N };
extern unsigned long long kvz_glue_count[KVZ_GLUE_CNT_N];
#define KVZ_GLUE_COUNT(family) __atomic_fetch_add(&kvz_glue_count[family], 1ull, __ATOMIC_RELAXED)
void kvz_glue_counters(unsigned long long out[KVZ_GLUE_CNT_N]);

void kvz_glue_session_close(void);  /* kvz_glue_search_inter.c: drop the encoder session (kvz_api encoder_open / encoder_close) */

void kvz_glue_die(const char *what);   /* no CPU fallback: report kvz_cuda_last_error() and abort */

#endif
