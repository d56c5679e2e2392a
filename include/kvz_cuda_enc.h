/*
 * include/kvz_cuda_enc.h -- ENCODER-SESSION face of libkvz_cuda.so: what a Kvazaar host whose pictures live in
 * HOST memory calls while it encodes (kvazaar_b200/host/ binds it into the reference encoder; INTEGRATION.md
 * section 2).  It wraps the batch face of kvz_cuda.h: the session keeps the source and the border-replicated
 * reference planes and the dense SAD maps of the frame in HBM, and answers one search_pu_inter()
 * (src/search_inter.c:1451-1524) per call -- the order in which those calls arrive is the host's CTU
 * wavefront (src/encoderstate.c:777-828), so every PU is searched with the candidates its already-decided
 * neighbours produce and the decisions equal the reference's.
 *
 * extern "C", plain pointers and sizes; returns follow kvz_cuda.h (1 ok / 0 fail, kvz_cuda_last_error()).
 * NO CPU fallback: without a usable B200 kvz_cuda_enc_open() returns NULL.
 */
#ifndef KVZ_CUDA_ENC_H_
#define KVZ_CUDA_ENC_H_

#include "kvz_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct kvz_cuda_enc kvz_cuda_enc;

/* the fields of encoder_control_t / kvz_config the motion search reads (src/search_inter.c:87-181, :1208-1256) */
typedef struct kvz_cuda_enc_cfg {
  int32_t device;               /* CUDA device of this process (one process per GPU) */
  int32_t width, height;        /* encoder_control->in.width / height: luma size of the whole picture */
  int32_t bitdepth;             /* 8 (KVZ_BIT_DEPTH; 10: not yet on this face) */
  int32_t ime_search;           /* enum kvz_cuda_ime_search from cfg.ime_algorithm */
  int32_t range;                /* search range of the full search: 8, 16, 32, 64 (search_inter.c:1208-1215); else 32 */
  int32_t me_early_termination; /* cfg.me_early_termination */
  int32_t me_max_steps;         /* cfg.me_max_steps */
  int32_t fme_level;            /* cfg.fme_level */
  int32_t owf, wpp;             /* cfg.owf, cfg.wpp (fracmv_within_tile) */
  int32_t sao_type, deblock_enable, mv_constraint;
  int32_t max_ref_lcu_right, max_ref_lcu_down;   /* encoder_control->max_inter_ref_lcu (encoder.c:240-241) */
  int32_t max_map_mbytes;       /* cap of the dense SAD maps of the owf + 1 frames in flight, MiB; 0 = default (98304) */
} kvz_cuda_enc_cfg;

/* Replaces nothing in the reference; called once next to kvz_strategyselector_init (src/kvazaar.c:81). */
KVZ_CUDA_API kvz_cuda_enc *kvz_cuda_enc_open(const kvz_cuda_enc_cfg *cfg);
KVZ_CUDA_API void kvz_cuda_enc_close(kvz_cuda_enc *enc);

/* one picture of state->frame->ref (src/imagelist.h:38-48) */
typedef struct kvz_cuda_enc_ref {
  const void *luma;             /* ref->images[i]->y: reconstructed luma in HOST memory; complete unless in_flight */
  int32_t stride;               /* samples per row */
  int32_t poc;                  /* ref->pocs[i] */
  uint64_t key;                 /* identity of the picture (its address): a picture already resident is not uploaded again */
  const int16_t *mv_hints;      /* quarter-pel (x, y) pairs sampled from the motion field ref->cu_arrays[i]: where the */
  int32_t n_mv_hints;           /* search windows of this frame will probably be centred (extra_mv, merge candidates) */
  int32_t in_flight;            /* --owf > 0: the picture is still being reconstructed by one of the owf previous encoder states
                                   (encoderstate.c:751-804).  Its 64x64 blocks are then uploaded, and this frame's SAD maps
                                   against it computed, CTU by CTU as the host's CTU dependencies guarantee them */
} kvz_cuda_enc_ref;

typedef struct kvz_cuda_enc_frame {
  uint64_t frame_id;            /* unique per (frame, tile): state->frame->num << 16 | state->tile->id */
  const void *src_luma;         /* source luma of the WHOLE picture in host memory (frame->source->base_image->y) */
  int32_t src_stride;
  uint64_t src_key;             /* identity of the source picture: tiles of one frame share one upload */
  int64_t src_pts;              /* frame->source->pts: a picture preloaded under this number is used without an upload */
  int32_t tile_x, tile_y, tile_w, tile_h;   /* state->tile->offset_x / y, state->tile->frame->width / height */
  uint8_t ref_LX[2][16];        /* state->frame->ref_LX */
  int32_t n_refs;               /* state->frame->ref->used_size */
  kvz_cuda_enc_ref refs[KVZ_CUDA_MAX_REFS];
} kvz_cuda_enc_frame;

/* Called once per (frame, tile) before its first search, from any thread: uploads the source and the references that
 * are not resident yet (references no longer listed are dropped), borders them, chooses one box of displacements
 * per reference (zero window + the windows the hints predict, at most 272 a side) and launches the dense SAD maps
 * of all references (kvz_cuda_ime_sad_box_maps_multi).  Does not wait for the device.  Returns a handle >= 0 for
 * kvz_cuda_enc_search_pu, or -1.  Replaces the per-candidate kvz_image_calc_sad calls of search_mv_full
 * (src/search_inter.c:886-956) for every PU of the frame. */
KVZ_CUDA_API int kvz_cuda_enc_frame_begin(kvz_cuda_enc *enc, const kvz_cuda_enc_frame *frame);

/* One search_pu_inter() reference loop (src/search_inter.c:1505-1511): `pu` with its merge candidates, `jobs[i]`
 * = the AMVP candidates / extra_mv of reference jobs[i].ref_idx (pu->num_jobs of them, ref_idx ascending;
 * pu->first_job is ignored), `results[i]` as kvz_cuda_me_search gives them.  lambda_sqrt_i =
 * (int)(state->lambda_sqrt + 0.5) of the calling CTU.  HOST pointers; blocks until the results are there.
 * Thread-safe: every calling thread owns a stream, calls of different threads overlap on the GPU. */
KVZ_CUDA_API int kvz_cuda_enc_search_pu(kvz_cuda_enc *enc, int frame_handle, int lambda_sqrt_i,
                                        const kvz_cuda_me_pu *pu, const kvz_cuda_me_job *jobs,
                                        kvz_cuda_me_result *results);

/* The same search split in two, so that a host can have searches in flight while it does its own work (residual
 * coding, intra search) and can launch a PU ahead of time with the candidates it EXPECTS the PU to get: begin sends
 * 1..8 PUs of one CTU in one launch and returns at once; end blocks until PU `pu_index` of that launch has its results.
 * `jobs` holds KVZ_CUDA_MAX_REFS records per PU (PU i's jobs at jobs[i * KVZ_CUDA_MAX_REFS ..], pus[i].num_jobs used).
 * A host that launched ahead must compare the PU and job records it sent with the real ones when the PU's turn comes
 * and search again when they differ -- the device computes exactly what it is given (kvazaar_b200/host/ does this, so
 * decisions stay the reference's).  Tickets belong to the calling thread; a thread's 64th-next begin reuses a ticket's
 * slot, after which end() on it fails ("expired").  Nothing in the reference corresponds to the split. */
KVZ_CUDA_API int kvz_cuda_enc_search_begin(kvz_cuda_enc *enc, int frame_handle, int lambda_sqrt_i, int n_pus,
                                           const kvz_cuda_me_pu *pus, const kvz_cuda_me_job *jobs, int *ticket);
KVZ_CUDA_API int kvz_cuda_enc_search_end(kvz_cuda_enc *enc, int ticket, int pu_index, kvz_cuda_me_result *results);
/* does not block: 1 when kvz_cuda_enc_search_end would return at once, 0 when the PU is still being searched, -1 when the
 * ticket has expired (or is not the calling thread's) */
KVZ_CUDA_API int kvz_cuda_enc_search_ready(kvz_cuda_enc *enc, int ticket, int pu_index);

/* Optional: put the source luma of the picture with presentation number `pts` into device memory ahead of time
 * (a host whose input is already on the device, or a benchmark that wants its inputs resident before the timed
 * region).  kvz_cuda_enc_frame_begin then uses it instead of uploading; it is dropped when its frame has begun.
 * Replaces nothing in the reference. */
KVZ_CUDA_API int kvz_cuda_enc_preload_source(kvz_cuda_enc *enc, int64_t pts, const void *luma, int32_t stride);

/* counters since open: stats[0] frames begun, [1] PU searches launched (including the ones launched ahead), [2] reference planes uploaded, [3] bytes of SAD maps
 * of the last frame, [4] map displacements of the last frame (sum over references of n_dx * n_dy),
 * [5] reference-jobs searched, [6] jobs not served entirely from a box (some window evaluated on demand: host-side
 * estimate from the candidate list) */
KVZ_CUDA_API void kvz_cuda_enc_stats(const kvz_cuda_enc *enc, unsigned long long stats[8]);

/* measurement counters since open (bench.py): m[0] bytes copied host -> device (planes), m[1] bytes of PU / job
 * arguments sent with the launches, m[2] bytes of results read back, m[3] dense SAD launches (ime_sad_multi_kernel)
 * whose duration is known, m[4] their summed duration in nanoseconds (CUDA events on the launching stream),
 * m[5] the pixel absolute differences those launches had to evaluate (8x8 level: 64 per 8x8 block and displacement,
 * SURVEY.md 8d), m[6] bytes of SAD maps they wrote, m[7] source planes taken from a preload */
KVZ_CUDA_API void kvz_cuda_enc_measurements(kvz_cuda_enc *enc, unsigned long long m[8]);

#ifdef __cplusplus
}
#endif
#endif /* KVZ_CUDA_ENC_H_ */
