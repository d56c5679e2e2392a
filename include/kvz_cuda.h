/*
 * include/kvz_cuda.h -- C ABI of libkvz_cuda.so, the B200 (sm_100a) implementation
 * of Kvazaar's inner encode loop.
 *
 * Every entry point is `extern "C"`, takes plain pointers and sizes, and returns
 * 1 on success / 0 on failure exactly like the reference's strategy registration
 * (src/strategyselector.c:216-256); kvz_cuda_last_error() then says why.  There
 * is NO CPU fallback: without a usable CUDA device every compute call fails.
 *
 * Two faces:
 *   (1) the PER-CALL face mirrors the reference's strategy typedefs byte for byte
 *       (host pointers in, host pointers out) so the functions can be registered
 *       with kvz_strategyselector_register() under strategy name "cuda"
 *       (priority 50) -- see INTEGRATION.md.  It exists for parity and for the
 *       kernel sweep; a launch per 8x8 block is not how throughput is obtained.
 *   (2) the BATCH face works on device-resident planes and arrays of jobs and is
 *       what a frame-level host loop calls (suggested names of SURVEY.md 8b).
 *
 * kvz_pixel is uint8_t (KVZ_BIT_DEPTH 8, src/kvazaar.h:72-77); entry points that
 * also exist for the 10-bit build carry a `bitdepth` argument and then take
 * uint16_t samples through `const void *`.
 */
#ifndef KVZ_CUDA_H_
#define KVZ_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define KVZ_CUDA_API __attribute__((visibility("default")))
#else
#define KVZ_CUDA_API
#endif

typedef uint8_t kvz_cuda_pixel;   /* kvz_pixel of the 8-bit build */

/* ------------------------------------------------------------------ */
/* lifecycle / plumbing                                                  */
/* ------------------------------------------------------------------ */

/* Select the device and create the library's streams and scratch buffers.
 * Replaces nothing in the reference; called from the glue's
 * kvz_strategy_register_*_cuda() (src/strategies/strategies-picture.c:66-88). */
KVZ_CUDA_API int kvz_cuda_init(int device);
KVZ_CUDA_API void kvz_cuda_shutdown(void);
KVZ_CUDA_API const char *kvz_cuda_last_error(void);
KVZ_CUDA_API int kvz_cuda_device_count(void);
/* number of kernels this library has launched since init (bench.py: gpu_launches) */
KVZ_CUDA_API unsigned long long kvz_cuda_launch_count(void);

/* Measured integer-pipe denominator for the SAD kernels' roofline: pixel
 * absolute differences per second a register-only VABSDIFF4.U8.ACC loop reaches on this
 * device.  Replaces nothing in the reference. */
KVZ_CUDA_API int kvz_cuda_measure_int_pipe_peak(double *abs_diffs_per_s);

/* device memory for host languages that have no CUDA binding of their own */
KVZ_CUDA_API void *kvz_cuda_malloc(size_t bytes);
KVZ_CUDA_API void kvz_cuda_free(void *dptr);
KVZ_CUDA_API void *kvz_cuda_malloc_host(size_t bytes);      /* pinned */
KVZ_CUDA_API void kvz_cuda_free_host(void *hptr);
KVZ_CUDA_API int kvz_cuda_memcpy_h2d(void *dst, const void *src, size_t bytes, void *stream);
KVZ_CUDA_API int kvz_cuda_memcpy_d2h(void *dst, const void *src, size_t bytes, void *stream);
KVZ_CUDA_API int kvz_cuda_memset(void *dst, int value, size_t bytes, void *stream);
KVZ_CUDA_API int kvz_cuda_stream_sync(void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: reference-picture preparation                             */
/* ------------------------------------------------------------------ */

/* Border a plane with `pad` replicated samples per side: reading the border is
 * the per-pixel coordinate clamp of image.c:320-452 / ipol-generic.c:761-781.
 * dst holds (w+2*pad) x (h+2*pad) samples, rows dst_pitch SAMPLES apart. */
KVZ_CUDA_API int kvz_cuda_pad_plane(const void *d_src, int w, int h, int src_pitch,
                                    void *d_dst, int pad, int dst_pitch, int bitdepth, void *stream);
/* border needed by kvz_cuda_ime_sad_maps() for a given search range */
KVZ_CUDA_API int kvz_cuda_ime_required_pad(int range);

/* ------------------------------------------------------------------ */
/* BATCH face: dense integer motion estimation  (rows A1-A3)             */
/* ------------------------------------------------------------------ */

typedef struct kvz_cuda_ime_desc {
  int32_t width, height;        /* luma size of the pictures */
  int32_t bitdepth;             /* 8 (10: see DESIGN.md) */
  int32_t range;                /* 8, 16, 32, 64 = --me full8..full64 (search_inter.c:1208-1215) */
  int32_t tile_x, tile_y, tile_w, tile_h;   /* tile searched; whole frame = 0,0,width,height */
  int32_t cur_pitch;            /* bytes per row of the current plane */
  int32_t ref_pitch;            /* bytes per row of the padded reference plane */
  int32_t ref_pad;              /* its border, >= kvz_cuda_ime_required_pad(range) + max(|center_x|, |center_y|) */
  int32_t center_x, center_y;   /* window centre in whole samples (any integers):
                                   0,0 = the zero window of search_mv_full (search_inter.c:891-895); another
                                   centre = the window round extra_mv / a merge candidate (search_inter.c:898-955) */
} kvz_cuda_ime_desc;

/* entries per row of the maps: roundup(2*range+1, 16) */
KVZ_CUDA_API int kvz_cuda_ime_map_pitch(int range);
/* bytes of the map of PU size 8<<level (level 0..3) for this descriptor */
KVZ_CUDA_API size_t kvz_cuda_ime_map_bytes(const kvz_cuda_ime_desc *d, int level);

/* For one (current, reference) pair: SAD of every 8x8/16x16/32x32/64x64 PU of every
 * CTU against the edge-clamped reference at every (dx,dy) in [-range,range]^2.
 * With K = 2*range+1 and n = 8 >> level, PU p = ctu * n*n + py*n + px owns K+1 rows of
 * map_pitch entries (uint16 for levels 0,1 and uint32 for 2,3):
 *   map[level][p][dy+range][dx+range]   for dx in [-range, range)   (rows 0..K-1)
 *   map[level][p][K][dy+range]          for dx = +range             (row K)
 * -- the last column is stored as a row so that every store of the kernels covers whole
 * 32-byte sectors; the remaining entries of a row (up to map_pitch) are unspecified.
 * Replaces the zero-window loop of search_mv_full (search_inter.c:891-895) over
 * check_mv_cost -> kvz_image_calc_sad -> kvz_reg_sad for ALL PUs of the frame. */
KVZ_CUDA_API int kvz_cuda_ime_sad_maps(const kvz_cuda_ime_desc *d,
                                       const void *d_cur, const void *d_ref_padded,
                                       uint16_t *d_sad8, uint16_t *d_sad16,
                                       uint32_t *d_sad32, uint32_t *d_sad64, void *stream);

/* The same for a rectangular box of displacements, dx in [x0, x0+n_dx), dy in [y0, y0+n_dy): one box that
 * covers all the windows search_mv_full visits for a reference (zero window, extra_mv window, merge-candidate
 * windows; search_inter.c:891-955) computes every displacement once, where separate windows overlap.
 * The maps are CHUNK-MAJOR: a chunk is 16 displacement columns (the last one 8 when n_dx is an odd multiple
 * of 8), and with P = PUs of the level, w_c = width of chunk c:
 *   entry (p, dy, dx) of a level lives at   16 * c * P * n_dy + (p * n_dy + (dy - y0)) * w_c + ((dx - x0) - 16 * c),
 *   c = (dx - x0) / 16   -- i.e. map[chunk][p][dy - y0][w_c]; each map holds exactly P * n_dy * n_dx entries.
 * (One CTA computes one (CTU, chunk): its output is one contiguous run per level; kvz_cuda_me_search reads a
 * PU's map as n_chunks contiguous runs.)
 * n_dx must be a multiple of 8 (round the box up: entries no window covers are ignored by
 * kvz_cuda_me_search), n_dy is free; both at most 272.
 * ref_pad must be >= 104 + max(|x0|, |x0+n_dx|, |y0|, |y0+n_dy|). */
typedef struct kvz_cuda_ime_box_desc {
  int32_t width, height, bitdepth;
  int32_t tile_x, tile_y, tile_w, tile_h;
  int32_t cur_pitch, ref_pitch, ref_pad;
  int32_t x0, y0, n_dx, n_dy;
} kvz_cuda_ime_box_desc;
KVZ_CUDA_API size_t kvz_cuda_ime_box_map_bytes(const kvz_cuda_ime_box_desc *d, int level);
KVZ_CUDA_API int kvz_cuda_ime_sad_box_maps(const kvz_cuda_ime_box_desc *d,
                                           const void *d_cur, const void *d_ref_padded,
                                           uint16_t *d_sad8, uint16_t *d_sad16,
                                           uint32_t *d_sad32, uint32_t *d_sad64, void *stream);

/* The box maps of up to 8 references of ONE current picture in a single launch (same results as one
 * kvz_cuda_ime_sad_box_maps call per reference; the picture/tile fields of every desc must agree).  One grid
 * instead of one per reference: no per-launch ramp and tail, heavy CTAs first. */
#define KVZ_CUDA_IME_MAX_MULTI 8
typedef struct kvz_cuda_ime_box_ref {
  kvz_cuda_ime_box_desc desc;
  const void *d_ref_padded;
  uint16_t *d_sad8, *d_sad16;
  uint32_t *d_sad32, *d_sad64;
} kvz_cuda_ime_box_ref;
KVZ_CUDA_API int kvz_cuda_ime_sad_box_maps_multi(const kvz_cuda_ime_box_ref *refs, int n_refs, const void *d_cur,
                                                 void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: per-PU motion decision  (rows A3, A4, A5, A7-A10)         */
/* ------------------------------------------------------------------ */

#define KVZ_CUDA_MAX_REFS 16      /* MAX_REF_PIC_COUNT (global.h) */

/* inter_merge_cand_t (inter.h:36-41) */
typedef struct kvz_cuda_merge_cand {
  uint8_t dir;                  /* 1 = L0, 2 = L1, 3 = bi */
  uint8_t ref[2];
  uint8_t reserved;
  int16_t mv[2][2];             /* quarter-pel */
} kvz_cuda_merge_cand;

/* one search_pu_inter() call (search_inter.c:1451-1524): a PU and its merge candidates */
typedef struct kvz_cuda_me_pu {
  int16_t x, y;                 /* luma origin, tile-relative (info->origin) */
  uint8_t w, h;                 /* multiples of 8, <= 64; other shapes (SMP / AMP sides 4, 12) are NOT searched: their jobs
                                   come back with cost 0xFFFFFFFF and took = 0xFE */
  uint8_t num_merge_cand;
  uint8_t num_jobs;             /* references searched: jobs[first_job .. first_job+num_jobs) in ref_idx order */
  uint32_t first_job;
  kvz_cuda_merge_cand merge_cand[5];
} kvz_cuda_me_pu;

/* one search_pu_inter_ref() call (search_inter.c:1134-1297) */
typedef struct kvz_cuda_me_job {
  int16_t mv_cand[2][2];        /* AMVP candidates for this reference (inter.c:1209) */
  int16_t extra_mv[2];          /* colocated CU's vector in the reference (search_inter.c:1188-1206) */
  uint8_t ref_idx;
  uint8_t reserved[3];
} kvz_cuda_me_job;

typedef struct kvz_cuda_me_result {
  int16_t mv[2];                /* info->best_mv, quarter-pel */
  uint32_t cost;                /* info->best_cost */
  uint32_t bitcost;             /* info->best_bitcost */
  uint8_t merged, merge_idx;    /* search_inter.c:1260-1273 */
  uint8_t mv_cand_idx;          /* select_mv_cand (search_inter.c:1277-1279) */
  uint8_t took;                 /* 1 if this reference lowered *inter_cost (search_inter.c:1281) */
} kvz_cuda_me_result;

#define KVZ_CUDA_MAX_MAPSETS 4

/* one result of kvz_cuda_ime_sad_maps() for (current, this reference) and the window centre it was made
 * for (n_dx == 0), or of kvz_cuda_ime_sad_box_maps() (n_dx, n_dy > 0; center_x, center_y = x0, y0 of the box) */
typedef struct kvz_cuda_me_mapset {
  int32_t center_x, center_y;
  int32_t n_dx, n_dy;
  const uint16_t *d_sad8, *d_sad16;
  const uint32_t *d_sad32, *d_sad64;
} kvz_cuda_me_mapset;

typedef struct kvz_cuda_me_ref {
  const void *d_ref_padded;     /* border-replicated luma of the reference picture */
  int32_t ref_pitch, ref_pad;   /* bytes per row; border >= 84 = 64 + 20 (and what the maps needed) */
  int32_t n_mapsets;            /* a job whose windows all lie inside a box set is searched from that box; else each */
  int32_t reserved;             /* window whose centre matches a window set is read from it and every other window
                                   is evaluated on demand (0 sets = everything on demand) */
  kvz_cuda_me_mapset mapsets[KVZ_CUDA_MAX_MAPSETS];
} kvz_cuda_me_ref;

typedef struct kvz_cuda_me_frame {
  int32_t width, height, bitdepth;
  int32_t range;                /* search_range of search_mv_full */
  int32_t tile_x, tile_y, tile_w, tile_h;
  const void *d_cur;            /* source luma */
  int32_t cur_pitch;
  int32_t lambda_sqrt_i;        /* (int32_t)(state->lambda_sqrt + 0.5), search_inter.c:411 */
  int32_t fme_level;            /* cfg.fme_level 0..4 */
  int32_t owf, wpp, sao_type, deblock_enable, mv_constraint;   /* inputs of fracmv_within_tile */
  int32_t max_ref_lcu_right, max_ref_lcu_down;                 /* ctrl->max_inter_ref_lcu (encoder.c:240-241) */
  uint8_t ref_LX[2][16];        /* state->frame->ref_LX */
  int32_t n_refs;
  kvz_cuda_me_ref refs[KVZ_CUDA_MAX_REFS];   /* indexed by ref_idx */
  /* integer search algorithm (cfg.ime_algorithm, search_inter.c:1208-1235); appended, 0 keeps the full search */
  int32_t ime_search;           /* enum kvz_cuda_ime_search */
  int32_t me_early_termination; /* cfg.me_early_termination: 0 off, 1 on, 2 sensitive (search_inter.c:415-460) */
  int32_t me_max_steps;         /* cfg.me_max_steps, -1 = unlimited (search_inter.c:715, 828) */
} kvz_cuda_me_frame;

enum kvz_cuda_ime_search {
  KVZ_CUDA_IME_FULL = 0,        /* search_mv_full over [-range, range]^2 windows (dense maps where present) */
  KVZ_CUDA_IME_HEXBS = 1,       /* hexagon_search (search_inter.c:690-778); probes are evaluated on demand, no maps */
  KVZ_CUDA_IME_DIA = 2,         /* diamond_search (search_inter.c:796-883) */
  KVZ_CUDA_IME_TZ = 3           /* tz_search (search_inter.c:595-671) with the reference's compiled-in parameters */
};

/* For each PU and each of its jobs: integer full search (search_mv_full), then fractional
 * refinement (search_frac) or SATD at the integer vector, the merge-index match and the AMVP
 * candidate choice, with *inter_cost carried over the PU's reference loop -- the body of
 * search_pu_inter_ref (search_inter.c:1134-1297) for all PUs of a batch at once.
 * d_pus, d_jobs, d_results are device arrays; results[i] belongs to jobs[i]. */
KVZ_CUDA_API int kvz_cuda_me_search(const kvz_cuda_me_frame *frame, const kvz_cuda_me_pu *d_pus, int n_pus,
                                    const kvz_cuda_me_job *d_jobs, kvz_cuda_me_result *d_results, void *stream);

/* The same results from ONE launch with one CTA per PU (integer search, then refinement, inside the CTA): the form
 * for a host that searches one PU -- or the PUs of one CTU anti-diagonal -- at a time, where launch and tail
 * latency matter more than throughput (kvz_cuda_enc_search_pu, kvz_cuda_enc.h, calls it with n_pus = 1).  A window
 * that no box map set holds is evaluated from a copy of the reference region in shared memory by the whole CTA
 * (square window sets, n_dx == 0, are not consulted here).  d_pus / d_jobs / d_results may be mapped host memory. */
KVZ_CUDA_API int kvz_cuda_me_search_fused(const kvz_cuda_me_frame *frame, const kvz_cuda_me_pu *d_pus, int n_pus,
                                          const kvz_cuda_me_job *d_jobs, kvz_cuda_me_result *d_results, void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: block costs  (rows A1, A2, A6, A7, A8, A16)               */
/* ------------------------------------------------------------------ */

enum kvz_cuda_cost_kind {
  KVZ_CUDA_COST_REG_SAD = 0,    /* reg_sad_func (strategies-picture.h:102): any w x h, no bit-depth shift */
  KVZ_CUDA_COST_SAD_NXN = 1,    /* cost_pixel_nxn_func sad_NxN: w = h = strides = N */
  KVZ_CUDA_COST_SATD_NXN = 2,   /* cost_pixel_nxn_func satd_NxN, N = 4..64 */
  KVZ_CUDA_COST_SATD_ANY = 3,   /* cost_pixel_any_size_func kvz_satd_any_size: w, h multiples of 4 */
  KVZ_CUDA_COST_SATD_QUAD = 4,  /* one prediction of kvz_satd_any_size_quad (its 4-wide strips are dropped) */
  KVZ_CUDA_COST_SSD = 5         /* pixels_calc_ssd_func: w x w */
};

typedef struct kvz_cuda_block_job {
  uint32_t a_offset, b_offset;  /* offsets of the two blocks in the buffers d_a, d_b, in SAMPLES (bytes at 8 bits;
                                   at bitdepth 10 the buffers hold uint16_t samples and strides count samples too) */
  int32_t a_stride, b_stride;   /* samples per row */
  int16_t w, h;
  int32_t kind;                 /* enum kvz_cuda_cost_kind */
} kvz_cuda_block_job;

/* out[i] = cost of jobs[i]; dual / quad variants are 2 / 4 jobs sharing one original block. */
KVZ_CUDA_API int kvz_cuda_block_costs(const void *d_a, const void *d_b, const kvz_cuda_block_job *d_jobs,
                                      int n_jobs, int bitdepth, uint32_t *d_out, void *stream);

enum kvz_cuda_image_cost_kind { KVZ_CUDA_IMAGE_SAD = 0, KVZ_CUDA_IMAGE_SATD = 1 };

typedef struct kvz_cuda_image_cost_job {
  int32_t pic_x, pic_y;         /* block in the source picture */
  int32_t ref_x, ref_y;         /* block in the reference, may lie partly or wholly outside */
  int16_t w, h;
  int32_t kind;                 /* enum kvz_cuda_image_cost_kind */
} kvz_cuda_image_cost_job;

/* kvz_image_calc_sad (image.c:455-479) / kvz_image_calc_satd (image.c:488-545) for a batch of
 * block pairs on two UNPADDED planes of width x height: the reference coordinates are clamped
 * per sample exactly like image.c:320-452. */
KVZ_CUDA_API int kvz_cuda_image_costs(const void *d_pic, int pic_pitch, const void *d_ref, int ref_pitch,
                                      int width, int height, const kvz_cuda_image_cost_job *d_jobs,
                                      int n_jobs, int bitdepth, uint32_t *d_out, void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: picture-level reductions              */
/* ------------------------------------------------------------------ */

typedef struct kvz_cuda_plane {
  const void *d_data;           /* first sample of the plane (device) */
  int32_t width, height;        /* samples */
  int32_t pitch;                /* bytes per row */
} kvz_cuda_plane;

/* Picture-hash SEI checksum of up to 3 planes in one launch: d_checksums[p] = what array_checksum_func
 * kvz_array_checksum (strategies/strategies-nal.h:42-47, strategies/generic/nal-generic.c:58-82; caller
 * kvz_image_checksum nal.c:60-76) computes for plane p before it unpacks the sum big-endian into
 * checksum_out[0..3] -- the host glue does that unpacking.  bitdepth 8: uint8 samples; > 8: uint16 samples,
 * both bytes hashed.  kvz_array_md5 is a serial chain over the plane and stays on the host. */
KVZ_CUDA_API int kvz_cuda_picture_checksum(const kvz_cuda_plane *planes, int n_planes, int bitdepth,
                                           uint32_t *d_checksums, void *stream);

/* d_sse[p] = sum over plane p of (src - rec)^2, the sums compute_psnr (encmain.c:101-131) turns into
 * 10 * log10(num_pixels * PIXEL_MAX^2 / sse) on the host; planes of src and rec must have equal sizes. */
KVZ_CUDA_API int kvz_cuda_picture_sse(const kvz_cuda_plane *src, const kvz_cuda_plane *rec, int n_planes,
                                      int bitdepth, uint64_t *d_sse, void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: deblocking filter                     */
/* ------------------------------------------------------------------ */

/* The fields of cu_info_t (cu.h:134-175) the filter reads; one record per 4x4 SCU of the frame (tile), raster
 * order, width / 4 records per row -- the device-side mirror of frame->cu_array. */
typedef struct kvz_cuda_scu {
  uint8_t type;                 /* cu_type_t: 1 CU_INTRA, 2 CU_INTER */
  uint8_t depth, tr_depth;      /* CU width = 64 >> depth, TU width = 64 >> tr_depth */
  uint8_t part_size;            /* part_mode_t 0..7 */
  uint8_t cbf_y;                /* cbf_is_set(cu->cbf, cu->tr_depth, COLOR_Y) */
  uint8_t mv_dir;               /* inter.mv_dir */
  uint8_t mv_ref[2];            /* inter.mv_ref */
  int16_t mv[2][2];             /* inter.mv */
  int8_t qp;                    /* cu->qp (read when cu_qp != 0) */
  uint8_t reserved[3];
} kvz_cuda_scu;                 /* 20 bytes */

typedef struct kvz_cuda_deblock_cfg {
  int32_t width, height;        /* state->tile->frame size in luma samples, multiples of 8 */
  int32_t qp;                   /* state->qp */
  int32_t cu_qp;                /* encoder_control->max_qp_delta_depth >= 0 (get_qp_y_pred, filter.c:263-285) */
  int32_t beta_offset_div2, tc_offset_div2;   /* cfg.deblock_beta, cfg.deblock_tc */
  int32_t slice_b;              /* state->frame->slicetype == KVZ_SLICE_B */
  int32_t reserved;
  uint8_t ref_LX[2][16];        /* state->frame->ref_LX (B-slice strength rule, filter.c:436-441) */
} kvz_cuda_deblock_cfg;

/* kvz_filter_deblock_lcu (filter.c:770-779) for every LCU of the picture: planes[0] luma and, with n_planes == 3,
 * the two chroma planes of a 4:2:0 picture, filtered IN PLACE.  Two launches: all vertical edges, then all
 * horizontal edges (the order the reference's LCU walk and its postponed rightmost columns implement). */
KVZ_CUDA_API int kvz_cuda_deblock_picture(const kvz_cuda_deblock_cfg *cfg, const kvz_cuda_scu *d_scus,
                                          const kvz_cuda_plane *planes, int n_planes, int bitdepth, void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: inter candidate derivation  (row A18)                     */
/* ------------------------------------------------------------------ */

typedef struct kvz_cuda_cand_frame {
  int32_t width, height;          /* state->tile->frame size in luma samples (spatial candidates) */
  int32_t in_width, in_height;    /* encoder_control->in.width / height (temporal candidate availability) */
  int32_t tile_offset_x, tile_offset_y;   /* state->tile->offset_x / y (colocated start vector) */
  int32_t poc;                    /* state->frame->poc */
  int32_t slice_b;                /* state->frame->slicetype == KVZ_SLICE_B */
  int32_t tmvp_enable;            /* cfg.tmvp_enable */
  int32_t n_refs;                 /* state->frame->ref->used_size */
  int32_t ref_pocs[16];           /* state->frame->ref->pocs */
  int32_t col_ref_pocs[2][16];    /* for c = ref_LX[0][0]: ref->images[c]->ref_pocs[ref->ref_LXs[c][list][idx]]
                                     (add_temporal_candidate, inter.c:1058-1066) */
  uint8_t ref_LX[2][16];          /* state->frame->ref_LX */
  uint8_t ref_LX_size[2];
  uint8_t reserved[2];
  int32_t ref_scu_pitch;          /* records per row of the references' SCU arrays (ref_cu_array->width / 4) */
  const kvz_cuda_scu *d_scus;     /* this frame (tile): width / 4 records per row, what is coded so far */
  const kvz_cuda_scu *d_ref_scus[16];   /* state->frame->ref->cu_arrays, indexed like ref_pocs */
} kvz_cuda_cand_frame;

/* For each PU d_pus[i] (in: x, y, w, h, first_job, num_jobs) fills merge_cand[5] / num_merge_cand like
 * kvz_inter_get_merge_cand (inter.c:1314-1446; the unused list of a uni-directional candidate reads mv 0 / ref 255
 * as inter_clear_cu_unused leaves it) and, for each of its jobs (in: ref_idx), mv_cand like kvz_inter_get_mv_cand
 * (inter.c:1209-1228) for the list and index that hold ref_idx (search_inter.c:1142-1162) and extra_mv like
 * search_pu_inter_ref (search_inter.c:1188-1206).  d_pu_flags (may be NULL): bit 0 = merge candidate A1 must not be
 * used, bit 1 = B1 must not be used (second PUs of Nx2N.. / 2NxN.., search_inter.c:1472-1477).  The neighbours are
 * read from d_scus at the picture positions the reference reads through the working lcu_t, so the array must hold
 * the decisions of every CU coded before the PU. */
KVZ_CUDA_API int kvz_cuda_inter_candidates(const kvz_cuda_cand_frame *frame, kvz_cuda_me_pu *d_pus, int n_pus,
                                           const uint8_t *d_pu_flags, kvz_cuda_me_job *d_jobs, void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: sample adaptive offset                */
/* ------------------------------------------------------------------ */

/* The fields of sao_info_t (sao.h:43-51) the filter reads, per CTU; luma CTUs and chroma CTUs have one array
 * each (frame->sao_luma / frame->sao_chroma), raster order over the picture's CTUs. */
typedef struct kvz_cuda_sao_params {
  int8_t type;                  /* sao_type: 0 none, 1 band, 2 edge */
  int8_t eo_class;              /* sao_eo_class 0..3 (neighbour offsets sao.h:59-64) */
  int8_t band_position[2];      /* [0] Y / U, [1] V */
  int8_t offsets[10];           /* [category] for Y / U, [5 + category] for V; band offsets at [1..4] / [6..9] */
  int8_t reserved[2];
} kvz_cuda_sao_params;          /* 16 bytes */

/* What the SAO search asks the pixels of one CTU of one colour plane for. */
typedef struct kvz_cuda_sao_stats {
  int32_t edge[4][2][5];        /* [eo_class]: cat_sum_cnt of calc_sao_edge_dir (sao-generic.c:90-117):
                                   [0][cat] = sum of (orig - rec), [1][cat] = count, block-interior samples */
  int32_t band[2][32];          /* sao_bands of calc_sao_bands (sao.c:245-260): [0][band] sum, [1][band] count */
} kvz_cuda_sao_stats;           /* 416 bytes */

/* One pass over the source (`orig`) and the deblocked reconstruction (`rec`) of a picture: plane 0 luma with
 * 64x64 CTUs, planes 1..2 chroma of a 4:2:0 picture with 32x32 CTUs (partial CTUs at the right / bottom, as
 * sao_search_luma / sao_search_chroma clip them, sao.c:577-637).  d_stats holds the records of plane 0's CTUs in
 * raster order, then plane 1's, then plane 2's.  Replaces the calls of calc_sao_edge_dir_func kvz_calc_sao_edge_dir
 * and calc_sao_bands in sao_search_edge_sao / sao_search_band_sao (sao.c:337-455) and, because
 *   kvz_sao_edge_ddistortion(eo_class, offsets) = sum_cat  n[cat] * h[cat]^2 - 2 * h[cat] * E[cat]   (h != 0)
 *   kvz_sao_band_ddistortion(pos, bands)       = sum_b<4  n[pos+b] * h[b]^2 - 2 * h[b] * E[pos+b]    (h != 0)
 * exactly (sao-generic.c:51-81, :158-182; the formula of sao.c:381-388), every sao_edge_ddistortion_func /
 * sao_band_ddistortion_func call of sao_search_best_mode (sao.c:465-575) as well. */
KVZ_CUDA_API int kvz_cuda_sao_ctu_stats(const kvz_cuda_plane *orig, const kvz_cuda_plane *rec, int n_planes,
                                        int bitdepth, kvz_cuda_sao_stats *d_stats, void *stream);

/* kvz_sao_reconstruct (sao.c:277-337) with sao_reconstruct_color_func kvz_sao_reconstruct_color
 * (sao-generic.c:120-155) for every CTU of the picture, i.e. the picture encoder_sao_reconstruct
 * (encoderstate.c:245-441) leaves once every LCU has been through it: rec_out = SAO(rec_in), each sample with the
 * parameters of the CTU containing it, neighbours read from rec_in, samples whose edge neighbour lies outside
 * the picture copied.  rec_out must not alias rec_in.  Plane / CTU conventions as above. */
KVZ_CUDA_API int kvz_cuda_sao_reconstruct(const kvz_cuda_plane *rec_in, const kvz_cuda_plane *rec_out, int n_planes,
                                          int bitdepth, const kvz_cuda_sao_params *d_luma,
                                          const kvz_cuda_sao_params *d_chroma, void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: transforms and quantisation  (rows A13, A14)              */
/* ------------------------------------------------------------------ */

/* `count` contiguous n x n int16 blocks (n = 4, 8, 16, 32).  dst4 != 0 selects the 4x4 DST that
 * kvz_get_dct_func / kvz_get_idct_func (strategies-dct.c:59-113) pick for intra luma 4x4.
 * Replaces dct_func kvz_dct_NxN / kvz_idct_NxN / kvz_fast_{forward,inverse}_dst_4x4
 * (strategies-dct.h:32-52; generic: dct-generic.c:206-617). */
KVZ_CUDA_API int kvz_cuda_transform_blocks(int n, int dst4, int inverse, int bitdepth, const int16_t *d_in,
                                           int16_t *d_out, int count, void *stream);

/* the fields of encoder_state_t / encoder_control_t that kvz_quant and kvz_dequant read */
typedef struct kvz_cuda_quant_cfg {
  int32_t qp;               /* state->qp */
  int32_t bitdepth;         /* encoder->bitdepth */
  int32_t slice_is_intra;   /* state->frame->slicetype == KVZ_SLICE_I (rounding offset 171 vs 85) */
  int32_t signhide;         /* cfg.signhide_enable */
  int32_t cu_is_intra;      /* block_type == CU_INTRA: scaling-list class; unused by the flat list */
} kvz_cuda_quant_cfg;

/* quant_func kvz_quant (strategies-quant.h:36; quant-generic.c:37-163) for `count` n x n blocks;
 * type 0 luma / 2 chroma, scan_idx 0 diagonal / 1 horizontal / 2 vertical.  Flat scaling list only. */
KVZ_CUDA_API int kvz_cuda_quant_blocks(const kvz_cuda_quant_cfg *cfg, int n, int type, int scan_idx,
                                       const int16_t *d_coef, int16_t *d_q_coef, int count, void *stream);
/* dequant_func kvz_dequant (quant-generic.c:279-321) */
KVZ_CUDA_API int kvz_cuda_dequant_blocks(const kvz_cuda_quant_cfg *cfg, int n, int type,
                                         const int16_t *d_q_coef, int16_t *d_coef, int count, void *stream);
/* coeff_abs_sum_func (quant-generic.c:323-330): out[i] = sum |coeffs[i * per_block ..]| */
KVZ_CUDA_API int kvz_cuda_coeff_abs_sums(const int16_t *d_coeffs, int per_block, int count, uint32_t *d_out,
                                         void *stream);

/* one kvz_quantize_residual() call (strategies-quant.h:38-42; quant-generic.c:180-273) */
typedef struct kvz_cuda_residual_job {
  uint32_t ref_offset, pred_offset;   /* byte offsets of ref_in / pred_in in d_ref / d_pred (stride in_stride) */
  uint32_t rec_offset;                /* byte offset of rec_out in d_rec (stride out_stride) */
  uint32_t coeff_offset;              /* element offset of coeff_out in d_coeff (width*width contiguous) */
  int32_t in_stride, out_stride;
  uint8_t width;                      /* 4, 8, 16, 32 */
  uint8_t color;                      /* color_t: 0 Y, 1 U, 2 V */
  uint8_t scan_idx;                   /* coeff_scan_order_t */
  uint8_t use_trskip;
  int8_t qp;                          /* state->qp of the CTU */
  uint8_t cu_is_intra;                /* cur_cu->type == CU_INTRA (selects the 4x4 DST) */
  uint8_t reserved[2];
} kvz_cuda_residual_job;

/* Fused residual -> transform -> quant -> dequant -> inverse transform -> reconstruction for a batch
 * of transform blocks; cfg->qp and cfg->cu_is_intra are taken from each job.  d_has_coeffs[i] is the
 * function's return value.  RDOQ (rdo.c:548-885) is not applied: this is the cfg.rdoq_enable == 0 path. */
KVZ_CUDA_API int kvz_cuda_quantize_residuals(const kvz_cuda_quant_cfg *cfg, const void *d_ref, const void *d_pred,
                                             void *d_rec, int16_t *d_coeff, const kvz_cuda_residual_job *d_jobs,
                                             int n_jobs, uint8_t *d_has_coeffs, void *stream);

/* ---- A15: rate-distortion optimised quantisation ---- */

/* The CABAC context states (cabac_ctx_t.uc_state, cabac.h:36-39) that kvz_rdoq reads from
 * state->cabac.ctx (cabac.h:53-90).  The host snapshots them once per CTU / wavefront step; any
 * number of blocks may share a snapshot. */
typedef struct kvz_cuda_rdoq_ctx {
  uint8_t cbf_luma[4], cbf_chroma[4];            /* qt_cbf_model_luma / qt_cbf_model_chroma */
  uint8_t sig_group[4];                          /* cu_sig_coeff_group_model */
  uint8_t sig_luma[27], sig_chroma[15];          /* cu_sig_model_luma / _chroma */
  uint8_t last_y_luma[15], last_y_chroma[15];    /* cu_ctx_last_y_luma / _chroma */
  uint8_t last_x_luma[15], last_x_chroma[15];    /* cu_ctx_last_x_luma / _chroma */
  uint8_t one_luma[16], one_chroma[8];           /* cu_one_model_luma / _chroma */
  uint8_t abs_luma[4], abs_chroma[2];            /* cu_abs_model_luma / _chroma */
  uint8_t root_cbf;                              /* cu_qt_root_cbf_model */
  uint8_t trskip_luma, trskip_chroma;               /* transform_skip_model_luma / _chroma (kvz_cuda_coeff_bits only) */
  uint8_t reserved;
} kvz_cuda_rdoq_ctx;                             /* 148 bytes */

/* one kvz_rdoq() call (rdo.c:548-551) */
typedef struct kvz_cuda_rdoq_job {
  uint32_t coef_offset;               /* element offset of the width*width block in d_coef and in d_q_coef */
  uint32_t ctx_index;                 /* entry of d_ctx holding state->cabac.ctx for this block */
  double lambda;                      /* state->lambda */
  uint8_t width;                      /* 4, 8, 16, 32 (height == width) */
  uint8_t type;                       /* 0 luma, 2 chroma */
  uint8_t scan_idx;                   /* coeff_scan_order_t; 1 and 2 only for width <= 8 (tables.h:72-77) */
  uint8_t tr_depth;
  int8_t qp;                          /* state->qp */
  uint8_t cu_is_intra;                /* block_type == CU_INTRA */
  uint8_t reserved[2];
} kvz_cuda_rdoq_job;                  /* 24 bytes */

/* kvz_rdoq (rdo.c:548-885) including kvz_rdoq_sign_hiding (rdo.c:405-540) for a batch of transform
 * blocks: d_q_coef receives dest_coeff.  cfg supplies bitdepth and signhide; qp and cu_is_intra come
 * from each job.  Flat scaling list.  All floating-point work is IEEE double in the reference's
 * operation order (no FMA contraction), so the chosen levels are bit-identical. */
KVZ_CUDA_API int kvz_cuda_rdoq_blocks(const kvz_cuda_quant_cfg *cfg, const kvz_cuda_rdoq_ctx *d_ctx,
                                      const int16_t *d_coef, int16_t *d_q_coef,
                                      const kvz_cuda_rdoq_job *d_jobs, int n_jobs, int max_width, void *stream);

typedef struct kvz_cuda_coeff_job {
  uint32_t coef_offset;               /* first coefficient of the block in d_coeffs, in int16 units */
  uint32_t ctx_index;                 /* which snapshot of d_ctx */
  uint16_t range;                     /* state->cabac.range when the cost is asked for (256..510) */
  uint8_t width;                      /* 4, 8, 16, 32 */
  uint8_t type;                       /* 0 luma, 2 chroma */
  uint8_t scan_idx;                   /* coeff_scan_order_t */
  uint8_t reserved[3];
} kvz_cuda_coeff_job;                 /* 16 bytes */

/* kvz_get_coeff_cost (rdo.c:208-222) on its CABAC path (state->qp >= cfg.fast_residual_cost_limit):
 * d_bits[i] = the bits kvz_encode_coeff_nxn (encode_coding_tree.c:103-345) spends on block i when run in counting
 * mode on a copy of the CABAC state, i.e. get_coeff_cabac_cost (rdo.c:158-197); 0 for an all-zero block.  The
 * count depends on the coder's range and the context states only (cabac.c:128-137 counts bytes and never looks at
 * `low`), both supplied per job; the snapshots are not modified.  signhide = cfg.signhide_enable,
 * trskip_enable = cfg.trskip_enable (the transform_skip_flag of 4x4 blocks, always coded as 0 here as in
 * rdo.c:188-194).  The fast estimate for low QPs, sum * (qp * COEFF_COST_QP_FACTOR + COEFF_COST_BIAS), is
 * kvz_cuda_coeff_abs_sums plus one multiplication on the host. */
KVZ_CUDA_API int kvz_cuda_coeff_bits(const kvz_cuda_rdoq_ctx *d_ctx, const int16_t *d_coeffs,
                                     const kvz_cuda_coeff_job *d_jobs, int n_jobs, int max_width, int signhide,
                                     int trskip_enable, uint32_t *d_bits, void *stream);

/* kvz_quantize_residual (quant-generic.c:180-273) with cfg.rdoq_enable: residual -> transform -> kvz_rdoq ->
 * dequant -> inverse transform -> reconstruction.  d_rdoq_jobs[i] describes the kvz_rdoq call of d_jobs[i]
 * (same width; coef_offset == coeff_offset; type 0 for colour Y, 2 otherwise; tr_depth as computed at
 * quant-generic.c:216-217).  d_work: scratch for the transform coefficients, as large as d_coeff.  With
 * rdoq_skip != 0 the 4x4 blocks use the scalar quantiser (cfg.rdoq_skip, quant-generic.c:213-215). */
KVZ_CUDA_API int kvz_cuda_quantize_residuals_rdoq(const kvz_cuda_quant_cfg *cfg, const void *d_ref, const void *d_pred,
                                                  void *d_rec, int16_t *d_coeff, const kvz_cuda_residual_job *d_jobs,
                                                  int n_jobs, uint8_t *d_has_coeffs, const kvz_cuda_rdoq_ctx *d_ctx,
                                                  const kvz_cuda_rdoq_job *d_rdoq_jobs, int max_width, int16_t *d_work,
                                                  int rdoq_skip, void *stream);

/* The same without any allocation and without a caller-supplied d_work.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  The transform-coefficient scratch is a file-scope device array of
 * KVZ_CUDA_NOALLOC_MAX_COEFFS int16 (2 MiB, held by every GPU that loads the library until exit); `coeff_elems` is the
 * number of int16 elements of d_coeff the jobs address (max over jobs of coeff_offset + width * width) and the call
 * fails with a clear kvz_cuda_last_error() when it exceeds that maximum.  The call makes no cudaMalloc / cudaFree and
 * uses no allocating library routine; inputs and outputs are still the caller's.  Every call shares the scratch, which
 * keeps the previous call's contents (nothing in it needs to start at zero: the first stage writes every coefficient
 * the RDOQ stage reads), so calls that OVERLAP ON THE GPU -- two streams, two host threads -- race; order them. */
#define KVZ_CUDA_NOALLOC_MAX_COEFFS (1u << 20)
KVZ_CUDA_API int kvz_cuda_quantize_residuals_rdoq_noalloc(const kvz_cuda_quant_cfg *cfg, const void *d_ref,
                                                          const void *d_pred, void *d_rec, int16_t *d_coeff,
                                                          size_t coeff_elems, const kvz_cuda_residual_job *d_jobs,
                                                          int n_jobs, uint8_t *d_has_coeffs, const kvz_cuda_rdoq_ctx *d_ctx,
                                                          const kvz_cuda_rdoq_job *d_rdoq_jobs, int max_width,
                                                          int rdoq_skip, void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: fractional-sample interpolation  (rows A9, A10, A11, A12) */
/* ------------------------------------------------------------------ */

enum kvz_cuda_interp_mode {
  KVZ_CUDA_INTERP_PIXEL = 0,    /* kvz_sample_quarterpel_luma / kvz_sample_octpel_chroma: clipped pixels */
  KVZ_CUDA_INTERP_14BIT = 1,    /* kvz_sample_14bit_*: int16 intermediate precision for bi-prediction */
  KVZ_CUDA_INTERP_FME = 2       /* one candidate block of kvz_filter_{hpel,qpel}_blocks_*_luma (int16 wrap kept) */
};

typedef struct kvz_cuda_interp_job {
  int32_t x, y;                 /* top-left whole sample of the block in the plane (xpos + tile offset) */
  int32_t mv_x, mv_y;           /* displacement: quarter-pel (luma) or eighth-pel (chroma), any integer */
  int16_t w, h;                 /* <= 64 (luma), <= 32 (chroma) */
  uint8_t chroma;               /* 0: 8-tap luma filter, 1: 4-tap chroma filter */
  uint8_t mode;                 /* enum kvz_cuda_interp_mode */
  uint8_t reserved[2];
  uint32_t dst_offset;          /* element offset in d_dst8 (modes 0, 2) or d_dst16 (mode 1) */
  int32_t dst_stride;
} kvz_cuda_interp_job;

/* Interpolates a batch of blocks of one UNPADDED plane (width x height, pitch bytes per row); samples outside
 * the plane are the clamped ones kvz_get_extended_block (ipol-generic.c:731-784) would have copied.
 * Replaces kvz_sample_quarterpel_luma, kvz_sample_octpel_chroma and their 14-bit variants
 * (strategies-ipol.h:44-48), and, four jobs per call, the ipol_blocks_func filters (strategies-ipol.h:36-38). */
KVZ_CUDA_API int kvz_cuda_interp_blocks(const void *d_plane, int pitch, int width, int height, int bitdepth,
                                        const kvz_cuda_interp_job *d_jobs, int n_jobs,
                                        void *d_dst8, int16_t *d_dst16, void *stream);

/* inter_recon_bipred_func on flat arrays of 14-bit samples: dst = clip((s0 + s1 + offset) >> (15 - bitdepth))
 * (picture-generic.c:553-570) */
KVZ_CUDA_API int kvz_cuda_bipred_blend(const int16_t *d_s0, const int16_t *d_s1, size_t n, int bitdepth,
                                       void *d_dst, void *stream);

/* ------------------------------------------------------------------ */
/* BATCH face: intra prediction and rough mode costs  (row A17)          */
/* ------------------------------------------------------------------ */

typedef struct kvz_cuda_intra_job {
  uint32_t ref_offset;          /* offset in d_refs (in SAMPLES) of this PU's reference samples: top[2w+1] then left[2w+1],
                                   [0] = the top-left corner (kvz_intra_references.ref, intra.h) */
  uint32_t orig_offset;         /* offset (samples) of the source block in d_orig; orig_stride in samples */
  int32_t orig_stride;
  uint32_t pred_offset;         /* offset (samples) in d_preds of this PU's 35 predictions (35 * w * w samples) */
  uint8_t log2_width;           /* 2..5 */
  uint8_t is_chroma;            /* color != COLOR_Y */
  uint8_t filter_boundary;      /* !(cfg->lossless && cfg->implicit_rdpcm), search_intra.c:420 */
  uint8_t has_orig;             /* 0: predictions only */
} kvz_cuda_intra_job;

/* For each PU: kvz_intra_predict (intra.c:281-332; angular_pred_func / intra_pred_planar_func of
 * strategies-intra.h:32-43 inside) for all 35 modes, and d_costs[(pu * 35 + mode) * 2 + {0, 1}] =
 * {satd_NxN, sad_NxN}(prediction, source) -- every cost search_intra_rough (search_intra.c:404-545) can ask for.
 * d_preds may be NULL; d_costs / d_orig may be NULL when no job has has_orig. */
KVZ_CUDA_API int kvz_cuda_intra_mode_costs(const void *d_refs, const void *d_orig, const kvz_cuda_intra_job *d_jobs,
                                           int n_jobs, int bitdepth, uint32_t *d_costs, void *d_preds, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* KVZ_CUDA_H_ */
