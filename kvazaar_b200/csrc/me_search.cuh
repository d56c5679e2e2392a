// kvazaar_b200/csrc/me_search.cuh -- launch interface of the per-PU motion decision kernel
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/kvz_cuda.h"

namespace kvz {

constexpr int kMePuArgsMax = 8;   // PUs one launch_me_search_pu can carry as kernel arguments
constexpr int kMeMaxBox = 272;     // largest box side (displacements) kvz_cuda_me_search reads

struct MeMapSet {
  int cx, cy;                // window centre the maps were made for; first displacement (x0, y0) of a box
  int n_dx, n_dy;            // box size, 0 = square window of the frame's range
  const uint16_t *sad8, *sad16;
  const uint32_t *sad32, *sad64;
};

struct MeRef {
  const uint8_t *plane;      // border-replicated reference luma
  int pitch, pad;            // bytes per row, border width
  int n_sets;
  MeMapSet sets[KVZ_CUDA_MAX_MAPSETS];   // dense maps of this (current, reference) pair
};

struct MeFrame {
  const uint8_t *cur;
  int cur_pitch;
  int width, height;
  int tile_x, tile_y, tile_w, tile_h;
  int ctus_x;
  int range, map_pitch;
  int ime_search, me_early_termination, me_max_steps;   // kvz_cuda_ime_search; cfg.me_early_termination / me_max_steps
  int lambda_sqrt_i;
  int fme_level, owf, wpp, sao_type, deblock_enable, mv_constraint, max_ref_lcu_right, max_ref_lcu_down;
  int n_refs;
  uint8_t ref_LX[2][16];
  MeRef refs[KVZ_CUDA_MAX_REFS];
};

int launch_me_search(const MeFrame &f, const kvz_cuda_me_pu *pus, const kvz_cuda_me_job *jobs,
                     kvz_cuda_me_result *results, int n_pus, cudaStream_t stream);

// the same, every PU by one CTA in one launch (integer search with any window evaluated from shared memory when no
// box holds it, then the refinement): the form for a few PUs at a time.  host_pu / host_jobs (n_pus <= kMePuArgsMax,
// KVZ_CUDA_MAX_REFS job records per PU): sent as kernel arguments; PU i's results go to results[i * KVZ_CUDA_MAX_REFS ..]
// and done_flag[i] (mapped host memory) receives done_seq after them, for a caller that polls instead of synchronising
int launch_me_search_pu(const MeFrame &f, const kvz_cuda_me_pu *pus, const kvz_cuda_me_job *jobs,
                        kvz_cuda_me_result *results, int n_pus, cudaStream_t stream, const kvz_cuda_me_pu *host_pu = nullptr,
                        const kvz_cuda_me_job *host_jobs = nullptr, uint32_t *done_flag = nullptr, uint32_t done_seq = 0);

int me_pu_phase_times(unsigned long long out[4][4]);

}  // namespace kvz
