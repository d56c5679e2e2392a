// kvazaar_b200/csrc/ime_sad.cuh -- launch interface of the dense SAD kernel
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace kvz {

struct ImeSadParams {
  const uint8_t *cur;      // current picture luma (device), un-padded
  const uint8_t *ref;      // reference luma with `pad` replicated border pixels on every side
  int width, height;       // picture size (luma)
  int cur_pitch, ref_pitch;// bytes per row
  int pad;                 // border of the padded reference plane
  int tile_x, tile_y, tile_w, tile_h;   // tile rectangle searched (whole frame: 0,0,width,height)
  int ctus_x, ctus_y;      // CTUs in the tile
  int x0, y0;              // first displacement: map[i][j] is the SAD at displacement (x0 + j, y0 + i)
  int n_dy;                // displacement rows computed, == 1 (mod 8)
  int skip;                // the first `skip` (0..7) computed rows are alignment only: map row i = computed row i + skip
  int n_chunks;            // chunks of 16 displacement columns computed by the chunk CTAs
  int n_dx_last;           // columns of the last chunk: 16 or 8
  int pu_rows;             // map rows per PU: n_dy, or n_dy + 1 when row n_dy holds the column dx = x0 + 16*n_chunks
  int map_pitch;           // entries per map row
  int chunk_major;         // box maps: map[chunk][pu][row][16 or 8] (see ime_sad.cu); 0: map[pu][row][map_pitch]
  uint16_t *sad8;          // [ctu*64 + by*8+bx][pu_rows][map_pitch]
  uint16_t *sad16;         // [ctu*16 + py*4+px][pu_rows][map_pitch]
  uint32_t *sad32;         // [ctu*4  + py*2+px][pu_rows][map_pitch]
  uint32_t *sad64;         // [ctu][pu_rows][map_pitch]
};

constexpr int kImeMaxMulti = 8;      // references per multi-reference launch

int ime_map_pitch(int range);
int launch_pad_plane(const void *src, int w, int h, int src_pitch, void *dst, int pad, int dst_pitch,
                     int bitdepth, cudaStream_t stream);
// square window [-range, range]^2 round (center_x, center_y): K = 2*range+1 rows of K-1 columns, the last
// column as row K; fills x0, y0, n_dy, n_chunks, pu_rows of `p`
int launch_ime_sad_dense(ImeSadParams p, int range, int center_x, int center_y, cudaStream_t stream);
// rectangular box: n_dx (multiple of 8) x n_dy displacements from (x0, y0), plain row-major maps
int launch_ime_sad_box(ImeSadParams p, int x0, int y0, int n_dx, int n_dy, cudaStream_t stream);

// 10-bit samples (uint16_t planes, pitches of `p` in bytes): ime_sad10.cu
int launch_ime_sad10_box(ImeSadParams p, int x0, int y0, int n_dx, int n_dy, int bitdepth, cudaStream_t stream);
// largest box height (displacement rows) of launch_ime_sad_box_multi
int ime_box_max_rows();
// the boxes of n <= kImeMaxMulti references of one current picture in ONE launch; boxes[r] = {x0, y0, n_dx, n_dy}
// ctu_only >= 0: only the maps of that CTU (what a host computes when the reference is still being reconstructed)
int launch_ime_sad_box_multi(const ImeSadParams *ps, const int (*boxes)[4], int n, cudaStream_t stream, int ctu_only = -1);
// after the frame rectangle (x0, y0, w, h) of a bordered plane's interior changed: re-replicate the border it feeds
int launch_pad_patch(void *plane, int W, int H, int pad, int pitch, int x0, int y0, int w, int h, int bitdepth,
                     cudaStream_t stream);

}  // namespace kvz
