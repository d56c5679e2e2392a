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
  int center_x, center_y;  // window centre: map[dy+R][dx+R] is the SAD at displacement (center_x+dx, center_y+dy)
  int map_pitch;           // entries per (pu, dy) row of every map: roundup(2R+1, 16)
  uint16_t *sad8;          // [ctu*64 + by*8+bx][2R+1][map_pitch]
  uint16_t *sad16;         // [ctu*16 + py*4+px][2R+1][map_pitch]
  uint32_t *sad32;         // [ctu*4  + py*2+px][2R+1][map_pitch]
  uint32_t *sad64;         // [ctu][2R+1][map_pitch]
};

int ime_map_pitch(int range);
int launch_pad_plane(const void *src, int w, int h, int src_pitch, void *dst, int pad, int dst_pitch,
                     int bitdepth, cudaStream_t stream);
int launch_ime_sad_dense(const ImeSadParams &p, int range, cudaStream_t stream);

}  // namespace kvz
