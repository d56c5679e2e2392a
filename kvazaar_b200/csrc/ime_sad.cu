// kvazaar_b200/csrc/ime_sad.cu
//
// Dense full-search SAD for one (current picture, reference picture) pair:
// for every 8x8 block of every CTU and every integer displacement (dx,dy) in
// [-R,R]^2 the kernel produces the SAD of the block against the edge-clamped
// reference, then sums 2x2 neighbours exactly into the 16x16, 32x32 and 64x64
// PU costs of the same displacement.
//
// Replaces, as one frame-level batch, the per-candidate call chain
//   check_mv_cost (search_inter.c:195-232)
//     -> kvz_image_calc_sad (image.c:455-479, clamp semantics image.c:320-452)
//       -> kvz_reg_sad (strategies/generic/picture-generic.c:86-99)
// for the zero-centred window of search_mv_full (search_inter.c:891-895).
//
// Design (sm_100a):
//  * the reference plane lives in HBM with replicated borders (pad_plane
//    kernel) so a tile fetch never needs per-pixel clamping: clamping the
//    coordinate == reading the replicated border;
//  * one CTA = one CTU x one chunk of 16 horizontal displacements.  The
//    (2R+64) x 80-byte reference tile and the 64x64 current CTU are staged in
//    shared memory once and reused for all 2R+1 vertical displacements;
//  * one thread = one 8x8 block x 4 consecutive dx.  It slides down the
//    reference rows; each row is loaded once (3 LDS), realigned to the 4 byte
//    shifts (6 SHF) and consumed by 8 current rows x 4 shifts x 2 words
//    = 64 VABSDIFF4.U8.ACC, i.e. 256 pixel differences per 9 non-SAD
//    instructions.  Eight (dy) x four (dx) accumulators rotate through a
//    statically indexed register window (loop unrolled by 8);
//  * every completed dy row is written as full 32-byte sectors (4 lanes x 8 B)
//    and staged in shared memory, where 4 warps (alternating by parity so the
//    load is even) form the 16/32/64 sums with packed adds and shuffles.
#include "common.cuh"
#include "ime_sad.cuh"

namespace kvz {

namespace {

constexpr int kThreads = 256;
constexpr int kTilePitch = 80;        // bytes per staged reference row (20 words)
constexpr int kChunkDx = 16;          // horizontal displacements per CTA

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t d;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

template <int R>
struct Smem {
  static constexpr int kRows = 2 * R + 64;
  static constexpr int kK = 2 * R + 1;
  alignas(16) uint8_t tile[kRows * kTilePitch];
  alignas(16) uint8_t cur[64 * 64];
  alignas(16) uint16_t stage[2][64][kChunkDx];   // [parity][block][dx in chunk]
  uint32_t sad64[kK][kChunkDx];
};

template <int R>
__global__ void __launch_bounds__(kThreads, 3)
ime_sad_dense_kernel(ImeSadParams p)
{
  constexpr int K = 2 * R + 1;
  constexpr int kGroupsValid = R / 2 + 1;            // groups of 4 dx; last holds dx == +R only
  __shared__ Smem<R> sm;

  const int chunk = blockIdx.x;
  const int ctu = blockIdx.y;
  const int ctu_cx = ctu % p.ctus_x, ctu_cy = ctu / p.ctus_x;
  const int ctu_x = p.tile_x + ctu_cx * 64;          // frame-absolute luma position
  const int ctu_y = p.tile_y + ctu_cy * 64;
  const int tid = threadIdx.x;

  // ---- stage the reference tile: rows ctu_y-R .. ctu_y+R+63, 80 bytes from
  //      column ctu_x - R + 16*chunk of the border-replicated plane ----
  {
    const uint8_t *src = p.ref + (size_t)(ctu_y + p.center_y - R + p.pad) * p.ref_pitch + (ctu_x + p.center_x - R + kChunkDx * chunk + p.pad);
    constexpr int kVecPerRow = kTilePitch / 8;
    // the window centre makes the source address arbitrary: fetch aligned words and realign
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(src);
    const int sh = (int)(a0 & 3) * 8;
    const uint8_t *src_al = reinterpret_cast<const uint8_t *>(a0 & ~(uintptr_t)3);
    for (int i = tid; i < Smem<R>::kRows * kVecPerRow; i += kThreads) {
      const int r = i / kVecPerRow, v = i % kVecPerRow;
      const uint32_t *w = reinterpret_cast<const uint32_t *>(src_al + (size_t)r * p.ref_pitch + v * 8);
      const uint32_t w0 = w[0], w1 = w[1], w2 = sh ? w[2] : 0u;
      *reinterpret_cast<uint2 *>(&sm.tile[r * kTilePitch + v * 8]) =
          make_uint2(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh));
    }
    // current CTU, coordinates clamped into the picture for partial CTUs
    for (int i = tid; i < 64 * 8; i += kThreads) {
      const int r = i >> 3, v = i & 7;
      const int y = min(ctu_y + r, p.height - 1);
      const int x = min(ctu_x + v * 8, p.width - 8);
      const uint2 val = *reinterpret_cast<const uint2 *>(p.cur + (size_t)y * p.cur_pitch + x);
      *reinterpret_cast<uint2 *>(&sm.cur[r * 64 + v * 8]) = val;
    }
    for (int i = tid; i < K * kChunkDx; i += kThreads) (&sm.sad64[0][0])[i] = 0;
    for (int i = tid; i < 2 * 64 * kChunkDx / 2; i += kThreads) reinterpret_cast<uint32_t *>(&sm.stage[0][0][0])[i] = 0;
  }
  __syncthreads();

  // ---- thread role in the compute phase ----
  const int warp = tid >> 5, lane = tid & 31;
  const int by = warp;                 // 8 warps <-> 8 block rows of the CTU
  const int bx = lane >> 2;            // 8 blocks per row
  const int gl = lane & 3;             // dx group inside the chunk
  const int g = chunk * 4 + gl;
  const int blk = by * 8 + bx;
  const bool blk_valid = (ctu_x + bx * 8 + 8 <= p.tile_x + p.tile_w) && (ctu_y + by * 8 + 8 <= p.tile_y + p.tile_h);
  const bool active = blk_valid && g < kGroupsValid;

  uint32_t cur[8][2];
#pragma unroll
  for (int y = 0; y < 8; y++) {
    const uint2 v = *reinterpret_cast<const uint2 *>(&sm.cur[(by * 8 + y) * 64 + bx * 8]);
    cur[y][0] = v.x; cur[y][1] = v.y;
  }
  uint32_t acc[8][4];
#pragma unroll
  for (int s = 0; s < 8; s++)
#pragma unroll
    for (int f = 0; f < 4; f++) acc[s][f] = 0;

  const uint8_t *rowp = &sm.tile[(by * 8) * kTilePitch + bx * 8 + gl * 4];
  // global destinations
  const size_t pu8 = (size_t)ctu * 64 + blk;
  uint16_t *out8 = p.sad8 + pu8 * K * p.map_pitch + g * 4;

  // reduction role (threads of 4 warps; which 4 alternates with k parity)
  const int rwarp = warp & 3;          // p32 index: (p32x, p32y)
  const int qx = lane >> 4, dxl = lane & 15;
  const int p16x = 2 * (rwarp & 1) + qx;
  const int p16y0 = 2 * (rwarp >> 1);
  uint16_t *out16a = p.sad16 + ((size_t)ctu * 16 + (p16y0 + 0) * 4 + p16x) * K * p.map_pitch + chunk * kChunkDx + dxl;
  uint16_t *out16b = p.sad16 + ((size_t)ctu * 16 + (p16y0 + 1) * 4 + p16x) * K * p.map_pitch + chunk * kChunkDx + dxl;
  uint32_t *out32 = p.sad32 + ((size_t)ctu * 4 + rwarp) * K * p.map_pitch + chunk * kChunkDx + dxl;

  constexpr int kIters = (2 * R + 8) / 8;
  for (int m = 0; m < kIters; m++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const int s = m * 8 + u;
      if (active) {
        const uint8_t *rp = rowp + s * kTilePitch;
        const uint32_t w0 = *reinterpret_cast<const uint32_t *>(rp);
        const uint32_t w1 = *reinterpret_cast<const uint32_t *>(rp + 4);
        const uint32_t w2 = *reinterpret_cast<const uint32_t *>(rp + 8);
        uint32_t lo[4], hi[4];
        lo[0] = w0; hi[0] = w1;
#pragma unroll
        for (int f = 1; f < 4; f++) {
          lo[f] = __funnelshift_r(w0, w1, 8 * f);
          hi[f] = __funnelshift_r(w1, w2, 8 * f);
        }
#pragma unroll
        for (int y = 0; y < 8; y++) {
          const int slot = (u - y) & 7;          // accumulator of k = s - y
#pragma unroll
          for (int f = 0; f < 4; f++) {
            acc[slot][f] = sad4(cur[y][0], lo[f], acc[slot][f]);
            acc[slot][f] = sad4(cur[y][1], hi[f], acc[slot][f]);
          }
        }
      }
      // row k = s - 7 is complete
      const int k = s - 7;
      const int slot = (u + 1) & 7;
      if (k >= 0 && k < K) {               // uniform across the CTA
        const uint32_t v0 = acc[slot][0] | (acc[slot][1] << 16);
        const uint32_t v1 = acc[slot][2] | (acc[slot][3] << 16);
        const int par = k & 1;
        if (active) {
          *reinterpret_cast<uint2 *>(out8 + (size_t)k * p.map_pitch) = make_uint2(v0, v1);
          *reinterpret_cast<uint2 *>(&sm.stage[par][blk][gl * 4]) = make_uint2(v0, v1);
        }
        __syncthreads();
        if ((warp >> 2) == par) {
          // SAD16 of two vertically adjacent 16x16 PUs, then 32x32, then 64x64
          const uint16_t (*st)[kChunkDx] = sm.stage[par];
          const int b0 = (2 * p16y0) * 8 + 2 * p16x;
          const uint32_t a = (uint32_t)st[b0][dxl] + st[b0 + 1][dxl] + st[b0 + 8][dxl] + st[b0 + 9][dxl];
          const uint32_t b = (uint32_t)st[b0 + 16][dxl] + st[b0 + 17][dxl] + st[b0 + 24][dxl] + st[b0 + 25][dxl];
          out16a[(size_t)k * p.map_pitch] = (uint16_t)a;
          out16b[(size_t)k * p.map_pitch] = (uint16_t)b;
          uint32_t s32 = a + b;
          s32 += __shfl_xor_sync(0xffffffffu, s32, 16);
          if (lane < 16) {
            out32[(size_t)k * p.map_pitch] = s32;
            atomicAdd(&sm.sad64[k][dxl], s32);
          }
        }
      }
#pragma unroll
      for (int f = 0; f < 4; f++) acc[slot][f] = 0;
    }
  }
  __syncthreads();
  // flush the 64x64 sums
  uint32_t *out64 = p.sad64 + (size_t)ctu * K * p.map_pitch + chunk * kChunkDx;
  for (int i = tid; i < K * kChunkDx; i += kThreads) {
    const int k = i / kChunkDx, d = i % kChunkDx;
    out64[(size_t)k * p.map_pitch + d] = sm.sad64[k][d];
  }
}

// ---- replicate-pad a plane (image.c:320-452 == clamp-to-edge) ----
template <typename PIX>
__global__ void pad_plane_kernel(const PIX *src, int w, int h, int src_pitch, PIX *dst, int pad, int dst_pitch)
{
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y;
  const int pw = w + 2 * pad;
  if (x >= pw) return;
  const int sx = clampi(x - pad, 0, w - 1), sy = clampi(y - pad, 0, h - 1);
  dst[(size_t)y * dst_pitch + x] = src[(size_t)sy * src_pitch + sx];
}

}  // namespace

int ime_map_pitch(int range) { return ((2 * range + 1) + 15) & ~15; }

int launch_pad_plane(const void *src, int w, int h, int src_pitch, void *dst, int pad, int dst_pitch,
                     int bitdepth, cudaStream_t stream)
{
  const int pw = w + 2 * pad, ph = h + 2 * pad;
  dim3 grid((pw + 255) / 256, ph);
  if (bitdepth == 8)
    pad_plane_kernel<uint8_t><<<grid, 256, 0, stream>>>((const uint8_t *)src, w, h, src_pitch, (uint8_t *)dst, pad, dst_pitch);
  else
    pad_plane_kernel<uint16_t><<<grid, 256, 0, stream>>>((const uint16_t *)src, w, h, src_pitch, (uint16_t *)dst, pad, dst_pitch);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_ime_sad_dense(const ImeSadParams &p, int range, cudaStream_t stream)
{
  const int groups = range / 2 + 1;
  const int chunks = (groups + 3) / 4;
  dim3 grid(chunks, p.ctus_x * p.ctus_y);
  switch (range) {
    case 8:  ime_sad_dense_kernel<8><<<grid, kThreads, 0, stream>>>(p); break;
    case 16: ime_sad_dense_kernel<16><<<grid, kThreads, 0, stream>>>(p); break;
    case 32: ime_sad_dense_kernel<32><<<grid, kThreads, 0, stream>>>(p); break;
    case 64: ime_sad_dense_kernel<64><<<grid, kThreads, 0, stream>>>(p); break;
    default: return fail_msg("ime_sad_dense: range must be 8, 16, 32 or 64");
  }
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
