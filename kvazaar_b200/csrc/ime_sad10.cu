// kvazaar_b200/csrc/ime_sad10.cu
//
// Dense box SAD maps for the 10-bit build of the reference (kvz_pixel = uint16_t, KVZ_BIT_DEPTH 10,
// kvazaar.h:72-77): the same (PU, displacement) costs as ime_sad.cu, i.e. the call chain
//   check_mv_cost (search_inter.c:195-232) -> kvz_image_calc_sad (image.c:455-479, clamp image.c:320-452,
//   result >> (KVZ_BIT_DEPTH - 8)) -> kvz_reg_sad (picture-generic.c:86-99)
// for every 8x8 block of every CTU, summed exactly into the 16x16 / 32x32 / 64x64 PUs, written in the
// chunk-major box layout of include/kvz_cuda.h so that kvz_cuda_me_search's integer scan reads them unchanged.
//
// There is no VABSDIFF for 16-bit lanes on sm_100a (vabsdiff2 is emulated with ~10 instructions); two samples
// per word cost three integer-pipe instructions here: VIMNMX.U16x2 (max), VIMNMX.U16x2 (min) and one IADD3
// (acc + max - min; a 16-bit lane holds at most 32 * 1023, no carry between the lanes).  That is 1.5 instructions
// per sample against 0.25 for 8-bit VABSDIFF4 -- the 10-bit roofline is 6x lower by construction.
//
// First version (parity first): one CTA = one CTU x 16 displacement columns, thread = one 8x8 block x 4 dx;
// every displacement row recomputes its eight reference rows from shared memory (no sliding accumulator window
// yet), the reference tile is staged twice (sample offsets 0 and 1) so that every operand is an aligned word.
#include "common.cuh"
#include "ime_sad.cuh"

namespace kvz {
namespace {

constexpr int kT = 256;
constexpr int kChunk = 16;
constexpr int kTileWords = 41;        // words per tile row and copy: 16 + 63 + 1 samples, odd pitch against bank conflicts

struct Head10 {
  uint32_t cur[64 * 32];              // the current CTU, two samples per word
  uint32_t s8[64 * kChunk];           // raw 8x8 sums of one displacement row
};

__host__ __device__ inline size_t smem10_bytes(int rows) { return sizeof(Head10) + (size_t)2 * rows * kTileWords * 4; }

__device__ __forceinline__ uint32_t sad2(uint32_t a, uint32_t b, uint32_t acc)
{
  return acc + __vmaxu2(a, b) - __vminu2(a, b);       // VIMNMX.U16x2 x2 + IADD3
}

__global__ void __launch_bounds__(kT)
ime_sad10_box_kernel(ImeSadParams p, int shift)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Head10 &H = *reinterpret_cast<Head10 *>(smem_raw);
  const int tid = threadIdx.x;
  const int n_dy = p.pu_rows, rows = n_dy + 63;
  uint32_t *tile0 = reinterpret_cast<uint32_t *>(smem_raw + sizeof(Head10));
  uint32_t *tile1 = tile0 + rows * kTileWords;
  const int chunk = blockIdx.x % p.n_chunks, ctu = blockIdx.x / p.n_chunks;
  const int nd = chunk == p.n_chunks - 1 ? p.n_dx_last : kChunk;
  const int ctu_x = p.tile_x + (ctu % p.ctus_x) * 64, ctu_y = p.tile_y + (ctu / p.ctus_x) * 64;
  const uint16_t *cur = reinterpret_cast<const uint16_t *>(p.cur);
  const uint16_t *ref = reinterpret_cast<const uint16_t *>(p.ref);
  const int cur_pitch = p.cur_pitch >> 1, ref_pitch = p.ref_pitch >> 1;       // samples

  // ---- stage the current CTU (coordinates clamped into the picture for partial CTUs) and the tile ----
  for (int i = tid; i < 64 * 32; i += kT) {
    const int r = i >> 5, w = i & 31;
    const int y = min(ctu_y + r, p.height - 1), x = min(ctu_x + 2 * w, p.width - 2);
    H.cur[i] = *reinterpret_cast<const uint32_t *>(cur + (size_t)y * cur_pitch + x);
  }
  {
    const uint16_t *src = ref + (size_t)(ctu_y + p.y0 + p.pad) * ref_pitch + (ctu_x + p.x0 + kChunk * chunk + p.pad);
    for (int i = tid; i < rows * 40; i += kT) {
      const int r = i / 40, k = i - r * 40;
      const uint16_t *s = src + (size_t)r * ref_pitch + 2 * k;
      const uint32_t a = s[0], b = s[1], c = s[2];
      tile0[r * kTileWords + k] = a | (b << 16);
      tile1[r * kTileWords + k] = b | (c << 16);
    }
  }
  __syncthreads();

  const int blk = tid >> 2, g = tid & 3, by = blk >> 3, bx = blk & 7;
  uint32_t c[8][4];
#pragma unroll
  for (int y = 0; y < 8; y++)
#pragma unroll
    for (int i = 0; i < 4; i++) c[y][i] = H.cur[(by * 8 + y) * 32 + bx * 4 + i];
  const int base = bx * 4 + 2 * g;                      // first tile word of this thread's four displacements
  const uint32_t PR = (uint32_t)n_dy, rs = (uint32_t)nd, nctu = (uint32_t)(p.ctus_x * p.ctus_y);
  const uint32_t before = (uint32_t)chunk * kChunk * PR;

  for (int dy = 0; dy < n_dy; dy++) {
    uint32_t acc[4] = { 0, 0, 0, 0 };
    if (4 * g < nd) {
#pragma unroll
      for (int y = 0; y < 8; y++) {
        const uint32_t *e = tile0 + (by * 8 + y + dy) * kTileWords + base;
        const uint32_t *o = tile1 + (by * 8 + y + dy) * kTileWords + base;
        uint32_t ev[5], ov[5];
#pragma unroll
        for (int i = 0; i < 5; i++) { ev[i] = e[i]; ov[i] = o[i]; }
#pragma unroll
        for (int i = 0; i < 4; i++) {
          acc[0] = sad2(c[y][i], ev[i], acc[0]);        // dx = 4g
          acc[1] = sad2(c[y][i], ov[i], acc[1]);        // dx = 4g + 1
          acc[2] = sad2(c[y][i], ev[i + 1], acc[2]);    // dx = 4g + 2
          acc[3] = sad2(c[y][i], ov[i + 1], acc[3]);    // dx = 4g + 3
        }
      }
      uint32_t sad[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        sad[j] = (acc[j] & 0xFFFFu) + (acc[j] >> 16);
        H.s8[blk * kChunk + 4 * g + j] = sad[j];
      }
      // 8x8 level: four entries = 8 bytes
      uint16_t *o8 = p.sad8 + (size_t)before * (nctu * 64) + ((size_t)(ctu * 64 + blk) * PR + dy) * rs + 4 * g;
      *reinterpret_cast<uint2 *>(o8) = make_uint2((sad[0] >> shift) | ((sad[1] >> shift) << 16),
                                                  (sad[2] >> shift) | ((sad[3] >> shift) << 16));
    }
    __syncthreads();
    {
      // 16x16: item = (PU 0..15, dx); 32x32: (PU 0..3, dx); 64x64: (dx)
      const int dx = tid & 15, pu = tid >> 4;
      if (dx < nd) {
        const int b0 = (pu >> 2) * 16 + (pu & 3) * 2;
        const uint32_t s16 = H.s8[b0 * kChunk + dx] + H.s8[(b0 + 1) * kChunk + dx] + H.s8[(b0 + 8) * kChunk + dx] +
                             H.s8[(b0 + 9) * kChunk + dx];
        p.sad16[(size_t)before * (nctu * 16) + ((size_t)(ctu * 16 + pu) * PR + dy) * rs + dx] = (uint16_t)(s16 >> shift);
        if (pu < 4) {
          uint32_t s32 = 0;
          const int q0 = (pu >> 1) * 32 + (pu & 1) * 4;
#pragma unroll
          for (int yy = 0; yy < 4; yy++)
#pragma unroll
            for (int xx = 0; xx < 4; xx++) s32 += H.s8[(q0 + yy * 8 + xx) * kChunk + dx];
          p.sad32[(size_t)before * (nctu * 4) + ((size_t)(ctu * 4 + pu) * PR + dy) * rs + dx] = s32 >> shift;
        } else if (pu == 4) {
          uint32_t s64 = 0;
          for (int b = 0; b < 64; b++) s64 += H.s8[b * kChunk + dx];
          p.sad64[(size_t)before * nctu + ((size_t)ctu * PR + dy) * rs + dx] = s64 >> shift;
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace

// rectangular box: n_dx (multiple of 8) x n_dy displacements from (x0, y0), chunk-major maps; pitches of `p` in bytes
int launch_ime_sad10_box(ImeSadParams p, int x0, int y0, int n_dx, int n_dy, int bitdepth, cudaStream_t stream)
{
  if (n_dx <= 0 || (n_dx & 7) || n_dy <= 0) return fail_msg("ime_sad10_box: n_dx must be a positive multiple of 8, n_dy positive");
  p.x0 = x0; p.y0 = y0; p.skip = 0; p.n_dy = n_dy; p.pu_rows = n_dy;
  p.n_chunks = (n_dx + 15) / 16; p.n_dx_last = n_dx - 16 * (p.n_chunks - 1);
  p.chunk_major = 1;
  const size_t smem = smem10_bytes(n_dy + 63);
  if (smem > 200 * 1024) return fail_msg("ime_sad10_box: too many displacement rows for the shared-memory tile");
  KVZ_CUDA_TRY(cudaFuncSetAttribute(ime_sad10_box_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ime_sad10_box_kernel<<<p.n_chunks * p.ctus_x * p.ctus_y, kT, smem, stream>>>(p, bitdepth - 8);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
