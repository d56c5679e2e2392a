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
// Design (sm_100a), v2:
//  * the reference plane lives in HBM with replicated borders (pad_plane kernel) so a
//    tile fetch never needs per-pixel clamping;
//  * one CTA = one CTU x one chunk of 16 horizontal displacements (plus one light CTA per
//    CTU for the 65th column).  The (2R+64)-row reference tile is staged in shared memory
//    FOUR times, pre-shifted by 0..3 bytes, so that every operand of the inner loop is an
//    aligned word: no funnel shifts compete with the SAD instructions for the ALU pipe;
//  * one thread = one 8x8 block x 4 displacements dx = g + 4j of byte-shift g.  It slides
//    down the tile rows: 3 LDS (5 words) feed 64 VABSDIFF4.U8.ACC per row.  Eight (dy)
//    x four (dx) accumulators rotate through a statically indexed register window; a new
//    accumulator starts from the zero register instead of being cleared; the first and
//    the last 8-row iterations are specialised so that no product outside the window is
//    ever issued;
//  * finished rows go to a shared-memory ring (4 STS.U16, no packing, no global address
//    arithmetic in the compute loop).  Every 8 rows all threads turn writer: 16-byte
//    vectors of the 8x8 map, and the 16/32/64 sums formed with packed u16 adds and
//    shuffles, each map written exactly once in full 32-byte sectors.
#include <type_traits>
#include "common.cuh"
#include <cuda.h>
#include "ime_sad.cuh"

namespace kvz {

namespace {

constexpr int kThreads = 256;
constexpr int kRowWords = 20;         // staged words per tile row and copy (80 bytes)
constexpr int kChunkDx = 16;          // horizontal displacements per CTA

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t d;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

constexpr int kRawWords = 22;           // aligned source words per tile row (84 bytes + up to 3 of misalignment)

__device__ __forceinline__ void cp_async4(void *smem_dst, const void *gmem_src)
{
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_wait_all()
{
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// ring-slot position of an 8x8 block: the low two bits are swizzled with the 32x32 PU index so that
// the four PUs read by one warp of the writer fall into different banks
__device__ __forceinline__ int stage_slot(int blk) { return blk ^ (((blk >> 2) & 1) | ((blk >> 4) & 2)); }

// geometry of one launch (runtime): n_dy displacement rows need n_dy + 63 tile rows
struct Geo {
  int n_dy, rows, copy_words, iters;
  __host__ __device__ explicit Geo(int n_dy_) : n_dy(n_dy_)
  {
    rows = n_dy + 63;
    copy_words = ((rows * kRowWords + 31) & ~31) + 8;          // == 8 (mod 32): the 4 copies hit disjoint banks
    iters = (n_dy + 7) >> 3;                                    // 8 tile rows each; the last one is partial
  }
  __host__ __device__ size_t smem_bytes() const
  {
    const size_t stage = 8 * 64 * kChunkDx * 2, raw = (size_t)rows * kRawWords * 4;   // raw overlays the ring before the first iteration
    return (size_t)4 * copy_words * 4 + 64 * 64 + (raw > stage ? raw : stage);
  }
};

struct SmemView {
  uint32_t *tile;       // [4][kCopyWords]
  uint8_t *cur;         // [64][64]
  uint16_t *stage;      // [8][64][16]
};

// ---- compute: 8 tile rows of a normal chunk ----
// PH 0: first iteration (rows 0..7: only products with u - y >= S exist -- S leading rows of the computed
// range are alignment only -- and only dy 0 completes, if S == 0), PH 1: middle, PH 2: last iteration
// (only products with u <= y exist).  NJ: displacement columns per thread (4, or 2 in a half chunk).
template <int PH, int S, int NJ>
__device__ __forceinline__ void rows8(const uint32_t *trow, const uint32_t (&cur)[8][2], uint32_t (&acc)[8][4],
                                      uint16_t *stg /* &stage[0][blk][g] */)
{
#ifdef KVZ_SAD_EXP_NOCOMPUTE      // development: time the writer phase alone
  if (PH >= 0) return;
#endif
#pragma unroll
  for (int u = 0; u < 8; u++) {
    if (PH == 0 && u < S) continue;
    const uint32_t *rp = trow + u * kRowWords;
    const uint2 a = *reinterpret_cast<const uint2 *>(rp);
    const uint2 b = NJ > 2 ? *reinterpret_cast<const uint2 *>(rp + 2) : make_uint2(rp[2], 0u);
    const uint32_t w[5] = { a.x, a.y, b.x, b.y, NJ > 2 ? rp[4] : 0u };
#pragma unroll
    for (int y = 0; y < 8; y++) {
      if (PH == 0 && u - y < S) continue;
      if (PH == 2 && u > y) continue;
      const int slot = (u - y) & 7;                 // accumulator of dy = 8m + u - y
#pragma unroll
      for (int j = 0; j < NJ; j++) {
        const uint32_t c0 = (y == 0) ? 0u : acc[slot][j];      // a new dy starts from the zero register
        acc[slot][j] = sad4(cur[y][1], w[j + 1], sad4(cur[y][0], w[j], c0));
      }
    }
    // dy = 8m + u - 7 is complete (always in PH 1/2; in PH 0 only for u == 7 and S == 0)
    if (PH != 0 || (u == 7 && S == 0)) {
      const int slot = (u + 1) & 7;
      uint16_t *o = stg + u * (64 * kChunkDx);
#pragma unroll
      for (int j = 0; j < NJ; j++) o[4 * j] = (uint16_t)acc[slot][j];
    }
  }
}

// ---- writer: the 8 dy rows finished by one iteration of a normal chunk ----
// Per-thread output positions (element offsets; every map of a launch has < 2^31 entries).
struct WriterPos {
  uint32_t o8;        // sad8 entry of (blk, dy 0, first dx of this thread's half)
  uint32_t o16, o32, o64;
  uint32_t src8;      // u16 index of the thread's 8 dx inside a ring slot
  uint32_t src16;     // u32 index of (first block of the 32x32 PU, dx pair) inside a ring slot
  uint32_t rs;        // entries between consecutive dy rows of a PU
  uint32_t sx[4];     // u32 offset of block (i ^ swizzle) inside a group of four ring blocks, i = 0..3
};

// Two map layouts:
//  * square window (chunk_major == 0): map[pu][row][map_pitch], chunk c owns columns 16c .. 16c+15 of every row;
//  * box (chunk_major != 0): CHUNK-MAJOR, map[chunk][pu][row][nd] with nd = 16 (8 for a half last chunk): what one
//    CTA writes for a PU is one contiguous run of n_dy * nd entries, and for the 64 blocks of its CTU one run of
//    64 * n_dy * nd entries (147 KB at 1080p/full32) -- DRAM sees long sequential writes instead of 32-byte pieces
//    a row pitch apart (measured: the writer phases alone 0.87 -> 0.52 ms per frame).
__device__ __forceinline__ WriterPos writer_pos(const ImeSadParams &p, int ctu, int chunk, int tid, int nd)
{
  const uint32_t PR = (uint32_t)p.pu_rows;
  WriterPos w;
  const int blk = (tid >> 1) & 63, half = tid & 1;
  const int lane = tid & 31, p32 = lane >> 3, dxp = lane & 7;
  w.src8 = stage_slot(blk) * kChunkDx + half * 8;
  w.src16 = ((p32 >> 1) * 32 + (p32 & 1) * 4) * (kChunkDx / 2) + dxp;
#pragma unroll
  for (int i = 0; i < 4; i++) w.sx[i] = (uint32_t)((i ^ p32) * (kChunkDx / 2));     // stage_slot() of the PU's blocks: low two bits ^ p32
  const uint32_t pu8 = (uint32_t)ctu * 64 + blk, pu16 = (uint32_t)ctu * 16 + (p32 >> 1) * 8 + (p32 & 1) * 2,
                 pu32 = (uint32_t)ctu * 4 + p32, pu64 = (uint32_t)ctu;
  if (p.chunk_major) {
    const uint32_t rs = (uint32_t)nd, nctu = (uint32_t)(p.ctus_x * p.ctus_y);
    const uint32_t before = (uint32_t)chunk * kChunkDx * PR;        // entries per PU in the chunks before this one
    w.rs = rs;
    w.o8 = before * (nctu * 64) + pu8 * PR * rs + half * 8;
    w.o16 = before * (nctu * 16) + pu16 * PR * rs + 2 * dxp;
    w.o32 = before * (nctu * 4) + pu32 * PR * rs + 2 * dxp;
    w.o64 = before * nctu + pu64 * PR * rs + 2 * dxp;
  } else {
    const uint32_t rs = (uint32_t)p.map_pitch;
    const uint32_t col = chunk * kChunkDx + 2 * dxp;
    w.rs = rs;
    w.o8 = pu8 * PR * rs + chunk * kChunkDx + half * 8;
    w.o16 = pu16 * PR * rs + col;
    w.o32 = pu32 * PR * rs + col;
    w.o64 = pu64 * PR * rs + col;
  }
  return w;
}

// nd: displacement columns of this chunk (16 or 8)
__device__ __forceinline__ void write_rows(const ImeSadParams &p, const WriterPos &wp, const uint16_t *stage, int m, int tid, int nd)
{
#ifdef KVZ_SAD_EXP_NOWRITE        // development: time the compute phase alone
  if (p.skip >= 0) return;
#endif
  const uint32_t PR = (uint32_t)p.pu_rows;
  const uint32_t rs = wp.rs;
  const int k0 = 8 * m - 7 - p.skip;              // map row of ring slot 0
  // (a) 8x8 map: thread = (block, half of the 16 dx), ring slots u0, u0+2, u0+4, u0+6
  {
    const int u0 = tid >> 7;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int u = u0 + 2 * i, k = k0 + u;
      if (k >= 0 && (tid & 1) * 8 < nd)
        *reinterpret_cast<uint4 *>(p.sad8 + wp.o8 + (uint32_t)k * rs) =
            *reinterpret_cast<const uint4 *>(stage + u * (64 * kChunkDx) + wp.src8);
    }
  }
  // (b) 16x16 / 32x32 / 64x64: thread = (ring slot, 32x32 PU, pair of dx)
  {
    const int u = tid >> 5, p32 = (tid >> 3) & 3;
    const int k = k0 + u;
    if (k >= 0) {                                   // uniform per warp
      const uint32_t *st = reinterpret_cast<const uint32_t *>(stage) + u * (64 * kChunkDx / 2) + wp.src16;
      const bool wr = 2 * (tid & 7) < nd;           // this pair of displacement columns exists
      uint32_t lo = 0, hi = 0;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        // 16x16 PU q of the 32x32 PU: blocks (2(q>>1), 2(q&1)) + {0,1}x{0,1}; the swizzled low two bits of a
        // block index come from wp.sx, the rest is an immediate offset
        const int b = (q >> 1) * 16 + (q & 1) * 2;
        auto at = [&](int blk) { return st[(blk & ~3) * (kChunkDx / 2) + wp.sx[blk & 3]]; };
        // packed u16 pairs: 4 x 16320 < 65536, no carry between the halves
        const uint32_t s16 = at(b) + at(b + 1) + at(b + 8) + at(b + 9);
        if (wr) *reinterpret_cast<uint32_t *>(p.sad16 + wp.o16 + ((uint32_t)((q >> 1) * 4 + (q & 1)) * PR + (uint32_t)k) * rs) = s16;
        lo += s16 & 0xFFFFu;
        hi += s16 >> 16;
      }
      if (wr) *reinterpret_cast<uint2 *>(p.sad32 + wp.o32 + (uint32_t)k * rs) = make_uint2(lo, hi);
      lo += __shfl_xor_sync(0xffffffffu, lo, 8);  hi += __shfl_xor_sync(0xffffffffu, hi, 8);
      lo += __shfl_xor_sync(0xffffffffu, lo, 16); hi += __shfl_xor_sync(0xffffffffu, hi, 16);
      if (p32 == 0 && wr) *reinterpret_cast<uint2 *>(p.sad64 + wp.o64 + (uint32_t)k * rs) = make_uint2(lo, hi);
    }
  }
}

__device__ __forceinline__ void stage_cur(const ImeSadParams &p, uint8_t *cur, int ctu_x, int ctu_y, int tid, int nthreads)
{
  // current CTU, coordinates clamped into the picture for partial CTUs
  for (int i = tid; i < 64 * 8; i += nthreads) {
    const int r = i >> 3, v = i & 7;
    const int y = min(ctu_y + r, p.height - 1);
    const int x = min(ctu_x + v * 8, p.width - 8);
    *reinterpret_cast<uint2 *>(&cur[r * 64 + v * 8]) = *reinterpret_cast<const uint2 *>(p.cur + (size_t)y * p.cur_pitch + x);
  }
}

// ---- the last column (dx = +R) of every map, stored as row K of each PU's map (entry dy + R) ----
// One CTA per CTU; thread = (row of 8x8 blocks, dy) with dy along the lanes, so that the K entries of a
// PU are written as one contiguous, fully covered row (a column of 2-byte entries at the row pitch
// would cost a DRAM read-modify-write per entry).
constexpr int kLcWords = 17;            // tile row pitch in words: odd, lanes (= rows) fall into distinct banks

template <int R>
struct LcSmem {
  uint32_t tile[(2 * R + 64) * kLcWords];
  alignas(16) uint8_t cur[64 * 64];
  uint4 part16[8][((2 * R + 1) + 15) & ~15];     // per block row: sums of horizontally adjacent 8x8 pairs
  uint2 part32[4][((2 * R + 1) + 15) & ~15];     // per 16x16 row: sums of horizontally adjacent 16x16 pairs
  uint32_t part64[2][((2 * R + 1) + 15) & ~15];
};

template <int R>
__device__ __forceinline__ void last_column(const ImeSadParams &p, unsigned char *smem_raw, int ctu)
{
  constexpr int K = 2 * R + 1;
  LcSmem<R> &sm = *reinterpret_cast<LcSmem<R> *>(smem_raw);
  const int tid = threadIdx.x;
  const int ctu_x = p.tile_x + (ctu % p.ctus_x) * 64, ctu_y = p.tile_y + (ctu / p.ctus_x) * 64;
  const int rs = p.map_pitch;
  {
    const uint8_t *src = p.ref + (size_t)(ctu_y + p.y0 + p.pad) * p.ref_pitch + (ctu_x + p.x0 + 2 * R + p.pad);
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(src);
    const int sh = (int)(a0 & 3) * 8;
    const uint8_t *src_al = reinterpret_cast<const uint8_t *>(a0 & ~(uintptr_t)3);
#pragma unroll 4
    for (int i = tid; i < (2 * R + 64) * 16; i += kThreads) {
      const int r = i >> 4, mw = i & 15;
      const uint32_t *w = reinterpret_cast<const uint32_t *>(src_al + (size_t)r * p.ref_pitch) + mw;
      sm.tile[r * kLcWords + mw] = __funnelshift_r(w[0], w[1], sh);
    }
    stage_cur(p, sm.cur, ctu_x, ctu_y, tid, kThreads);
  }
  __syncthreads();
  // 8x8: item = (block row by, entry k); entries K .. pitch-1 of the row are written as zero
  for (int item = tid; item < 8 * rs; item += kThreads) {
    const int by = item / rs, k = item - by * rs;
    uint32_t s8[8];
#pragma unroll
    for (int bx = 0; bx < 8; bx++) {
      uint32_t acc = 0;
      if (k < K) {
#pragma unroll
        for (int y = 0; y < 8; y++) {
          const uint2 c = *reinterpret_cast<const uint2 *>(&sm.cur[(by * 8 + y) * 64 + bx * 8]);
          const uint32_t *t = &sm.tile[(by * 8 + y + k) * kLcWords + 2 * bx];
          acc = sad4(c.y, t[1], sad4(c.x, t[0], acc));
        }
      }
      s8[bx] = acc;
      p.sad8[(((size_t)ctu * 64 + by * 8 + bx) * (K + 1) + K) * rs + k] = (uint16_t)acc;
    }
    sm.part16[by][k] = make_uint4(s8[0] + s8[1], s8[2] + s8[3], s8[4] + s8[5], s8[6] + s8[7]);
  }
  __syncthreads();
  for (int item = tid; item < 4 * rs; item += kThreads) {
    const int r16 = item / rs, k = item - r16 * rs;
    const uint4 a = sm.part16[2 * r16][k], b = sm.part16[2 * r16 + 1][k];
    const uint32_t s16[4] = { a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w };
#pragma unroll
    for (int q = 0; q < 4; q++)
      p.sad16[(((size_t)ctu * 16 + r16 * 4 + q) * (K + 1) + K) * rs + k] = (uint16_t)s16[q];
    sm.part32[r16][k] = make_uint2(s16[0] + s16[1], s16[2] + s16[3]);
  }
  __syncthreads();
  for (int item = tid; item < 2 * rs; item += kThreads) {
    const int r32 = item / rs, k = item - r32 * rs;
    const uint2 a = sm.part32[2 * r32][k], b = sm.part32[2 * r32 + 1][k];
    p.sad32[(((size_t)ctu * 4 + r32 * 2 + 0) * (K + 1) + K) * rs + k] = a.x + b.x;
    p.sad32[(((size_t)ctu * 4 + r32 * 2 + 1) * (K + 1) + K) * rs + k] = a.y + b.y;
    sm.part64[r32][k] = a.x + b.x + a.y + b.y;
  }
  __syncthreads();
  for (int k = tid; k < rs; k += kThreads)
    p.sad64[((size_t)ctu * (K + 1) + K) * rs + k] = sm.part64[0][k] + sm.part64[1][k];
}

#ifndef KVZ_SAD_MINB
#define KVZ_SAD_MINB 2
#endif
// one CTA = 16 dx (one chunk) x all dy of one CTU.  R > 0: square window, the first n_ctu CTAs do the last
// column instead (they are short and latency bound: started first, they overlap the rest); R == 0: box
// named barrier over `count` threads (count a multiple of 32); barrier 0 with the whole CTA == __syncthreads()
__device__ __forceinline__ void bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }

// Stages the reference tile of (ctu, chunk) as four byte-shifted copies and the current CTU; kThreads threads
// (tid = 0..kThreads-1) that share named barrier `bar`.  `raw` is scratch for rows * kRawWords words.
__device__ __forceinline__ void stage_tile(const ImeSadParams &p, const Geo &geo, uint32_t *tile, uint32_t *raw, uint8_t *cur_s,
                                           int ctu, int chunk, int tid, int bar)
{
  const int CW = geo.copy_words;
  const int ctu_x = p.tile_x + (ctu % p.ctus_x) * 64, ctu_y = p.tile_y + (ctu / p.ctus_x) * 64;
  // tile rows ctu_y+y0 .., 84 bytes from column ctu_x+x0+16*chunk
  const uint8_t *src = p.ref + (size_t)(ctu_y + p.y0 + p.pad) * p.ref_pitch + (ctu_x + p.x0 + kChunkDx * chunk + p.pad);
  const uintptr_t a0 = reinterpret_cast<uintptr_t>(src);
  const int sh = (int)(a0 & 3) * 8;
  const uint8_t *src_al = reinterpret_cast<const uint8_t *>(a0 & ~(uintptr_t)3);
  // (1) raw rows (22 aligned words each) straight into shared memory, all copies in flight at once
  for (int i = tid; i < geo.rows * kRawWords; i += kThreads) {
    const int r = i / kRawWords, mw = i - r * kRawWords;
    cp_async4(raw + i, src_al + (size_t)r * p.ref_pitch + 4 * mw);
  }
  stage_cur(p, cur_s, ctu_x, ctu_y, tid, kThreads);
  cp_async_wait_all();
  bar_sync(bar, kThreads);
  // (2) the four byte-shifted copies
  for (int i = tid; i < geo.rows * kRowWords; i += kThreads) {
    const int r = i / kRowWords, mw = i - r * kRowWords;
    const uint32_t *w = raw + r * kRawWords + mw;
    const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
    const uint32_t t0 = __funnelshift_r(w0, w1, sh), t1 = __funnelshift_r(w1, w2, sh);
    tile[0 * CW + i] = t0;
    tile[1 * CW + i] = __funnelshift_r(t0, t1, 8);
    tile[2 * CW + i] = __funnelshift_r(t0, t1, 16);
    tile[3 * CW + i] = __funnelshift_r(t0, t1, 24);
  }
  bar_sync(bar, kThreads);
}

// ---- the same staging with the raw rows fetched by ONE TMA tile load (the default; KVZ_CUDA_SAD_TMA=0 switches it off) ----
// cp.async.bulk.tensor.2d brings the box {kTmaRowBytes bytes x rows} that starts at the 16-byte boundary at or before
// the window's first sample into `raw` (128-byte aligned, row pitch kTmaRowBytes) and signals an mbarrier with the
// byte count; thread 0 issues it, every compute thread waits on the barrier's phase.  The four byte-shifted copies are
// then made as before, the alignment shift now being 0..15 bytes.
// Measured on B200: the innermost TMA coordinate must be a multiple of 16 BYTES -- with uint8 elements any other x
// raises "an illegal instruction was encountered" (tools/tma_probe.py) -- so TMA cannot deliver the four copies
// pre-shifted by 0..3 bytes (four loads at x + 0..3), and 128-byte aligned copy bases would also put all four copies
// on the same banks (they are staggered by 8 banks on purpose, see Geo).
constexpr int kTmaRowBytes = 112;          // 84 bytes of window + up to 15 of alignment, rounded up to 16
constexpr int kTmaRowWords = kTmaRowBytes / 4;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void stage_tile_tma(const ImeSadParams &p, const CUtensorMap *tmap, const Geo &geo, uint32_t *tile,
                                               uint32_t *raw, uint8_t *cur_s, uint64_t *mbar, int ctu, int chunk, int tid, int bar)
{
  const int CW = geo.copy_words;
  const int ctu_x = p.tile_x + (ctu % p.ctus_x) * 64, ctu_y = p.tile_y + (ctu / p.ctus_x) * 64;
  const int gx = ctu_x + p.x0 + kChunkDx * chunk + p.pad, gy = ctu_y + p.y0 + p.pad;
  const int woff = (gx & 15) >> 2, sh = (gx & 3) * 8;
  if (tid == 0) {
    const uint32_t mb = smem_u32(mbar), dst = smem_u32(raw);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(geo.rows * kTmaRowBytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(mb), "r"(gx & ~15), "r"(gy) : "memory");
  }
  stage_cur(p, cur_s, ctu_x, ctu_y, tid, kThreads);
  bar_sync(bar, kThreads);                                   // the mbarrier is initialised, the current CTU staged
  {
    const uint32_t mb = smem_u32(mbar);
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                   : "=r"(done) : "r"(mb), "r"(0) : "memory");
  }
  for (int i = tid; i < geo.rows * kRowWords; i += kThreads) {
    const int r = i / kRowWords, mw = i - r * kRowWords;
    const uint32_t *w = raw + r * kTmaRowWords + woff + mw;
    const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
    const uint32_t t0 = __funnelshift_r(w0, w1, sh), t1 = __funnelshift_r(w1, w2, sh);
    tile[0 * CW + i] = t0;
    tile[1 * CW + i] = __funnelshift_r(t0, t1, 8);
    tile[2 * CW + i] = __funnelshift_r(t0, t1, 16);
    tile[3 * CW + i] = __funnelshift_r(t0, t1, 24);
  }
  bar_sync(bar, kThreads);
}

template <int NJ>
__device__ __forceinline__ void first_rows8(int skip, const uint32_t *trow, const uint32_t (&cur)[8][2], uint32_t (&acc)[8][4], uint16_t *stg)
{
  switch (skip) {
    case 0: rows8<0, 0, NJ>(trow, cur, acc, stg); break;
    case 1: rows8<0, 1, NJ>(trow, cur, acc, stg); break;
    case 2: rows8<0, 2, NJ>(trow, cur, acc, stg); break;
    case 3: rows8<0, 3, NJ>(trow, cur, acc, stg); break;
    case 4: rows8<0, 4, NJ>(trow, cur, acc, stg); break;
    case 5: rows8<0, 5, NJ>(trow, cur, acc, stg); break;
    case 6: rows8<0, 6, NJ>(trow, cur, acc, stg); break;
    default: rows8<0, 7, NJ>(trow, cur, acc, stg); break;
  }
}

// one chunk CTA (kThreads threads): 16 (or 8) dx x all dy of one CTU; compute and writer phases alternate
__device__ __forceinline__ void chunk_cta(const ImeSadParams &p, unsigned char *smem_raw, int ctu, int chunk)
{
  const int tid = threadIdx.x;
  const Geo geo(p.n_dy);
  const int CW = geo.copy_words;
  uint32_t *tile = reinterpret_cast<uint32_t *>(smem_raw);
  uint8_t *cur_s = smem_raw + (size_t)4 * CW * 4;
  uint16_t *stage = reinterpret_cast<uint16_t *>(cur_s + 64 * 64);
  stage_tile(p, geo, tile, reinterpret_cast<uint32_t *>(stage), cur_s, ctu, chunk, tid, 0);
  const int warp = tid >> 5, lane = tid & 31;
  const int by = warp, bx = lane >> 2, g = lane & 3, blk = by * 8 + bx;
  uint32_t cur[8][2];
#pragma unroll
  for (int y = 0; y < 8; y++) {
    const uint2 v = *reinterpret_cast<const uint2 *>(&cur_s[(by * 8 + y) * 64 + bx * 8]);
    cur[y][0] = v.x; cur[y][1] = v.y;
  }
  uint32_t acc[8][4];
  const uint32_t *trow = tile + g * CW + (by * 8) * kRowWords + 2 * bx;
  uint16_t *stg = stage + stage_slot(blk) * kChunkDx + g;
  const int nd = chunk == p.n_chunks - 1 ? p.n_dx_last : kChunkDx;
  const WriterPos wp = writer_pos(p, ctu, chunk, tid, nd);
  auto run = [&](auto nj_tag) {
    constexpr int NJ = decltype(nj_tag)::value;
    first_rows8<NJ>(p.skip, trow, cur, acc, stg);
    __syncthreads();
    if (p.skip == 0) write_rows(p, wp, stage, 0, tid, nd);
    __syncthreads();
    for (int m = 1; m < geo.iters - 1; m++) {
      rows8<1, 0, NJ>(trow + m * 8 * kRowWords, cur, acc, stg);
      __syncthreads();
      write_rows(p, wp, stage, m, tid, nd);
      __syncthreads();
    }
    rows8<2, 0, NJ>(trow + (geo.iters - 1) * 8 * kRowWords, cur, acc, stg);
    __syncthreads();
    write_rows(p, wp, stage, geo.iters - 1, tid, nd);
  };
  if (nd == kChunkDx) run(std::integral_constant<int, 4>());
  else run(std::integral_constant<int, 2>());
}

// ---- warp-specialised chunk CTA: kThreads compute threads + kWriterThreads writer threads ----
// Measured on B200 (1080p, four boxes): the compute phases alone take 0.80 ms per frame and the writer phases
// alone 0.87 ms, but alternating them inside a CTA gives 1.05 ms even with two CTAs per SM: whenever both CTAs
// write, the integer pipe idles.  Here the compute warps never stop: finished rows go to one of two ring
// buffers, and two writer warps drain the other one (16-byte vectors of the 8x8 map, 16/32/64 sums) behind
// named barriers.  full[b]: compute arrive -> writers wait; empty[b]: writers arrive -> compute wait.
#ifndef KVZ_SAD_WRITERS
#define KVZ_SAD_WRITERS 128
#endif
constexpr int kWriterThreads = KVZ_SAD_WRITERS;
constexpr int kWsThreads = kThreads + kWriterThreads;
constexpr int kRingEntries = 8 * 64 * kChunkDx;               // uint16 entries of one ring buffer
#ifndef KVZ_SAD_RING
#define KVZ_SAD_RING 2
#endif
constexpr int kRing = KVZ_SAD_RING;                           // ring buffers between compute and writer warps
enum { kBarCompute = 1, kBarFull0 = 2, kBarEmpty0 = 2 + kRing };

__host__ __device__ inline size_t ws_smem_bytes(const Geo &geo, bool tma = false)
{
  const size_t raw = (size_t)geo.rows * (tma ? 112 : kRawWords * 4), ring = (size_t)kRing * kRingEntries * 2;
  return (size_t)4 * geo.copy_words * 4 + 64 * 64 + (raw > ring ? raw : ring);
}

template <bool TMA>
__device__ __forceinline__ void chunk_cta_ws(const ImeSadParams &p, unsigned char *smem_raw, int ctu, int chunk,
                                             const CUtensorMap *tmap = nullptr, uint64_t *mbar = nullptr)
{
  const int tid = threadIdx.x;
  const Geo geo(p.n_dy);
  const int CW = geo.copy_words;
  uint32_t *tile = reinterpret_cast<uint32_t *>(smem_raw);
  uint8_t *cur_s = smem_raw + (size_t)4 * CW * 4;
  uint16_t *stage = reinterpret_cast<uint16_t *>(cur_s + 64 * 64);        // [2][kRingEntries]
  const int nd = chunk == p.n_chunks - 1 ? p.n_dx_last : kChunkDx;
  const int iters = geo.iters;
  if (tid >= kThreads) {
    // ---- writer warps: each thread stands in for four threads of the 256-thread writer layout ----
    const int wt = tid - kThreads;
    constexpr int kV = kThreads / kWriterThreads;
    WriterPos wp[kV];
#pragma unroll
    for (int v = 0; v < kV; v++) wp[v] = writer_pos(p, ctu, chunk, wt + v * kWriterThreads, nd);
    for (int m = 0, b = 0; m < iters; m++, b = b + 1 == kRing ? 0 : b + 1) {
      bar_sync(kBarFull0 + b, kWsThreads);
      if (m > 0 || p.skip == 0) {
#pragma unroll
        for (int v = 0; v < kV; v++) write_rows(p, wp[v], stage + b * kRingEntries, m, wt + v * kWriterThreads, nd);
      }
      if (m + kRing < iters) bar_arrive(kBarEmpty0 + b, kWsThreads);
    }
    return;
  }
  // ---- compute warps ----
  if (TMA) stage_tile_tma(p, tmap, geo, tile, reinterpret_cast<uint32_t *>(stage), cur_s, mbar, ctu, chunk, tid, kBarCompute);
  else stage_tile(p, geo, tile, reinterpret_cast<uint32_t *>(stage), cur_s, ctu, chunk, tid, kBarCompute);
  const int warp = tid >> 5, lane = tid & 31;
  const int by = warp, bx = lane >> 2, g = lane & 3, blk = by * 8 + bx;
  uint32_t cur[8][2];
#pragma unroll
  for (int y = 0; y < 8; y++) {
    const uint2 v = *reinterpret_cast<const uint2 *>(&cur_s[(by * 8 + y) * 64 + bx * 8]);
    cur[y][0] = v.x; cur[y][1] = v.y;
  }
  uint32_t acc[8][4];
  const uint32_t *trow = tile + g * CW + (by * 8) * kRowWords + 2 * bx;
  uint16_t *stg = stage + stage_slot(blk) * kChunkDx + g;
  auto run = [&](auto nj_tag) {
    constexpr int NJ = decltype(nj_tag)::value;
    first_rows8<NJ>(p.skip, trow, cur, acc, stg);                 // the raw rows (overlaid on the ring) are dead by now
    bar_arrive(kBarFull0, kWsThreads);
    int b = kRing > 1 ? 1 : 0;
    for (int m = 1; m < iters - 1; m++) {
      if (m >= kRing) bar_sync(kBarEmpty0 + b, kWsThreads);
      rows8<1, 0, NJ>(trow + m * 8 * kRowWords, cur, acc, stg + b * kRingEntries);
      bar_arrive(kBarFull0 + b, kWsThreads);
      b = b + 1 == kRing ? 0 : b + 1;
    }
    if (iters > 1) {                                              // (a single iteration is complete after the first one)
      const int m = iters - 1;
      if (m >= kRing) bar_sync(kBarEmpty0 + b, kWsThreads);
      rows8<2, 0, NJ>(trow + m * 8 * kRowWords, cur, acc, stg + b * kRingEntries);
      bar_arrive(kBarFull0 + b, kWsThreads);
    }
  };
  if (nd == kChunkDx) run(std::integral_constant<int, 4>());
  else run(std::integral_constant<int, 2>());
}

template <int R>
__global__ void __launch_bounds__(kThreads, KVZ_SAD_MINB)
ime_sad_dense_kernel(ImeSadParams p)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n_ctu = R > 0 ? p.ctus_x * p.ctus_y : 0;
  if (R > 0 && (int)blockIdx.x < n_ctu) {
    if constexpr (R > 0) last_column<R>(p, smem_raw, blockIdx.x);
  } else {
    const int bid = blockIdx.x - n_ctu;
    chunk_cta(p, smem_raw, bid / p.n_chunks, bid % p.n_chunks);
  }
}

// The box maps of SEVERAL references of one current picture in one launch: one grid instead of one per
// reference removes the per-launch ramp and tail (8.6 waves each at 1080p), and the CTAs are ordered heavy
// first -- the full 16-column chunks of every reference, then the 8-column remainders, which fill the tail.
struct ImeSadMulti {
  ImeSadParams p[kImeMaxMulti];
  int full_end[kImeMaxMulti];     // cumulative CTA count after reference r
  int n;
  int ctu_only;                   // >= 0: the maps of this one CTU only (grid = the chunks of every reference)
};

__global__ void __launch_bounds__(kWsThreads, 2)
ime_sad_multi_kernel(const __grid_constant__ ImeSadMulti m)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int bid = blockIdx.x;
  int r = 0;
  while (bid >= m.full_end[r]) r++;
  const int local = bid - (r ? m.full_end[r - 1] : 0);
  const int ctu = m.ctu_only >= 0 ? m.ctu_only : local / m.p[r].n_chunks;
  const int chunk = m.ctu_only >= 0 ? local : local - ctu * m.p[r].n_chunks;
  chunk_cta_ws<false>(m.p[r], smem_raw, ctu, chunk);
}

// the same kernel, raw tile rows by cp.async.bulk.tensor (one tensor map per reference plane)
struct ImeSadMultiTma {
  ImeSadMulti m;
  CUtensorMap maps[kImeMaxMulti];
};

__global__ void __launch_bounds__(kWsThreads, 2)
ime_sad_multi_tma_kernel(const __grid_constant__ ImeSadMultiTma t)
{
  extern __shared__ __align__(128) unsigned char smem_tma[];
  __shared__ __align__(8) uint64_t mbar;
  const ImeSadMulti &m = t.m;
  const int bid = blockIdx.x;
  int r = 0;
  while (bid >= m.full_end[r]) r++;
  const int local = bid - (r ? m.full_end[r - 1] : 0);
  const int ctu = m.ctu_only >= 0 ? m.ctu_only : local / m.p[r].n_chunks;
  const int chunk = m.ctu_only >= 0 ? local : local - ctu * m.p[r].n_chunks;
  chunk_cta_ws<true>(m.p[r], smem_tma, ctu, chunk, &t.maps[r], &mbar);
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

// Re-replicate the border of an already bordered plane after the frame rectangle [x0, x0 + w) x [y0, y0 + h) of its
// interior changed: every border sample whose clamped source lies in the rectangle.  In place.
template <typename PIX>
__global__ void pad_patch_kernel(PIX *plane, int W, int H, int pad, int pitch, int ex0, int ey0, int ew, int eh)
{
  const int x = ex0 + blockIdx.x * blockDim.x + threadIdx.x, y = ey0 + blockIdx.y;
  if (x >= ex0 + ew || y >= ey0 + eh) return;
  if (x >= 0 && x < W && y >= 0 && y < H) return;                  // interior: already there
  const int sx = clampi(x, 0, W - 1), sy = clampi(y, 0, H - 1);
  plane[(size_t)(y + pad) * pitch + (x + pad)] = plane[(size_t)(sy + pad) * pitch + (sx + pad)];
}

}  // namespace

int launch_pad_patch(void *plane, int W, int H, int pad, int pitch, int x0, int y0, int w, int h, int bitdepth,
                     cudaStream_t stream)
{
  if (x0 > 0 && y0 > 0 && x0 + w < W && y0 + h < H) return 1;      // touches no edge
  const int ex0 = x0 == 0 ? -pad : x0, ey0 = y0 == 0 ? -pad : y0;
  const int ex1 = x0 + w == W ? W + pad : x0 + w, ey1 = y0 + h == H ? H + pad : y0 + h;
  dim3 grid((ex1 - ex0 + 127) / 128, ey1 - ey0);
  if (bitdepth == 8) pad_patch_kernel<uint8_t><<<grid, 128, 0, stream>>>((uint8_t *)plane, W, H, pad, pitch, ex0, ey0, ex1 - ex0, ey1 - ey0);
  else pad_patch_kernel<uint16_t><<<grid, 128, 0, stream>>>((uint16_t *)plane, W, H, pad, pitch, ex0, ey0, ex1 - ex0, ey1 - ey0);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

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

template <int R>
static int launch_r(const ImeSadParams &p, cudaStream_t stream)
{
  auto kern = ime_sad_dense_kernel<R>;
  const Geo geo(p.n_dy);
  size_t smem = geo.smem_bytes();
  if (R > 0 && sizeof(LcSmem<(R > 0 ? R : 8)>) > smem) smem = sizeof(LcSmem<(R > 0 ? R : 8)>);   // last-column CTAs share the launch
  if (smem > 113 * 1024) return fail_msg("ime_sad: too many displacement rows for the shared-memory tile");
  KVZ_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int n_ctu = p.ctus_x * p.ctus_y;
  kern<<<(p.n_chunks + (R > 0 ? 1 : 0)) * n_ctu, kThreads, smem, stream>>>(p);   // per CTU: [dx = +R,] then chunks of 16 dx
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_ime_sad_dense(ImeSadParams p, int range, int center_x, int center_y, cudaStream_t stream)
{
  p.x0 = center_x - range; p.y0 = center_y - range;
  p.n_dy = 2 * range + 1; p.n_chunks = range / 8; p.pu_rows = p.n_dy + 1;
  p.skip = 0; p.n_dx_last = kChunkDx; p.chunk_major = 0;
  switch (range) {
    case 8:  return launch_r<8>(p, stream);
    case 16: return launch_r<16>(p, stream);
    case 32: return launch_r<32>(p, stream);
    case 64: return launch_r<64>(p, stream);
    default: return fail_msg("ime_sad_dense: range must be 8, 16, 32 or 64");
  }
}

static int box_geometry(ImeSadParams &p, int x0, int y0, int n_dx, int n_dy)
{
  if (n_dx <= 0 || (n_dx & 7) || n_dy <= 0) return fail_msg("ime_sad_box: n_dx must be a positive multiple of 8, n_dy positive");
  // the kernel computes 8k + 1 rows; the surplus goes in front (alignment rows that are never written)
  p.skip = (8 - ((n_dy - 1) & 7)) & 7;
  p.x0 = x0; p.y0 = y0 - p.skip; p.n_dy = n_dy + p.skip; p.pu_rows = n_dy;
  p.n_chunks = (n_dx + 15) / 16; p.n_dx_last = n_dx - 16 * (p.n_chunks - 1);
  p.chunk_major = 1;
  return 1;
}

int launch_ime_sad_box(ImeSadParams p, int x0, int y0, int n_dx, int n_dy, cudaStream_t stream)
{
  if (!box_geometry(p, x0, y0, n_dx, n_dy)) return 0;
  return launch_r<0>(p, stream);
}

// the largest n_dy launch_ime_sad_box_multi accepts: the CTA stages n_dy (+ up to 7 alignment rows) + 63 reference rows
int ime_box_max_rows()
{
  int best = 1;
  for (int n = 1; n <= 272; n++) {
    const int skip = (8 - ((n - 1) & 7)) & 7;
    if (ws_smem_bytes(Geo(n + skip)) <= 113 * 1024) best = n;
  }
  return best;
}

typedef CUresult (*TmaEncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// the driver's cuTensorMapEncodeTiled, looked up once; null when this driver does not have it (then: cp.async staging)
static TmaEncodeFn tma_encode_fn()
{
  static const TmaEncodeFn fn = [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      p = nullptr;
    }
    return reinterpret_cast<TmaEncodeFn>(p);
  }();
  return fn;
}

// The same launch with the raw tile rows fetched by TMA.  One tensor map per reference: the bordered
// plane as a 2-D byte tensor {pitch, rows}, box {112 bytes, n_dy + 63 rows} (at most 256 rows per box: 240 here).
static int launch_multi_tma(const ImeSadMulti &m, cudaStream_t stream)
{
  const TmaEncodeFn encode = tma_encode_fn();
  if (!encode) return fail_msg("ime_sad (TMA): cuTensorMapEncodeTiled is not available");
  ImeSadMultiTma t;
  t.m = m;
  size_t smem = 0;
  for (int r = 0; r < m.n; r++) {
    const ImeSadParams &p = m.p[r];
    const Geo geo(p.n_dy);
    if (geo.rows > 256) return fail_msg("ime_sad (TMA): more than 256 tile rows");
    if ((reinterpret_cast<uintptr_t>(p.ref) & 15) || (p.ref_pitch & 15)) return fail_msg("ime_sad (TMA): plane base and pitch must be multiples of 16");
    const cuuint64_t dims[2] = { (cuuint64_t)p.ref_pitch, (cuuint64_t)(p.height + 2 * p.pad) };
    const cuuint64_t strides[1] = { (cuuint64_t)p.ref_pitch };
    const cuuint32_t box[2] = { (cuuint32_t)kTmaRowBytes, (cuuint32_t)geo.rows };
    const cuuint32_t estr[2] = { 1, 1 };
    const CUresult cr = encode(&t.maps[r], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t *>(p.ref), dims, strides, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) return fail_msg("ime_sad (TMA): cuTensorMapEncodeTiled failed");
    const size_t sb = ws_smem_bytes(geo, true);
    if (sb > smem) smem = sb;
  }
  if (smem > 113 * 1024) return fail_msg("ime_sad (TMA): too many displacement rows for the shared-memory tile");
  static size_t opted = 0;
  static std::mutex opt_lock;
  KVZ_CUDA_TRY(opt_in_dynamic_smem(ime_sad_multi_tma_kernel, smem, opted, opt_lock));
  ime_sad_multi_tma_kernel<<<m.full_end[m.n - 1], kWsThreads, smem, stream>>>(t);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_ime_sad_box_multi(const ImeSadParams *ps, const int (*boxes)[4], int n, cudaStream_t stream, int ctu_only)
{
  if (n <= 0 || n > kImeMaxMulti) return fail_msg("ime_sad_box_multi: 1..8 references per launch");
  ImeSadMulti m;
  memset(&m, 0, sizeof(m));
  m.n = n;
  m.ctu_only = ctu_only;
  size_t smem = 0;
  int total = 0;
  for (int r = 0; r < n; r++) {
    m.p[r] = ps[r];
    if (!box_geometry(m.p[r], boxes[r][0], boxes[r][1], boxes[r][2], boxes[r][3])) return 0;
    const size_t sb = ws_smem_bytes(Geo(m.p[r].n_dy));
    if (sb > smem) smem = sb;
    const int n_ctu = m.p[r].ctus_x * m.p[r].ctus_y;
    if (ctu_only >= n_ctu) return fail_msg("ime_sad_box_multi: CTU index outside the tile");
    total += (ctu_only >= 0 ? 1 : n_ctu) * m.p[r].n_chunks;
    m.full_end[r] = total;
  }
  if (smem > 113 * 1024) return fail_msg("ime_sad: too many displacement rows for the shared-memory tile");
  // raw tile rows by TMA (stage_tile_tma; measured 2 % faster than the cp.async staging at 1080p) whenever every
  // plane can be described by a tensor map: base and pitch multiples of 16 bytes.  KVZ_CUDA_SAD_TMA=0: cp.async.
  static const bool allow_tma = !(getenv("KVZ_CUDA_SAD_TMA") && atoi(getenv("KVZ_CUDA_SAD_TMA")) == 0);
  bool use_tma = allow_tma && tma_encode_fn() != nullptr;
  for (int r = 0; r < n; r++)
    use_tma = use_tma && !(reinterpret_cast<uintptr_t>(m.p[r].ref) & 15) && !(m.p[r].ref_pitch & 15) && Geo(m.p[r].n_dy).rows <= 256;
  if (use_tma) return launch_multi_tma(m, stream);
  static size_t opted = 0;                        // opt in once per size (any host thread may launch this)
  static std::mutex opt_lock;
  KVZ_CUDA_TRY(opt_in_dynamic_smem(ime_sad_multi_kernel, smem, opted, opt_lock));
  ime_sad_multi_kernel<<<total, kWsThreads, smem, stream>>>(m);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
