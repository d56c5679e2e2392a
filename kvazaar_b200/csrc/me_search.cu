// kvazaar_b200/csrc/me_search.cu
//
// Per-PU motion decision on device: for every PU of a batch and every reference
// picture it is searched against, reproduce search_pu_inter_ref()
// (search_inter.c:1134-1297) minus the candidate derivation (the AMVP / merge
// candidates are inputs).  Two kernels:
//
//   me_scan_kernel   integer search = search_mv_full (search_inter.c:886-956), one warp per
//      (PU, reference) job -- the jobs of a PU do not depend on each other here.  Windows
//      whose centre has a dense SAD map set (ime_sad.cu) are read from it, the others are
//      evaluated on demand from the border-replicated reference plane.  Every candidate
//      costs SAD + calc_mvd_cost (search_inter.c:373-412) and the winner is the FIRST
//      candidate in the reference's visiting order with the smallest cost (check_mv_cost
//      accepts on strict '<', search_inter.c:195-232): a lexicographic arg-min over
//      (cost, window, row, column).  The winner is parked in the job's result record.
//   me_refine_kernel  per PU, references in order (the `*inter_cost` carry of
//      search_pu_inter, search_inter.c:1505-1511, makes this part sequential):
//      fractional refinement = search_frac (search_inter.c:965-1128) with the 8-tap
//      filters of kvz_filter_{hpel,qpel}_blocks_*_luma (ipol-generic.c:192-658, int16
//      intermediates included) and kvz_satd_any_size_quad, four candidates per step,
//      gated by `best_cost < *inter_cost`; otherwise SATD at the integer vector
//      (search_inter.c:1244-1256); then the merge-index match and select_mv_cand
//      (search_inter.c:1258-1297).  A warp works on a PU up to 16x16 (four PUs per
//      CTA), the whole 256-thread CTA on a larger one.
#include <stdlib.h>

#include "common.cuh"
#include "me_search.cuh"
#include "satd.cuh"

namespace kvz {
namespace {

constexpr int kCtaThreads = 128;
#ifndef KVZ_BIG_GROUP
#define KVZ_BIG_GROUP 256
#endif
constexpr int kBigGroup = KVZ_BIG_GROUP;    // threads working on one PU larger than 16x16

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t d;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// HEVC luma DCT-IF (filter.c:54-60): { 0, 0, 0, 64, 0, 0, 0, 0 }, { -1, 4, -10, 58, 17, -5, 1, 0 },
// { -1, 4, -11, 40, 40, -11, 4, -1 }, { 0, 1, -5, 17, 58, -10, 4, -1 }, packed as signed bytes
// (t0..t3) and (t4..t7) for dp4a / dp2a
#define KVZ_PACK4(a, b, c, d) ((uint32_t)((a) & 0xFF) | ((uint32_t)((b) & 0xFF) << 8) | ((uint32_t)((c) & 0xFF) << 16) | ((uint32_t)((d) & 0xFF) << 24))
__device__ __constant__ uint32_t c_fir4[4][2] = {
  { KVZ_PACK4(0, 0, 0, 64), KVZ_PACK4(0, 0, 0, 0) }, { KVZ_PACK4(-1, 4, -10, 58), KVZ_PACK4(17, -5, 1, 0) },
  { KVZ_PACK4(-1, 4, -11, 40), KVZ_PACK4(40, -11, 4, -1) }, { KVZ_PACK4(0, 1, -5, 17), KVZ_PACK4(58, -10, 4, -1) } };

// c + sum of 4 (unsigned byte of a) x (signed byte of b)
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c)
{
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
// c + a.lo16 * b.byte0 + a.hi16 * b.byte1   (signed), and the same with bytes 2, 3
__device__ __forceinline__ int dp2a_lo_ss(uint32_t a, uint32_t b, int c)
{
  int d;
  asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ int dp2a_hi_ss(uint32_t a, uint32_t b, int c)
{
  int d;
  asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// 2*floor(log2(v+2)): get_ep_ex_golomb_bitcost (search_inter.c:235-253)
__device__ __forceinline__ uint32_t golomb_bits(uint32_t v) { return 2u * (31u - __clz(v + 2u)); }

__device__ __forceinline__ uint32_t mvd_bits(int dx, int dy) { return golomb_bits(abs(dx)) + golomb_bits(abs(dy)); }

// ---- thread group: a warp or the whole CTA ----
template <int G>
struct Group {
  int tid;        // 0..G-1 inside the group
  int id;         // group index inside the CTA
  int bar = 0;    // groups of several warps that are not the whole CTA: their named barrier (1..15); 0 = __syncthreads
  __device__ __forceinline__ void sync() const
  {
    if (G == 32) __syncwarp();
    else if (bar == 0) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(G) : "memory");
  }
};

__device__ __forceinline__ uint32_t pack_mv(int x, int y) { return ((uint32_t)x << 16) | ((uint32_t)y & 0xFFFFu); }

struct JobCtx {
  int x, y, w, h;            // PU: tile-relative origin, size
  int ref_idx;
  int16_t mv_cand[2][2];
  int lambda;
  // merge candidates that point into this reference with one vector (inter_merge_cand_t, dir != 3):
  // packed quarter-pel vector of candidate i, valid if bit i of mvalid
  uint32_t mkey[5];
  uint32_t mvalid;
};

__device__ __forceinline__ void make_ctx(JobCtx &j, const MeFrame &f, const kvz_cuda_me_pu &pu, const kvz_cuda_merge_cand *merge,
                                         const kvz_cuda_me_job &job)
{
  j.x = pu.x; j.y = pu.y; j.w = pu.w; j.h = pu.h;
  j.ref_idx = job.ref_idx;
  j.mv_cand[0][0] = job.mv_cand[0][0]; j.mv_cand[0][1] = job.mv_cand[0][1];
  j.mv_cand[1][0] = job.mv_cand[1][0]; j.mv_cand[1][1] = job.mv_cand[1][1];
  j.lambda = f.lambda_sqrt_i;
  j.mvalid = 0;
#pragma unroll
  for (int i = 0; i < 5; i++) {
    j.mkey[i] = 0;
    if (i < pu.num_merge_cand) {
      const kvz_cuda_merge_cand m = merge[i];
      if (m.dir != 3 && f.ref_LX[m.dir - 1][m.ref[m.dir - 1]] == job.ref_idx) {
        j.mkey[i] = pack_mv(m.mv[m.dir - 1][0], m.mv[m.dir - 1][1]);
        j.mvalid |= 1u << i;
      }
    }
  }
}

// fracmv_within_tile (search_inter.c:87-172); x,y quarter-pel
#ifndef KVZ_REFINE_WT
#define KVZ_REFINE_WT within_tile
#endif
__device__ __forceinline__ bool within_tile_inl(const MeFrame &f, const JobCtx &j, int x, int y)
{
  const bool frac_luma = (x % 4 != 0) || (y % 4 != 0);
  const bool frac_chroma = (x % 8 != 0) || (y % 8 != 0);
  if (f.owf && f.wpp) {
    int margin = frac_luma ? 4 : (frac_chroma ? 2 : 0);
    if (f.sao_type) margin += 10;               // SAO_DELAY_PX (global.h:175)
    else if (f.deblock_enable) margin += 8;     // DEBLOCK_DELAY_PX (global.h:163)
    const int lcu_x = j.x / 64, lcu_y = j.y / 64;
    const int dx = ((j.x + j.w + margin) * 4 + x) / (64 << 2) - lcu_x;
    const int dy = ((j.y + j.h + margin) * 4 + y) / (64 << 2) - lcu_y;
    if (dy > f.max_ref_lcu_down) return false;
    if (dx + dy > f.max_ref_lcu_down + f.max_ref_lcu_right) return false;
  }
  if (f.mv_constraint == 0) return true;        // KVZ_MV_CONSTRAIN_NONE
  int margin = 0;
  if (f.mv_constraint == 4) margin = frac_luma ? 16 : (frac_chroma ? 8 : 0);
  const int ax = j.x * 4 + x, ay = j.y * 4 + y;
  const int from_right = (f.tile_w << 2) - (ax + (j.w << 2));
  const int from_bottom = (f.tile_h << 2) - (ay + (j.h << 2));
  return ax >= margin && ay >= margin && from_right >= margin && from_bottom >= margin;
}

// out-of-line form for the integer scan, where the check is rare and inlining it costs registers
__device__ __noinline__ bool within_tile(const MeFrame &f, const JobCtx &j, int x, int y) { return within_tile_inl(f, j, x, y); }

// calc_mvd_cost (search_inter.c:373-412): bits for vector (x,y) << mv_shift
__device__ __forceinline__ uint32_t mvd_cost_bits(const JobCtx &j, int x, int y, int mv_shift)
{
  x *= 1 << mv_shift;
  y *= 1 << mv_shift;
  const uint32_t q = pack_mv(x, y);
  // the first merge candidate with this vector wins: its index is the bit count
  uint32_t hit = 0;
#pragma unroll
  for (int i = 4; i >= 0; i--)
    if (((j.mvalid >> i) & 1u) && j.mkey[i] == q) hit = (uint32_t)i | 0x100u;
  if (hit) return hit & 0xFFu;
  const uint32_t c0 = mvd_bits(x - j.mv_cand[0][0], y - j.mv_cand[0][1]);
  const uint32_t c1 = mvd_bits(x - j.mv_cand[1][0], y - j.mv_cand[1][1]);
  return min(c0, c1);
}

// SAD of the PU (rows `cur_pitch` apart) against the reference block whose top-left sample is
// (rx, ry) in frame coordinates: kvz_image_calc_sad (image.c:455-479).  The block origin is
// clamped into the replicated border, which reads the same samples as the per-pixel
// coordinate clamp as long as pad >= PU size.
__device__ __noinline__ uint32_t sad_on_demand(const MeFrame &f, const MeRef &r, const uint8_t *cur, int cur_pitch, int w, int h,
                                               int rx, int ry)
{
  rx = clampi(rx, -r.pad, f.width + r.pad - w);
  ry = clampi(ry, -r.pad, f.height + r.pad - h);
  const uint8_t *p = r.plane + (size_t)(ry + r.pad) * r.pitch + (rx + r.pad);
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const int sh = (int)(a & 3) * 8;
  const uint32_t *rw = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
  const int wpitch = r.pitch >> 2;
  uint32_t acc = 0;
  for (int yy = 0; yy < h; yy++) {
    const uint32_t *cw = reinterpret_cast<const uint32_t *>(cur + (size_t)yy * cur_pitch);
    uint32_t prev = __ldg(rw);
    for (int c = 0; c < (w >> 2); c++) {
      uint32_t v = prev;
      if (sh) {
        const uint32_t nxt = __ldg(rw + c + 1);
        v = __funnelshift_r(prev, nxt, sh);
        prev = nxt;
      } else if (c + 1 < (w >> 2)) {
        prev = __ldg(rw + c + 1);
      }
      acc = sad4(__ldg(cw + c), v, acc);
    }
    rw += wpitch;
  }
  return acc;
}

// the same over the rows row0, row0 + rstep, ... only (a lane group shares one candidate)
__device__ __noinline__ uint32_t sad_on_demand_rows(const MeFrame &f, const MeRef &r, const uint8_t *cur, int cur_pitch, int w, int h,
                                                    int rx, int ry, int row0, int rstep)
{
  rx = clampi(rx, -r.pad, f.width + r.pad - w);
  ry = clampi(ry, -r.pad, f.height + r.pad - h);
  const uint8_t *p = r.plane + (size_t)(ry + r.pad + row0) * r.pitch + (rx + r.pad);
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const int sh = (int)(a & 3) * 8;
  const uint32_t *rw = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
  const int wpitch = r.pitch >> 2;
  uint32_t acc = 0;
  for (int yy = row0; yy < h; yy += rstep) {
    const uint32_t *cw = reinterpret_cast<const uint32_t *>(cur + (size_t)yy * cur_pitch);
    uint32_t prev = __ldg(rw);
    for (int c = 0; c < (w >> 2); c++) {
      uint32_t v = prev;
      if (sh) {
        const uint32_t nxt = __ldg(rw + c + 1);
        v = __funnelshift_r(prev, nxt, sh);
        prev = nxt;
      } else if (c + 1 < (w >> 2)) {
        prev = __ldg(rw + c + 1);
      }
      acc = sad4(__ldg(cw + c), v, acc);
    }
    rw += wpitch * rstep;
  }
  return acc;
}

__device__ __forceinline__ unsigned long long warp_min(unsigned long long v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = __shfl_xor_sync(0xffffffffu, v, o);
    v = t < v ? t : v;
  }
  return v;
}

// per-window separable tables of the AMVP bit cost: bits = min_c(gx[c][col] + gy[c][row])
struct BitTables {
  alignas(16) uint2 gx[132];     // lambda * bits of column i for AMVP candidate 0 (.x) and 1 (.y)
  alignas(16) uint2 gy[132];     // same for row i
};

// ======================================================================================
// integer search: one warp per (PU, reference) job
// ======================================================================================
constexpr int kScanWarps = 4;
#ifndef KVZ_SCAN_DEPTH
#define KVZ_SCAN_DEPTH 2      // map rows in flight per lane (uint16 maps)
#endif
#ifndef KVZ_SCAN_MINB
#define KVZ_SCAN_MINB 6
#endif

constexpr int kScanRing = 8;             // 16-byte pieces in flight per lane of the box scan

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src)
{
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}

// development timers of the box scan (KVZ_CUDA_STATS): [0] table set-up, [1] seed, [2] stream, [3] scans timed
__device__ unsigned long long g_box_scan_ns[4];
__device__ int g_box_scan_timing;
__device__ int g_pu_use_seed = 1;       // KVZ_CUDA_PU_SEED=0 switches the early-rejection seed of me_pu_kernel off (experiments)
__device__ __forceinline__ unsigned long long gtime_ns()
{
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// per-warp scratch of the integer search
template <bool BOX> struct ScanSmemT;
template <> struct ScanSmemT<false> {         // per-window scan: static
  BitTables bt;
  int hit_pos[5];             // window position (row * K + col) of a merge-override entry
  int hit_bits[5];
  int n_hit;
};
template <> struct ScanSmemT<true> {          // box scan: carved out of dynamic shared memory, sized by the largest box
  uint4 *ring;                // [kScanRing][32] per-lane stream of map vectors (cp.async)
  // column table, piece-major: pieces 0,1 = candidate-0 bits of entries 0..3 / 4..7 of a vector, 2,3 = candidate 1,
  // so that lanes on consecutive vectors read consecutive 16-byte pieces (no bank conflicts)
  uint4 *gxv;                 // [4][nv]
  int nv;
  uint2 *gy;                  // lambda * bits of box row i for AMVP candidate 0 (.x) and 1 (.y)
  uint32_t *rowinfo;          // per box row: covered column run lo | hi << 9, bit 18 = several runs, bits 19.. = windows
  int *hit_pos;               // row << 16 | col of the box
  int *hit_bits;
  int &n_hit;
  __device__ __forceinline__ uint2 gx(int col) const
  {
    const uint32_t *pv = reinterpret_cast<const uint32_t *>(&gxv[((col >> 2) & 1) * nv + (col >> 3)]) + (col & 3);
    return make_uint2(pv[0], pv[2 * nv * 4]);
  }
  __device__ __forceinline__ void set_gx(int col, uint32_t a0, uint32_t a1) const
  {
    uint32_t *pv = reinterpret_cast<uint32_t *>(&gxv[((col >> 2) & 1) * nv + (col >> 3)]) + (col & 3);
    pv[0] = a0; pv[2 * nv * 4] = a1;
  }
};

// one search window of search_mv_full
struct WinCtx {
  const MeFrame *f;
  const JobCtx *j;
  const ScanSmemT<false> *sm;
  int cx, cy;               // centre (whole samples)
  int R, K;
  int n_hit;
  bool free_mv;             // no tile / WPP restriction applies
};

// exact cost of one candidate: tile restriction and merge override included; ~0u if excluded
__device__ __noinline__ uint32_t exact_cost(const WinCtx &w, int row, int col, uint32_t sad)
{
  const int mx = w.cx - w.R + col, my = w.cy - w.R + row;
  if (!w.free_mv && !within_tile(*w.f, *w.j, mx * 4, my * 4)) return 0xFFFFFFFFu;
  const uint2 bx = w.sm->bt.gx[col], by = w.sm->bt.gy[row];
  uint32_t mvd = min(bx.x + by.x, bx.y + by.y);
  const int pos = row * w.K + col;
  for (int q = 0; q < w.n_hit; q++)
    if (w.sm->hit_pos[q] == pos) mvd = (uint32_t)w.sm->hit_bits[q] * (uint32_t)w.j->lambda;
  return sad + mvd;
}

#ifndef KVZ_BOX_INLINE
#define KVZ_BOX_INLINE __forceinline__
#endif
// union scan of a box: what the rare paths need
struct BoxCtx {
  const MeFrame *f;
  const JobCtx *j;
  const ScanSmemT<true> *sm;
  const int *wcx, *wcy;
  int nwin, n_hit;
  int bx0, by0, R, K;
  bool free_mv;
};

// exact cost of the box entry (row, col): tile restriction and merge override included; ~0u if excluded
__device__ KVZ_BOX_INLINE uint32_t box_price(const BoxCtx &b, int row, int col, uint32_t sad)
{
  const int dx = b.bx0 + col, dy = b.by0 + row;
  if (!b.free_mv && !within_tile(*b.f, *b.j, dx * 4, dy * 4)) return 0xFFFFFFFFu;
  const uint2 gx = b.sm->gx(col), gy = b.sm->gy[row];
  uint32_t mvd = min(gx.x + gy.x, gx.y + gy.y);
  for (int q = 0; q < b.n_hit; q++)
    if (b.sm->hit_pos[q] == ((row << 16) | col)) mvd = (uint32_t)b.sm->hit_bits[q] * (uint32_t)b.j->lambda;
  return sad + mvd;
}

// (first covering window << 16) | position inside that window, of the box entry pos = row << 16 | col
__device__ KVZ_BOX_INLINE uint32_t box_key(const BoxCtx &b, uint32_t pos)
{
  const int dx = b.bx0 + (int)(pos & 0xFFFFu), dy = b.by0 + (int)(pos >> 16);
  for (int w = 0; w < b.nwin; w++)
    if (abs(dx - b.wcx[w]) <= b.R && abs(dy - b.wcy[w]) <= b.R)
      return ((uint32_t)w << 16) | (uint32_t)((dy - b.wcy[w] + b.R) * b.K + dx - b.wcx[w] + b.R);
  return 0xFFFFFFFFu;
}

// two box entries of equal cost: the one search_mv_full meets first
__device__ KVZ_BOX_INLINE uint32_t box_tie(const BoxCtx &b, uint32_t pos_a, uint32_t pos_b)
{
  return box_key(b, pos_b) < box_key(b, pos_a) ? pos_b : pos_a;
}

// The window centres search_mv_full visits, in order (search_inter.c:891-955): the zero vector, extra_mv >> 2 unless a
// merge candidate rounds to it (mv_in_merge, :260-273), then every uni-directional merge candidate >> 2 that is not zero.
__device__ __forceinline__ int build_windows(const kvz_cuda_me_pu &pu, const kvz_cuda_merge_cand *merge,
                                             const kvz_cuda_me_job &job, int (&wcx)[7], int (&wcy)[7])
{
  int nwin = 1;
  wcx[0] = 0; wcy[0] = 0;
  const int ex = job.extra_mv[0] >> 2, ey = job.extra_mv[1] >> 2;
  bool in_merge = false;
  for (int i = 0; i < pu.num_merge_cand; i++) {
    const kvz_cuda_merge_cand &m = merge[i];
    if (m.dir == 3) continue;
    if (((m.mv[m.dir - 1][0] + 2) >> 2) == ex && ((m.mv[m.dir - 1][1] + 2) >> 2) == ey) in_merge = true;
  }
  if (!in_merge) { wcx[nwin] = ex; wcy[nwin] = ey; nwin++; }
  for (int i = 0; i < pu.num_merge_cand; i++) {
    const kvz_cuda_merge_cand &m = merge[i];
    if (m.dir == 3) continue;
    const int mx = m.mv[m.dir - 1][0] >> 2, my = m.mv[m.dir - 1][1] >> 2;
    if (mx == 0 && my == 0) continue;
    wcx[nwin] = mx; wcy[nwin] = my; nwin++;
  }
  return nwin;
}

// The winner of an integer search, key = cost << 32 | window << 16 | position in the window (~0: no candidate),
// parked in the job's result record for the refinement
__device__ void park_winner(unsigned long long best, const MeFrame &f, const kvz_cuda_me_pu &pu, const kvz_cuda_merge_cand *merge,
                            const kvz_cuda_me_job &job, kvz_cuda_me_result *out)
{
  JobCtx j;
  make_ctx(j, f, pu, merge, job);
  const int R = f.range, K = 2 * R + 1;
  int wcx[7], wcy[7];
  build_windows(pu, merge, job, wcx, wcy);
  int best_mvx = 0, best_mvy = 0;            // quarter-pel (info->best_mv starts at 0,0)
  uint32_t best_cost = 0xFFFFFFFFu, best_bits = 0;
  if (best != ~0ull) {
    const int wi = (int)((best >> 16) & 0xFFFF), pos = (int)(best & 0xFFFF);
    best_mvx = (wcx[wi] - R + pos % K) * 4;
    best_mvy = (wcy[wi] - R + pos / K) * 4;
    best_cost = (uint32_t)(best >> 32);
    best_bits = mvd_cost_bits(j, best_mvx >> 2, best_mvy >> 2, 2);
  }
  kvz_cuda_me_result r;
  r.mv[0] = (int16_t)best_mvx; r.mv[1] = (int16_t)best_mvy;
  r.cost = best_cost; r.bitcost = best_bits;
  r.merged = 0; r.merge_idx = 0; r.mv_cand_idx = 0; r.took = 0;
  *out = r;
}

// ======================================================================================
// heuristic integer searches (hexagon_search search_inter.c:690-778, diamond_search :796-883,
// select_starting_point :282-307, early_terminate :415-460): one warp per (PU, reference) job.
// The probes of one step are independent candidates, so a lane group evaluates probe i (lane i holds its coordinates) and the step's
// outcome -- check_mv_cost's strict '<' over the probes in order -- is the (cost, index) minimum.
// ======================================================================================
struct HeurState {
  uint32_t best_cost, best_bits;
  int mvx, mvy;                     // whole samples
};

// Lane i < n HOLDS candidate i = (px, py); the warp is cut into groups of 32 / 8 = 4 lanes (8 lanes when n <= 4),
// group g evaluates candidate g with its lanes taking every 4th (8th) row of the PU.  Returns the index of the
// candidate that became the new best, or -1 if none beats the current best (all lanes return the same value).
__device__ int probe_set(const MeFrame &f, const MeRef &ref, const JobCtx &j, const uint8_t *cur, int fx0, int fy0,
                         bool free_mv, int lane, int n, int px, int py, HeurState &h)
{
  const int gl = n <= 4 ? 3 : 2, G = 1 << gl;                  // log2 of the lanes per candidate
  const int pr = lane >> gl, sub = lane & (G - 1);
  const int qx = __shfl_sync(0xffffffffu, px, pr), qy = __shfl_sync(0xffffffffu, py, pr);
  const bool live = pr < n && (free_mv || within_tile(f, j, qx * 4, qy * 4));
  uint32_t sad = live ? sad_on_demand_rows(f, ref, cur, f.cur_pitch, j.w, j.h, fx0 + qx, fy0 + qy, sub, G) : 0u;
  for (int o = G >> 1; o > 0; o >>= 1) sad += __shfl_xor_sync(0xffffffffu, sad, o);
  uint32_t cost = 0xFFFFFFFFu, bits = 0;
  if (live && sad < h.best_cost) {                             // check_mv_cost (search_inter.c:211)
    bits = mvd_cost_bits(j, qx, qy, 2);
    cost = sad + bits * (uint32_t)j.lambda;
  }
  unsigned long long key = ((unsigned long long)cost << 8) | (unsigned)pr;
  key = warp_min(key);
  const uint32_t wcost = (uint32_t)(key >> 8);
  const int win = (int)(key & 0xFFu);
  if (wcost >= h.best_cost) return -1;
  h.best_cost = wcost;
  h.best_bits = __shfl_sync(0xffffffffu, bits, win << gl);
  h.mvx = __shfl_sync(0xffffffffu, px, win);
  h.mvy = __shfl_sync(0xffffffffu, py, win);
  return win;
}

__device__ void heuristic_job(int lane, const MeFrame &f, const kvz_cuda_me_pu &pu, const kvz_cuda_merge_cand *merge,
                              const kvz_cuda_me_job &job, kvz_cuda_me_result *out)
{
  JobCtx j;
  make_ctx(j, f, pu, merge, job);
  const MeRef &ref = f.refs[job.ref_idx];
  const bool free_mv = !(f.owf && f.wpp) && f.mv_constraint == 0;
  const int fx0 = f.tile_x + j.x, fy0 = f.tile_y + j.y;
  const uint8_t *cur = f.cur + (size_t)fy0 * f.cur_pitch + fx0;
  HeurState h = { 0xFFFFFFFFu, 0u, 0, 0 };

  // ---- select_starting_point: zero vector, extra_mv, rounded uni-directional merge candidates ----
  {
    int n = 1, mx = 0, my = 0;                                 // lane 0: (0, 0)
    const int ex = job.extra_mv[0] >> 2, ey = job.extra_mv[1] >> 2;
    bool in_merge = false;
    for (int i = 0; i < pu.num_merge_cand; i++) {
      const kvz_cuda_merge_cand &m = merge[i];
      if (m.dir == 3) continue;
      in_merge |= ((m.mv[m.dir - 1][0] + 2) >> 2) == ex && ((m.mv[m.dir - 1][1] + 2) >> 2) == ey;
    }
    if ((ex != 0 || ey != 0) && !in_merge) { if (lane == n) { mx = ex; my = ey; } n++; }
    for (int i = 0; i < pu.num_merge_cand; i++) {
      const kvz_cuda_merge_cand &m = merge[i];
      if (m.dir == 3) continue;
      const int x = (m.mv[m.dir - 1][0] + 2) >> 2, y = (m.mv[m.dir - 1][1] + 2) >> 2;
      if (x == 0 && y == 0) continue;
      if (lane == n) { mx = x; my = y; }
      n++;
    }
    probe_set(f, ref, j, cur, fx0, fy0, free_mv, lane, n, mx, my, h);
  }

  bool done = false;
  // ---- early_terminate: two rounds of a 4- then 3-point cross ----
  if (f.me_early_termination) {
    const int ptx[7] = { 0, -1, 0, 1, 0, -1, 0 }, pty[7] = { -1, 0, 1, 0, -1, 0, 0 };
    int cx = h.mvx, cy = h.mvy, first = 0, last = 3;
    for (int k = 0; k < 2 && !done; k++) {
      const double threshold = f.me_early_termination == 2 ? (double)h.best_cost * 0.95 : (double)h.best_cost;
      const int idx = min(first + lane, 6);
      const int w = probe_set(f, ref, j, cur, fx0, fy0, free_mv, lane, last - first + 1, cx + ptx[idx], cy + pty[idx], h);
      const int best = w < 0 ? 6 : first + w;
      cx += ptx[best]; cy += pty[best];
      if ((double)h.best_cost >= threshold) done = true;
      first = (best + 3) % 4;
      last = first + 2;
    }
  }

  if (!done && f.ime_search == KVZ_CUDA_IME_HEXBS) {
    const int bx[9] = { 0, 1, 2, 1, -1, -2, -1, 1, 2 }, by[9] = { 0, -2, 0, 2, 2, 0, -2, -2, 0 };
    const int sx[9] = { 0, 0, -1, 1, 0, -1, 1, -1, 1 }, sy[9] = { 0, -1, 0, 0, 1, -1, -1, 1, 1 };
    uint32_t steps = (uint32_t)f.me_max_steps;
    int cx = h.mvx, cy = h.mvy;
    int li = min(lane + 1, 8);
    int w = probe_set(f, ref, j, cur, fx0, fy0, free_mv, lane, 6, cx + bx[li], cy + by[li], h);
    int best = w < 0 ? 0 : w + 1;
    while (best != 0 && steps != 0) {
      if (steps > 0) steps -= 1;
      const int start = best == 1 ? 6 : (best == 8 ? 1 : best - 1);
      cx += bx[best]; cy += by[best];
      li = min(start + lane, 8);
      w = probe_set(f, ref, j, cur, fx0, fy0, free_mv, lane, 3, cx + bx[li], cy + by[li], h);
      best = w < 0 ? 0 : start + w;
    }
    li = min(lane + 1, 8);
    probe_set(f, ref, j, cur, fx0, fy0, free_mv, lane, 8, cx + sx[li], cy + sy[li], h);
  } else if (!done && f.ime_search == KVZ_CUDA_IME_TZ) {
    // tz_search (search_inter.c:595-671) with its compiled-in parameters: range 96, 8-point diamond
    // patterns (4 points at distance 1, tz_pattern_search :463-577), no raster step, star refinement
    const int range = 96;
    int best_dist = 0;
    auto pattern = [&](int dist, int cx, int cy) {
      const int hd = dist / 2, k = lane & 7;
      const int ox = k == 1 ? dist : k == 3 ? -dist : k == 4 || k == 5 ? hd : k >= 6 ? -hd : 0;
      const int oy = k == 0 ? dist : k == 2 ? -dist : k == 4 || k == 7 ? hd : k == 5 || k == 6 ? -hd : 0;
      if (probe_set(f, ref, j, cur, fx0, fy0, free_mv, lane, dist == 1 ? 4 : 8, cx + ox, cy + oy, h) >= 0) best_dist = dist;
    };
    int sx = h.mvx, sy = h.mvy, idle = 0;
    for (int d = 1; d <= range; d *= 2) {
      pattern(d, sx, sy);
      if (best_dist != d) idle++;
      if (idle >= 3) break;
    }
    if (sx != 0 || sy != 0) {
      idle = 0;
      for (int d = 1; d <= range / 2; d *= 2) {
        pattern(d, 0, 0);
        if (best_dist != d) idle++;
        if (idle >= 3) break;
      }
    }
    while (best_dist > 0) {
      best_dist = 0;
      sx = h.mvx; sy = h.mvy;
      for (int d = 1; d <= range; d *= 2) pattern(d, sx, sy);
    }
  } else if (!done) {
    const int dx[5] = { 0, 1, 0, -1, 0 }, dy[5] = { -1, 0, 1, 0, 0 };
    uint32_t steps = (uint32_t)f.me_max_steps;
    int cx = h.mvx, cy = h.mvy;
    int li = min(lane, 4);
    int w = probe_set(f, ref, j, cur, fx0, fy0, free_mv, lane, 5, cx + dx[li], cy + dy[li], h);
    int best = w < 0 ? 4 : w;
    if (best != 4) {
      cx += dx[best]; cy += dy[best];
      int from = 4;
      bool better;
      do {
        better = false;
        if (steps > 0) steps -= 1;
        // the three points that are not where the search came from, in index order
        int k = lane, idx = -1;
        for (int i = 0; i < 4; i++) {
          if (i == from) continue;
          if (k == 0 && idx < 0) idx = i;
          k--;
        }
        const int n = from == 4 ? 4 : 3;
        if (idx < 0) idx = 4;
        w = probe_set(f, ref, j, cur, fx0, fy0, free_mv, lane, n, cx + dx[idx], cy + dy[idx], h);
        if (w >= 0) {
          best = __shfl_sync(0xffffffffu, idx, w);
          better = true;
          cx += dx[best]; cy += dy[best];
          from = best ^ 3;                                     // search_inter.c:874 (as the oracle restates it)
        }
      } while (better && steps != 0);
    }
  }

  if (lane == 0) {
    kvz_cuda_me_result r;
    r.mv[0] = (int16_t)(h.mvx * 4); r.mv[1] = (int16_t)(h.mvy * 4);
    r.cost = h.best_cost; r.bitcost = h.best_bits;
    r.merged = 0; r.merge_idx = 0; r.mv_cand_idx = 0; r.took = 0;
    *out = r;
  }
}

// returns false when the job belongs to the other instantiation (nothing written except the 0xFF mark of BOX)
// BOX only: `nsub` warps share one job, warp `sub` of them streams every nsub-th step of the box and leaves its best
// key in *part instead of parking a winner (the caller combines the parts, park_winner).
template <bool BOX>
__device__ bool scan_job(int lane, ScanSmemT<BOX> &sm, const MeFrame &f, const kvz_cuda_me_pu &pu, const kvz_cuda_merge_cand *merge,
                         const kvz_cuda_me_job &job, kvz_cuda_me_result *out, bool after_box, int sub = 0, int nsub = 1,
                         unsigned long long *part = nullptr)
{
  constexpr int G = 32;
  // the box instantiation ran first and marked the jobs it left for this one (took = 0xFF)
  if (!BOX && after_box && __shfl_sync(0xffffffffu, lane == 0 ? (int)out->took : 0, 0) != 0xFF) return false;
  JobCtx j;
  make_ctx(j, f, pu, merge, job);
  const MeRef &ref = f.refs[job.ref_idx];
  const int R = f.range, K = 2 * R + 1;
  const bool free_mv = !(f.owf && f.wpp) && f.mv_constraint == 0;
  const int fx0 = f.tile_x + j.x, fy0 = f.tile_y + j.y;       // frame-absolute PU origin
  const uint8_t *cur = f.cur + (size_t)fy0 * f.cur_pitch + fx0;

  // ---- window list of search_mv_full ----
  int wcx[7], wcy[7];
  const int nwin = build_windows(pu, merge, job, wcx, wcy);

  // level of the dense maps that holds this PU, if any
  int level = -1;
  if (j.w == j.h && (j.w == 8 || j.w == 16 || j.w == 32 || j.w == 64) && (j.x % j.w) == 0 && (j.y % j.h) == 0)
    level = j.w == 8 ? 0 : j.w == 16 ? 1 : j.w == 32 ? 2 : 3;

  unsigned long long best = ~0ull;

  // ---- a box map set that covers every window of this job: one pass over the union ----
  int bi = -1;
  if (level >= 0) {
    int lox = wcx[0], hix = wcx[0], loy = wcy[0], hiy = wcy[0];
    for (int w = 1; w < nwin; w++) {
      lox = min(lox, wcx[w]); hix = max(hix, wcx[w]);
      loy = min(loy, wcy[w]); hiy = max(hiy, wcy[w]);
    }
    for (int m = 0; m < ref.n_sets; m++) {
      const MeMapSet &ms = ref.sets[m];
      const void *mp = level == 0 ? (const void *)ms.sad8 : level == 1 ? (const void *)ms.sad16
                     : level == 2 ? (const void *)ms.sad32 : (const void *)ms.sad64;
      if (ms.n_dx > 0 && mp != nullptr && lox - R >= ms.cx && hix + R < ms.cx + ms.n_dx && loy - R >= ms.cy &&
          hiy + R < ms.cy + ms.n_dy) { bi = m; break; }
    }
  }
  if ((bi >= 0) != BOX) {                       // the other instantiation of the kernel owns this job
    if (BOX && lane == 0) out->took = 0xFF;
    return false;
  }
  if constexpr (BOX) {
    // Every displacement of the union of the windows is priced once.  A displacement belongs to the
    // FIRST window of the list that covers it (a later window meets it again with the same cost and
    // loses the tie), so the key of a candidate is (cost, that window, its position in that window).
    const MeMapSet &ms = ref.sets[bi];
    const bool timed = g_box_scan_timing && lane == 0 && sub == 0;
    unsigned long long tb0 = 0, tb1 = 0, tb2 = 0;
    if (timed) tb0 = gtime_ns();
    BoxCtx bc;
    bc.f = &f; bc.j = &j; bc.sm = &sm; bc.wcx = wcx; bc.wcy = wcy; bc.nwin = nwin;
    bc.bx0 = ms.cx; bc.by0 = ms.cy; bc.R = R; bc.K = K; bc.free_mv = free_mv;
    uint32_t wbest = 0xFFFFFFFFu, wpos = 0;          // best cost so far and its box entry (row << 16 | col)
    const int bx0 = ms.cx, by0 = ms.cy, n_dx = ms.n_dx, n_dy = ms.n_dy;
    bool close_x;
    // the bounding rectangle of the job's windows inside the box: rows r0 .. r0 + nr - 1, chunks c0 .. c1 (16 columns
    // each).  Only that part of the box is streamed and only its rows / columns get table entries.
    int r0, nr, c0, c1;
    {
      int lox = wcx[0], hix = wcx[0], loy = wcy[0], hiy = wcy[0];
      for (int w = 1; w < nwin; w++) {
        lox = min(lox, wcx[w]); hix = max(hix, wcx[w]);
        loy = min(loy, wcy[w]); hiy = max(hiy, wcy[w]);
      }
      close_x = hix - lox <= K;
      r0 = loy - R - by0; nr = hiy - loy + 2 * R + 1;
      c0 = (lox - R - bx0) >> 4; c1 = (hix + R - bx0) >> 4;
    }
    const int n = 8 >> level;
    const int ctu = (j.y >> 6) * f.ctus_x + (j.x >> 6);
    const size_t prow = (size_t)ctu * n * n + ((j.y & 63) / j.h) * n + (j.x & 63) / j.w;
    // the first map vectors are requested before the tables are built: their latency hides behind that work
    // Vectors of 8 entries in row-major order, 32 per step.  A lane streams its vectors through a ring in
    // shared memory with cp.async: kScanRing 16-byte pieces (8 vectors of uint16 entries, 4 of uint32) are
    // in flight per lane while one is priced -- the scan is bound by HBM latency, not by registers.
    // Box maps are CHUNK-MAJOR (ime_sad.cu): map[chunk][pu][row][16], the last chunk 8 wide when n_dx is an odd
    // multiple of 8.  Vector index i runs chunk by chunk, inside a chunk row-major (2 vectors per row, 1 in a half
    // chunk): the vectors of a chunk are contiguous in memory.  The traversal order is free -- ties are settled
    // by key, not by visiting order.
    const int nch = (n_dx + 15) >> 4, last_vpc = (n_dx & 8) ? 1 : 2;
    const int pus_level = f.ctus_x * ((f.tile_h + 63) >> 6) * n * n;
    const bool wide = level >= 2;                                   // uint32 entries: two pieces per vector
    const uint4 *mp = wide ? reinterpret_cast<const uint4 *>(level == 2 ? ms.sad32 : ms.sad64)
                           : reinterpret_cast<const uint4 *>(level == 0 ? ms.sad8 : ms.sad16);
    const int depth = wide ? kScanRing / 2 : kScanRing;             // vectors in flight
    const int vpc1 = (c1 == nch - 1) ? last_vpc : 2;                // vectors per row in chunk c1
    const int total = nr * (2 * (c1 - c0) + vpc1);
    // cursor over the vectors of this lane: chunk c, index ic inside the rows r0.. of the PU's run of that chunk,
    // off = position in the map (in vectors of 8 entries)
    struct Cursor { int c, ic; size_t off; };
    const size_t chunk_stride = (size_t)2 * pus_level * n_dy;      // vectors of a full chunk, all PUs
    const size_t pu_run = (size_t)prow * n_dy + r0;                // rows before the first streamed row of this PU, per chunk
    auto run_len = [&](int c) { return nr * ((c == nch - 1) ? last_vpc : 2); };
    auto run_off = [&](int c) { return (size_t)c * chunk_stride + pu_run * ((c == nch - 1) ? last_vpc : 2); };
    auto advance = [&](Cursor &k, int by) {
      k.ic += by; k.off += by;
      while (k.c < c1 && k.ic >= run_len(k.c)) {
        k.ic -= run_len(k.c); k.c++;
        k.off = run_off(k.c) + k.ic;
      }
    };
    Cursor ci = { c0, 0, run_off(c0) }, cc = ci;                    // issue / consume
    const int first = lane + G * sub, stride = G * nsub;              // this warp's vectors: first, first + stride, ...
    advance(ci, first); advance(cc, first);
    int issued = first;
    auto issue = [&](int slot) {
      if (issued < total) {
        const size_t v = ci.off;
        if (wide) {
          cp_async16(&sm.ring[(2 * slot) * 32 + lane], mp + 2 * v);
          cp_async16(&sm.ring[(2 * slot + 1) * 32 + lane], mp + 2 * v + 1);
        } else {
          cp_async16(&sm.ring[slot * 32 + lane], mp + v);
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      issued += stride; advance(ci, stride);
    };
    for (int t = 0; t < depth; t++) issue(t);
    __syncwarp();
    for (int i = 16 * c0 + lane; i < min(n_dx, 16 * (c1 + 1)); i += G) {
      const int qx = (bx0 + i) * 4;
      const uint32_t lam = (uint32_t)j.lambda;
      sm.set_gx(i, lam * golomb_bits(abs(qx - j.mv_cand[0][0])), lam * golomb_bits(abs(qx - j.mv_cand[1][0])));
    }
    for (int i = r0 + lane; i < r0 + nr; i += G) {
      const int qy = (by0 + i) * 4;
      const uint32_t lam = (uint32_t)j.lambda;
      sm.gy[i] = make_uint2(lam * golomb_bits(abs(qy - j.mv_cand[0][1])), lam * golomb_bits(abs(qy - j.mv_cand[1][1])));
      // columns of this row that some window covers: [lo, hi] when they form one run (the usual case:
      // window centres less than a window apart), else the per-window masks are used
      int lo = 0x1FF, hi = -1;
      uint32_t wm = 0;
      bool one_run = true;
      for (int w = 0; w < nwin; w++) {
        if (abs(by0 + i - wcy[w]) > R) continue;
        wm |= 1u << w;
        lo = min(lo, wcx[w] - R - bx0); hi = max(hi, wcx[w] + R - bx0);
        if (close_x) continue;                     // centres within one window width: any subset forms one run
        // otherwise a window that is not the rightmost one needs a neighbour starting inside or right after it
        bool rightmost = true, linked = false;
        for (int v = 0; v < nwin; v++) {
          if (abs(by0 + i - wcy[v]) > R || wcx[v] <= wcx[w]) continue;
          rightmost = false;
          linked |= wcx[v] - wcx[w] <= K;
        }
        one_run &= rightmost || linked;
      }
      sm.rowinfo[i] = hi < 0 ? 0xFFFFFFFFu : ((uint32_t)lo | ((uint32_t)hi << 9) | (one_run ? 0u : 1u << 18) | (wm << 19));
    }
    if (lane == 0) {
      int n = 0;
#pragma unroll
      for (int i = 0; i < 5; i++) {
        if (!((j.mvalid >> i) & 1u)) continue;
        const int mvx = (int16_t)(j.mkey[i] >> 16), mvy = (int16_t)(j.mkey[i] & 0xFFFFu);
        if ((mvx & 3) || (mvy & 3)) continue;
        const int col = (mvx >> 2) - bx0, row = (mvy >> 2) - by0;
        if (col < 0 || col >= n_dx || row < 0 || row >= n_dy) continue;
        const int pos = (row << 16) | col;
        bool dup = false;
        for (int q = 0; q < n; q++) dup |= sm.hit_pos[q] == pos;     // first matching candidate wins
        if (!dup) { sm.hit_pos[n] = pos; sm.hit_bits[n] = i; n++; }
      }
      sm.n_hit = n;
    }
    __syncwarp();
    const int n_hit = sm.n_hit;
    bc.n_hit = n_hit;
    if (timed) tb1 = gtime_ns();

    const int z_r0 = -R - by0, z_c0 = -R - bx0;     // the zero window (window 0) inside the box
    int hitv[5];                                    // vectors that hold a merge-override entry: row << 13 | vector
#pragma unroll
    for (int q = 0; q < 5; q++) hitv[q] = q < n_hit ? sm.hit_pos[q] >> 3 : -1;
    auto price = [&](int row, int vec, const uint32_t (&sad)[8]) {
      const int c0 = vec * 8;
      const uint32_t ri = sm.rowinfo[row];
      if (ri == 0xFFFFFFFFu) return;
      uint32_t mask;
      if (ri & (1u << 18)) {
        mask = 0;
        for (int w = 0; w < nwin; w++) {
          if (!((ri >> (19 + w)) & 1u)) continue;
          const int lo = wcx[w] - R - bx0 - c0, hi = lo + 2 * R;
          if (hi < 0 || lo > 7) continue;
          mask |= (0xFFu << max(lo, 0)) & (0xFFu >> (7 - min(hi, 7)));
        }
        mask &= 0xFFu;
      } else {
        const int lo = (int)(ri & 0x1FFu) - c0, hi = (int)((ri >> 9) & 0x1FFu) - c0;
        mask = (hi < 0 || lo > 7) ? 0u : (0xFFu << max(lo, 0)) & (0xFFu >> (7 - min(hi, 7))) & 0xFFu;
      }
      if (mask == 0) return;
      const int rv = (row << 13) | vec;
      const bool special = !free_mv || rv == hitv[0] || rv == hitv[1] || rv == hitv[2] || rv == hitv[3] || rv == hitv[4];
      // candidate (cost, entry): strict improvement is cheap, an equal cost is settled by the visiting order
      auto claim = [&](uint32_t cost, uint32_t pos) {
        if (cost < wbest) { wbest = cost; wpos = pos; }
        else if (cost == wbest && cost != 0xFFFFFFFFu) wpos = box_tie(bc, wpos, pos);
      };
      if (special) {
#pragma unroll
        for (int e = 0; e < 8; e++)
          if ((mask >> e) & 1u) claim(box_price(bc, row, c0 + e, sad[e]), ((uint32_t)row << 16) | (uint32_t)(c0 + e));
        return;
      }
      const uint2 gy = sm.gy[row];
      const int d = (int)gy.y - (int)gy.x;
      const uint4 a0l = sm.gxv[vec], a0h = sm.gxv[sm.nv + vec], a1l = sm.gxv[2 * sm.nv + vec], a1h = sm.gxv[3 * sm.nv + vec];
      const int a0[8] = { (int)a0l.x, (int)a0l.y, (int)a0l.z, (int)a0l.w, (int)a0h.x, (int)a0h.y, (int)a0h.z, (int)a0h.w };
      const int a1[8] = { (int)a1l.x, (int)a1l.y, (int)a1l.z, (int)a1l.w, (int)a1h.x, (int)a1h.y, (int)a1h.z, (int)a1h.w };
      int k[8];
      // min(a0 + b0, a1 + b1) - b0 = min(a0, a1 + d): signed, b0 is added back below
#pragma unroll
      for (int e = 0; e < 8; e++) k[e] = (int)sad[e] + min(a0[e], a1[e] + d);
      if (mask != 0xFFu) {
#pragma unroll
        for (int e = 0; e < 8; e++) k[e] = ((mask >> e) & 1u) ? k[e] : 0x0FFFFFFF;
      }
      const int m = min(min(min(k[0], k[1]), min(k[2], k[3])), min(min(k[4], k[5]), min(k[6], k[7])));
      const uint32_t cost = (uint32_t)(m + (int)gy.x);
      if (cost > wbest) return;
      uint32_t eq = 0;                                // entries of the vector that reach the minimum
#pragma unroll
      for (int e = 0; e < 8; e++) eq |= (uint32_t)(k[e] == m) << e;
      eq &= mask;
      // inside the zero window (or with a single window) the row-major order of the box is the visiting order
      const bool plain = nwin == 1 || ((unsigned)(row - z_r0) <= (unsigned)(2 * R) && c0 >= z_c0 && c0 + 7 <= z_c0 + 2 * R);
      if (cost < wbest && (plain || (eq & (eq - 1)) == 0)) {
        wbest = cost; wpos = ((uint32_t)row << 16) | (uint32_t)(c0 + __ffs((int)eq) - 1);
        return;
      }
      for (; eq; eq &= eq - 1) claim(cost, ((uint32_t)row << 16) | (uint32_t)(c0 + __ffs((int)eq) - 1));
    };

    // Early rejection: the exact cost of a few likely candidates (the window centres and the two AMVP
    // predictors, where they are searched) bounds the minimum from above; a vector whose smallest SAD
    // already exceeds that bound (bit costs are >= 0) can hold neither the winner nor a tie with it.
    uint32_t seed = 0xFFFFFFFFu;
    // (measured in the encode, 1080p, B200: with the seed 5.4 / 7.2 / 7.8 frames/s at --owf 0 / 2 / 4, without 5.2 / 6.1 / 7.1)
    if (part == nullptr || g_pu_use_seed) {
      int sx = 0, sy = 0;
      bool have = false;
      if (lane < nwin) { sx = wcx[lane]; sy = wcy[lane]; have = true; }
      else if (lane == 8 || lane == 9) { sx = (j.mv_cand[lane - 8][0] + 2) >> 2; sy = (j.mv_cand[lane - 8][1] + 2) >> 2; have = true; }
      if (have) {
        const int col = sx - bx0, row = sy - by0;
        const uint32_t pos = ((uint32_t)row << 16) | (uint32_t)col;
        if (col >= 0 && col < n_dx && row >= 0 && row < n_dy && box_key(bc, pos) != 0xFFFFFFFFu) {
          const int sc = col >> 4, sw = (sc == nch - 1 && last_vpc == 1) ? 8 : 16;
          const size_t at = (size_t)sc * 16 * pus_level * n_dy + (prow * (size_t)n_dy + row) * sw + (col - 16 * sc);
          const uint32_t sad = level < 2 ? (uint32_t)__ldg((level == 0 ? ms.sad8 : ms.sad16) + at) & 0xFFFFu
                                         : __ldg((level == 2 ? ms.sad32 : ms.sad64) + at);
          seed = box_price(bc, row, col, sad);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) seed = min(seed, __shfl_xor_sync(0xffffffffu, seed, o));
    }
    uint32_t bound = seed;                          // min(seed, wbest)
    if (timed) tb2 = gtime_ns();
    int slot = 0;
    for (int i = first; i < total; i += stride) {
      if (wide) asm volatile("cp.async.wait_group %0;" ::"n"(kScanRing / 2 - 1) : "memory");
      else asm volatile("cp.async.wait_group %0;" ::"n"(kScanRing - 1) : "memory");
      uint32_t sad[8];
      if (wide) {
        const uint4 a = sm.ring[(2 * slot) * 32 + lane], b = sm.ring[(2 * slot + 1) * 32 + lane];
        sad[0] = a.x; sad[1] = a.y; sad[2] = a.z; sad[3] = a.w; sad[4] = b.x; sad[5] = b.y; sad[6] = b.z; sad[7] = b.w;
      } else {
        const uint4 a = sm.ring[slot * 32 + lane];
        sad[0] = a.x & 0xFFFFu; sad[1] = a.x >> 16; sad[2] = a.y & 0xFFFFu; sad[3] = a.y >> 16;
        sad[4] = a.z & 0xFFFFu; sad[5] = a.z >> 16; sad[6] = a.w & 0xFFFFu; sad[7] = a.w >> 16;
      }
      issue(slot);                                                  // the slot just read is free again
      slot = slot + 1 == depth ? 0 : slot + 1;
      const uint32_t smin = min(min(min(sad[0], sad[1]), min(sad[2], sad[3])), min(min(sad[4], sad[5]), min(sad[6], sad[7])));
      if (smin <= bound) {
        const bool half = cc.c == nch - 1 && last_vpc == 1;
        price(r0 + (half ? cc.ic : cc.ic >> 1), 2 * cc.c + (half ? 0 : (cc.ic & 1)), sad);
        bound = min(bound, wbest);
      }
      advance(cc, stride);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (wbest != 0xFFFFFFFFu) best = ((unsigned long long)wbest << 32) | box_key(bc, wpos);
    if (timed) {
      const unsigned long long tb3 = gtime_ns();
      atomicAdd(&g_box_scan_ns[0], tb1 - tb0); atomicAdd(&g_box_scan_ns[1], tb2 - tb1);
      atomicAdd(&g_box_scan_ns[2], tb3 - tb2); atomicAdd(&g_box_scan_ns[3], 1ull);
    }
  } else {
  if constexpr (!BOX) {
  for (int wi = 0; wi < nwin; wi++) {
    const int cx = wcx[wi], cy = wcy[wi];
    // separable AMVP bit tables and the merge overrides of this window
    __syncwarp();
    for (int i = lane; i < K; i += G) {
      const int qx = (cx - R + i) * 4, qy = (cy - R + i) * 4;
      const uint32_t lam = (uint32_t)j.lambda;
      sm.bt.gx[i] = make_uint2(lam * golomb_bits(abs(qx - j.mv_cand[0][0])), lam * golomb_bits(abs(qx - j.mv_cand[1][0])));
      sm.bt.gy[i] = make_uint2(lam * golomb_bits(abs(qy - j.mv_cand[0][1])), lam * golomb_bits(abs(qy - j.mv_cand[1][1])));
    }
    if (lane == 0) {
      int n = 0;
#pragma unroll
      for (int i = 0; i < 5; i++) {
        if (!((j.mvalid >> i) & 1u)) continue;
        const int mvx = (int16_t)(j.mkey[i] >> 16), mvy = (int16_t)(j.mkey[i] & 0xFFFFu);
        if ((mvx & 3) || (mvy & 3)) continue;
        const int col = (mvx >> 2) - (cx - R), row = (mvy >> 2) - (cy - R);
        if (col < 0 || col >= K || row < 0 || row >= K) continue;
        const int pos = row * K + col;
        bool dup = false;
        for (int q = 0; q < n; q++) dup |= sm.hit_pos[q] == pos;     // first matching candidate wins
        if (!dup) { sm.hit_pos[n] = pos; sm.hit_bits[n] = i; n++; }
      }
      sm.n_hit = n;
    }
    __syncwarp();
    WinCtx wc;
    wc.f = &f; wc.j = &j; wc.sm = &sm; wc.cx = cx; wc.cy = cy; wc.R = R; wc.K = K; wc.n_hit = sm.n_hit; wc.free_mv = free_mv;
    const int n_hit = wc.n_hit;
    uint32_t wbest = 0xFFFFFFFFu;
    int wpos = 0;

    // a map set made for this window centre?  (duplicates of earlier windows need no skipping:
    // with equal cost the earlier window's key is smaller)
    int si = -1;
    if (level >= 0) {
      for (int m = 0; m < ref.n_sets; m++) {
        const MeMapSet &ms = ref.sets[m];
        const void *mp = level == 0 ? (const void *)ms.sad8 : level == 1 ? (const void *)ms.sad16
                       : level == 2 ? (const void *)ms.sad32 : (const void *)ms.sad64;
        if (ms.cx == cx && ms.cy == cy && mp != nullptr) { si = m; break; }
      }
    }
    if (si >= 0) {
      const MeMapSet &ms = ref.sets[si];
      // Window straight from the dense maps: rows 0..K-1 hold K-1 = 8*vpr columns as vectors of 8
      // entries, row K holds the last column.  A lane owns one vector column (8 window columns) and
      // walks down the rows, so it meets its candidates in visiting order and a strict '<' keeps the
      // first of equal costs.
      const int n = 8 >> level;
      const int ctu = (j.y >> 6) * f.ctus_x + (j.x >> 6);
      const size_t prow = (size_t)ctu * n * n + ((j.y & 63) / j.h) * n + (j.x & 63) / j.w;
      const int vpr = K >> 3;                              // 2, 4, 8 or 16
      const int lv = 31 - __clz(vpr);
      const int vec = lane & (vpr - 1), c0 = vec * 8;
      const int rstep = G >> lv;
      // rows of this lane's vector column that hold a merge-override position
      int hrow[5];
#pragma unroll
      for (int q = 0; q < 5; q++) {
        const int hp = q < n_hit ? sm.hit_pos[q] : -1;
        const int hr = hp / K, hc = hp - hr * K;
        hrow[q] = (hp >= 0 && hc < K - 1 && (hc >> 3) == vec) ? hr : -1;
      }
      // column terms of this lane's 8 columns, index in the vector folded into the low 3 bits:
      // cost << 3 | e  =  (sad << 3) + min(a0, a1 + d) + (b0 << 3),  d = (b1 - b0) << 3
      int a0[8], a1[8];
#pragma unroll
      for (int e = 0; e < 8; e++) {
        const uint2 bx = sm.bt.gx[c0 + e];
        a0[e] = (int)(bx.x << 3) | e;
        a1[e] = (int)(bx.y << 3) | e;
      }
      auto fast8 = [&](int row, const int (&s8)[8]) {       // s8 = sad << 3
        const uint2 by = sm.bt.gy[row];
        const int d = ((int)by.y - (int)by.x) * 8;
        int k[8];
#pragma unroll
        for (int e = 0; e < 8; e++) k[e] = s8[e] + min(a0[e], a1[e] + d);
        const int m = min(min(min(k[0], k[1]), min(k[2], k[3])), min(min(k[4], k[5]), min(k[6], k[7])));
        const uint32_t cost = (uint32_t)((m >> 3) + (int)by.x);
        if (cost < wbest) { wbest = cost; wpos = row * K + c0 + (m & 7); }
      };
      auto exact8 = [&](int row, const uint32_t (&sad)[8]) {
#pragma unroll
        for (int e = 0; e < 8; e++) {
          const uint32_t cost = exact_cost(wc, row, c0 + e, sad[e]);
          if (cost < wbest) { wbest = cost; wpos = row * K + c0 + e; }
        }
      };
      auto is_special = [&](int row) {
        return !free_mv || row == hrow[0] || row == hrow[1] || row == hrow[2] || row == hrow[3] || row == hrow[4];
      };
      if (level < 2) {
        // 8x8 / 16x16 PUs: uint16 maps, four rows in flight per lane
        const uint16_t *mp = (level == 0 ? ms.sad8 : ms.sad16) + prow * (K + 1) * f.map_pitch + c0;
        for (int r0 = lane >> lv; r0 < K; r0 += KVZ_SCAN_DEPTH * rstep) {
          uint4 v[KVZ_SCAN_DEPTH];
#pragma unroll
          for (int u = 0; u < KVZ_SCAN_DEPTH; u++) {
            const int row = r0 + u * rstep;
            v[u] = row < K ? __ldg(reinterpret_cast<const uint4 *>(mp + (size_t)row * f.map_pitch)) : make_uint4(0, 0, 0, 0);
          }
#pragma unroll
          for (int u = 0; u < KVZ_SCAN_DEPTH; u++) {
            const int row = r0 + u * rstep;
            if (row >= K) break;
            const uint32_t w4[4] = { v[u].x, v[u].y, v[u].z, v[u].w };
            if (is_special(row)) {
              uint32_t sad[8];
#pragma unroll
              for (int e = 0; e < 8; e++) sad[e] = (e & 1) ? w4[e >> 1] >> 16 : w4[e >> 1] & 0xFFFFu;
              exact8(row, sad);
            } else {
              int s8[8];
#pragma unroll
              for (int e = 0; e < 8; e++) s8[e] = (int)((e & 1) ? (w4[e >> 1] >> 13) & 0x7FFF8u : (w4[e >> 1] << 3) & 0x7FFF8u);
              fast8(row, s8);
            }
          }
        }
      } else {
        // 32x32 / 64x64 PUs: uint32 maps, two rows in flight per lane
        const uint32_t *mp = (level == 2 ? ms.sad32 : ms.sad64) + prow * (K + 1) * f.map_pitch + c0;
        for (int r0 = lane >> lv; r0 < K; r0 += 2 * rstep) {
          uint4 v[2][2];
#pragma unroll
          for (int u = 0; u < 2; u++) {
            const int row = r0 + u * rstep;
            const uint4 *q = reinterpret_cast<const uint4 *>(mp + (size_t)row * f.map_pitch);
            v[u][0] = row < K ? __ldg(q) : make_uint4(0, 0, 0, 0);
            v[u][1] = row < K ? __ldg(q + 1) : make_uint4(0, 0, 0, 0);
          }
#pragma unroll
          for (int u = 0; u < 2; u++) {
            const int row = r0 + u * rstep;
            if (row >= K) break;
            const uint32_t sad[8] = { v[u][0].x, v[u][0].y, v[u][0].z, v[u][0].w, v[u][1].x, v[u][1].y, v[u][1].z, v[u][1].w };
            if (is_special(row)) {
              exact8(row, sad);
            } else {
              int s8[8];
#pragma unroll
              for (int e = 0; e < 8; e++) s8[e] = (int)(sad[e] << 3);
              fast8(row, s8);
            }
          }
        }
      }
      // last column: row K of the map, entry = window row.  Within a lane the positions of the last
      // column are not all later than those of the vector column, so both the cost and the
      // position decide.
      {
        const uint2 bxl = sm.bt.gx[K - 1];
        bool hit_last = false;                      // a merge override inside the last column?
        for (int q = 0; q < n_hit; q++) hit_last |= (sm.hit_pos[q] % K) == K - 1;
        const size_t lrow = (prow * (K + 1) + K) * f.map_pitch;
        for (int row = lane; row < K; row += G) {
          const uint32_t sad = level < 2 ? (uint32_t)__ldg((level == 0 ? ms.sad8 : ms.sad16) + lrow + row)
                                         : __ldg((level == 2 ? ms.sad32 : ms.sad64) + lrow + row);
          uint32_t cost;
          if (free_mv && !hit_last) {
            const uint2 by = sm.bt.gy[row];
            cost = sad + min(bxl.x + by.x, bxl.y + by.y);
          } else {
            cost = exact_cost(wc, row, K - 1, sad);
          }
          const int pos = row * K + K - 1;
          if (cost < wbest || (cost == wbest && cost != 0xFFFFFFFFu && pos < wpos)) { wbest = cost; wpos = pos; }
        }
      }
    } else {
      // on demand, skipping what an earlier window of this list already covered
      for (int p = lane; p < K * K; p += G) {
        const int row = p / K, col = p % K;
        const int mx = cx - R + col, my = cy - R + row;
        bool covered = false;
        for (int q = 0; q < wi; q++)
          covered |= (mx >= wcx[q] - R && mx <= wcx[q] + R && my >= wcy[q] - R && my <= wcy[q] + R);
        if (covered) continue;
        if (!free_mv && !within_tile(f, j, mx * 4, my * 4)) continue;
        const uint32_t cost = exact_cost(wc, row, col, sad_on_demand(f, ref, cur, f.cur_pitch, j.w, j.h, fx0 + mx, fy0 + my));
        if (cost < wbest) { wbest = cost; wpos = p; }
      }
    }
    if (wbest != 0xFFFFFFFFu) {
      const unsigned long long key = ((unsigned long long)wbest << 32) | (unsigned)((wi << 16) | wpos);
      best = key < best ? key : best;
    }
  }
  }
  }
  best = warp_min(best);
  if (part != nullptr) {                       // one of several warps on this job: the caller combines and parks
    if (lane == 0) *part = best;
    return true;
  }
  // ---- winner of the integer search, parked in the job's result record ----
  if (lane == 0) park_winner(best, f, pu, merge, job, out);
  return true;
}


// bytes of dynamic shared memory one warp of the box scan needs for boxes up to n_dx x n_dy
__host__ __device__ inline size_t box_scan_warp_bytes(int n_dx, int n_dy)
{
  const size_t b = (size_t)kScanRing * 32 * 16 + (size_t)4 * (n_dx / 8) * 16 + (size_t)n_dy * 8 + (((size_t)n_dy * 4 + 15) & ~(size_t)15) + 48;
  return (b + 15) & ~(size_t)15;
}

// BOX: the jobs whose windows all lie inside a box map set (single pass over the union); !BOX: the others.
// box_dx, box_dy: the largest box sides among the frame's map sets (BOX only)
template <bool BOX>
__global__ void __launch_bounds__(kScanWarps * 32, KVZ_SCAN_MINB)
me_scan_kernel(const __grid_constant__ MeFrame f, const kvz_cuda_me_pu *__restrict__ pus,
               const kvz_cuda_me_job *__restrict__ jobs, kvz_cuda_me_result *__restrict__ results, int n_pus,
               int box_dx, int box_dy)
{
  extern __shared__ __align__(16) unsigned char scan_dyn[];
  __shared__ kvz_cuda_merge_cand merge[5];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const kvz_cuda_me_pu &pu = pus[blockIdx.x];
  if (threadIdx.x < 5) merge[threadIdx.x] = pu.merge_cand[threadIdx.x];
  __syncthreads();
  if constexpr (BOX) {
    unsigned char *base = scan_dyn + warp * box_scan_warp_bytes(box_dx, box_dy);
    uint4 *ring = reinterpret_cast<uint4 *>(base);
    uint4 *gxv = ring + kScanRing * 32;
    const int nv = box_dx / 8;
    uint2 *gy = reinterpret_cast<uint2 *>(gxv + 4 * nv);
    uint32_t *rowinfo = reinterpret_cast<uint32_t *>(gy + box_dy);
    int *hits = reinterpret_cast<int *>(reinterpret_cast<unsigned char *>(rowinfo) + (((size_t)box_dy * 4 + 15) & ~(size_t)15));
    ScanSmemT<true> sm = { ring, gxv, nv, gy, rowinfo, hits, hits + 5, hits[10] };
    for (int q = warp; q < pu.num_jobs; q += kScanWarps)
      scan_job<true>(lane, sm, f, pu, merge, jobs[pu.first_job + q], results + pu.first_job + q, false);
  } else {
    __shared__ ScanSmemT<false> sm[kScanWarps];
    for (int q = warp; q < pu.num_jobs; q += kScanWarps) {
      if (f.ime_search != KVZ_CUDA_IME_FULL) heuristic_job(lane, f, pu, merge, jobs[pu.first_job + q], results + pu.first_job + q);
      else scan_job<false>(lane, sm[warp], f, pu, merge, jobs[pu.first_job + q], results + pu.first_job + q, box_dx > 0);
    }
  }
}

// ======================================================================================
// refinement: fractional search / integer SATD, merge match, candidate choice, carry
// ======================================================================================
template <int G, int S>
struct RefineSmem {
  // pitches chosen so that the 8 lanes of a tile (8 consecutive rows) hit distinct banks
  static constexpr int kWinPitch = S + 20;                     // bytes
  static constexpr int kCurPitch = 72;                         // bytes
  alignas(16) uint8_t cur[S * kCurPitch];                      // PU rows
  alignas(16) uint8_t win[(S + 9) * kWinPitch + 16];           // source window round the integer vector
  alignas(16) uint32_t pp[3][(S + 7) * (S + 4)];               // horizontally filtered planes, vertical int16 pairs, pitch w + 4
  kvz_cuda_merge_cand merge[5];
  uint32_t cost4[4];
};

// Both outcomes of one job's refinement, computed before the *inter_cost of the references in front of it is known
// (me_pu_kernel refines the references of a PU side by side): what search_pu_inter_ref leaves when the gate
// `best_cost < *inter_cost` (search_inter.c:1241) lets search_frac run, and when it does not.
struct RefineBoth {
  int16_t int_mv[2];            // winner of the integer search
  uint32_t int_cost, int_bits;  // its SAD-based cost (the gate's left-hand side) and bits
  uint32_t satd_cost;           // gate closed: SATD at the integer vector + bits * lambda (search_inter.c:1244-1256)
  int16_t fme_mv[2];            // gate open: search_frac's result
  uint32_t fme_cost, fme_bits;
};

// SPEC: compute both outcomes into *both and return; the carry is resolved by refine_finish
template <int G, int S, bool SPEC = false>
__device__ void refine_job(const Group<G> &g, RefineSmem<G, S> &sm, const MeFrame &f, const kvz_cuda_me_pu &pu,
                           const kvz_cuda_me_job &job, uint32_t &inter_cost, kvz_cuda_me_result *out,
                           RefineBoth *both = nullptr)
{
  JobCtx j;
  make_ctx(j, f, pu, sm.merge, job);
  const MeRef &ref = f.refs[job.ref_idx];
  const bool free_mv = !(f.owf && f.wpp) && f.mv_constraint == 0;
  const int fx0 = f.tile_x + j.x, fy0 = f.tile_y + j.y;       // frame-absolute PU origin

  // winner of the integer search (me_scan_kernel)
  int best_mvx = out->mv[0], best_mvy = out->mv[1];
  uint32_t best_cost = out->cost, best_bits = out->bitcost;

  const bool do_fme = f.fme_level > 0 && (SPEC ? best_cost < 0xFFFFFFFFu : best_cost < inter_cost);
  const bool do_satd = !do_fme && best_cost < 0xFFFFFFFFu;
  if (SPEC && g.tid == 0) {
    both->int_mv[0] = (int16_t)best_mvx; both->int_mv[1] = (int16_t)best_mvy;
    both->int_cost = best_cost; both->int_bits = best_bits;
    both->satd_cost = 0xFFFFFFFFu; both->fme_cost = 0xFFFFFFFFu;
  }
  if (do_fme || do_satd) {
    // ---- source window round the integer vector: columns -4..w+3(+pad), rows -4..h+4 ----
    // The window origin is clamped into the replicated border: with pad >= w + 20 this reads the same
    // samples as the per-pixel clamp of kvz_get_extended_block (ipol-generic.c:761-781).
    const int imx = best_mvx >> 2, imy = best_mvy >> 2;
    constexpr int WP = RefineSmem<G, S>::kWinPitch, CP = RefineSmem<G, S>::kCurPitch;
    {
      const int ox = clampi(fx0 + imx - 4, -ref.pad, f.width + ref.pad - (j.w + 20));
      const int oy = clampi(fy0 + imy - 4, -ref.pad, f.height + ref.pad - (j.h + 9));
      const uint8_t *p0 = ref.plane + (size_t)(oy + ref.pad) * ref.pitch + (ox + ref.pad);
      const uintptr_t a0 = reinterpret_cast<uintptr_t>(p0);
      const int sh = (int)(a0 & 3) * 8;
      const uint32_t *rw = reinterpret_cast<const uint32_t *>(a0 & ~(uintptr_t)3);
      const int wpr = (j.w + 12) >> 2, wpitch = ref.pitch >> 2;      // words per window row (w + 8 used, 4 spare)
      g.sync();
      for (int i = g.tid; i < (j.h + 9) * wpr; i += G) {
        const int r = i / wpr, c = i - r * wpr;
        const uint32_t *q = rw + (size_t)r * wpitch + c;
        const uint32_t lo = __ldg(q), hi = __ldg(q + 1);
        *reinterpret_cast<uint32_t *>(&sm.win[r * WP + c * 4]) = __funnelshift_r(lo, hi, sh);
      }
      g.sync();
    }

    const int tiles_x = j.w >> 3, tiles = tiles_x * (j.h >> 3);
    const int lane8 = g.tid & 7;
    // tile index -> (candidate, tile row, tile column); shifts when the PU is a power of two on both axes
    const bool pow2 = (tiles_x & (tiles_x - 1)) == 0 && (tiles & (tiles - 1)) == 0;
    const int lg_tx = 31 - __clz(tiles_x), lg_t = 31 - __clz(tiles);
    auto split = [&](int v, int &k, int &ty, int &tx) {      // v = k * tiles + ty * tiles_x + tx
      if (pow2) { k = v >> lg_t; const int t = v & (tiles - 1); ty = t >> lg_tx; tx = t & (tiles_x - 1); }
      else { k = v / tiles; const int t = v - k * tiles; ty = t / tiles_x; tx = t - ty * tiles_x; }
    };

    // SATD at the integer position: kvz_satd_any_size (strategies-picture.h:62-100), 8x8 tiles
    if (g.tid < 4) sm.cost4[g.tid] = 0;
    g.sync();
    for (int u0 = 0; u0 < tiles * 8; u0 += G) {
      const int u = u0 + g.tid;
      const bool on = u < tiles * 8;
      int k0, ty, tx;
      split(on ? (u >> 3) : 0, k0, ty, tx);
      int d[8];
      const uint2 cw = *reinterpret_cast<const uint2 *>(sm.cur + (ty * 8 + lane8) * CP + tx * 8);
      const uint8_t *pr = sm.win + (ty * 8 + lane8 + 4) * WP + tx * 8 + 4;
      const uint32_t p0 = *reinterpret_cast<const uint32_t *>(pr), p1 = *reinterpret_cast<const uint32_t *>(pr + 4);
#pragma unroll
      for (int x = 0; x < 4; x++) {
        d[x] = (int)((cw.x >> (8 * x)) & 0xFF) - (int)((p0 >> (8 * x)) & 0xFF);
        d[x + 4] = (int)((cw.y >> (8 * x)) & 0xFF) - (int)((p1 >> (8 * x)) & 0xFF);
      }
      const uint32_t s = satd8_rows(d, lane8);
      if (on && lane8 == 0) atomicAdd(&sm.cost4[0], s);
    }
    g.sync();
    const uint32_t satd_int = sm.cost4[0];
    if (SPEC && g.tid == 0) both->satd_cost = satd_int + best_bits * (uint32_t)j.lambda;

    if (do_satd) {
      // search_inter.c:1244-1256
      best_cost = satd_int + best_bits * (uint32_t)j.lambda;
    } else {
      // ---- search_frac ----
      // Horizontally filtered planes are shared between the candidates that use them
      // (kvz_filter_hpel/qpel_blocks_*_luma, ipol-generic.c:192-658, do the same): plane (fx, sx)
      // holds H[r][x] = sum_t fir[fx][t] * win[r][x + sx + 1 + t] for rows r = 0..h+7 (source rows
      // -4..h+3) as int16, stored as vertical pairs PP[r][x] = (H[r][x], H[r+1][x]) so that the
      // vertical 8-tap filter is four 2-way dot products.  Three planes are alive at a time.
      const int PW = j.w + 4;                               // words per plane row (w + 4: conflict-free 16-byte rows)
      auto make_plane = [&](int slot, int fx, int sx) {
        const uint32_t t_lo = c_fir4[fx][0], t_hi = c_fir4[fx][1];
        const int o = sx + 1;                                // byte offset inside the aligned 12-byte run
        const int strips = j.w >> 2;
        uint32_t *dst = sm.pp[slot];
        for (int it = g.tid; it < (j.h + 7) * strips; it += G) {
          const int r = it / strips, x = (it - r * strips) * 4;
          int hv[2][4];
#pragma unroll
          for (int rr = 0; rr < 2; rr++) {
            const uint32_t *wr = reinterpret_cast<const uint32_t *>(&sm.win[(r + rr) * WP + x]);
            const uint32_t w0 = wr[0], w1 = wr[1], w2 = wr[2];
#pragma unroll
            for (int q = 0; q < 4; q++) {
              const uint32_t lo = __funnelshift_rc(w0, w1, 8 * (o + q)), hi = __funnelshift_rc(w1, w2, 8 * (o + q));
              hv[rr][q] = dp4a_us(hi, t_hi, dp4a_us(lo, t_lo, 0));
            }
          }
          uint4 o4;
          o4.x = __byte_perm((uint32_t)hv[0][0], (uint32_t)hv[1][0], 0x5410);
          o4.y = __byte_perm((uint32_t)hv[0][1], (uint32_t)hv[1][1], 0x5410);
          o4.z = __byte_perm((uint32_t)hv[0][2], (uint32_t)hv[1][2], 0x5410);
          o4.w = __byte_perm((uint32_t)hv[0][3], (uint32_t)hv[1][3], 0x5410);
          *reinterpret_cast<uint4 *>(&dst[r * PW + x]) = o4;
        }
      };
      int slot_l = 0, slot_r = 1, slot_0 = 2, slot_a = -1, slot_b = -1;   // planes (2,-1), (2,0), (0,0), px=-1, px=+1
      make_plane(slot_l, 2, -1);
      make_plane(slot_r, 2, 0);
      make_plane(slot_0, 0, 0);

      const int sqx[9] = { 0, -1, 1, 0, 0, -1, 1, -1, 1 };
      const int sqy[9] = { 0, 0, 0, -1, 1, -1, -1, 1, 1 };
      int mvx = imx, mvy = imy;
      uint32_t bits0 = mvd_cost_bits(j, mvx, mvy, 2);
      uint32_t fbest_cost = satd_int + bits0 * (uint32_t)j.lambda;
      uint32_t fbest_bits = bits0;
      int best_index = 0, hx = 0, hy = 0, i = 1;
      mvx *= 2; mvy *= 2;
      for (int step = 0; step < f.fme_level; step++) {
        const int mv_shift = step < 2 ? 1 : 0;
        if (step == 2) {
          // quarter-pel planes round the best half-pel position; they take the slots of the half-pel
          // planes that the remaining candidates do not read
          g.sync();
          if (hx < 0) { slot_a = slot_r; slot_b = slot_0; }
          else if (hx > 0) { slot_a = slot_l; slot_b = slot_0; }
          else { slot_a = slot_l; slot_b = slot_r; }
          make_plane(slot_a, (2 * hx - 1) & 3, (2 * hx - 1) >> 2);
          make_plane(slot_b, (2 * hx + 1) & 3, (2 * hx + 1) >> 2);
        }
        // quarter-pel displacement of the 4 candidates relative to the integer vector
        int qx[4], qy[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int px = sqx[(step & 1) ? 5 + k : 1 + k], py = sqy[(step & 1) ? 5 + k : 1 + k];
          if (step < 2) { qx[k] = 2 * px; qy[k] = 2 * py; }
          else { qx[k] = 2 * hx + px; qy[k] = 2 * hy + py; }
        }
        g.sync();
        if (g.tid < 4) sm.cost4[g.tid] = 0;
        g.sync();
        // vertical pass + SATD: one 8x8 tile row per lane, 8 lanes per tile
        const int units = 4 * tiles * 8;
        for (int u0 = 0; u0 < units; u0 += G) {
          const int u = u0 + g.tid;
          const bool on = u < units;
          int k, ty, tx;
          split(on ? (u >> 3) : 0, k, ty, tx);
          const int cqx = qx[k], cqy = qy[k];
          const int slot = (cqx & 1) ? (cqx == 2 * hx - 1 ? slot_a : slot_b) : (cqx < 0 ? slot_l : cqx == 0 ? slot_0 : slot_r);
          const int fy = cqy & 3, sy = cqy >> 2;
          const uint32_t t_lo = c_fir4[fy][0], t_hi = c_fir4[fy][1];
          const uint32_t *pp = sm.pp[slot] + (ty * 8 + lane8 + sy + 1) * PW + tx * 8;
          int acc[8];
#pragma unroll
          for (int x = 0; x < 8; x++) acc[x] = 0;
#pragma unroll
          for (int i2 = 0; i2 < 4; i2++) {
            const uint4 va = *reinterpret_cast<const uint4 *>(pp + 2 * i2 * PW);
            const uint4 vb = *reinterpret_cast<const uint4 *>(pp + 2 * i2 * PW + 4);
            const uint32_t wv[8] = { va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w };
            const uint32_t tp = i2 < 2 ? t_lo : t_hi;
#pragma unroll
            for (int x = 0; x < 8; x++)
              acc[x] = (i2 & 1) ? dp2a_hi_ss(wv[x], tp, acc[x]) : dp2a_lo_ss(wv[x], tp, acc[x]);
          }
          int d[8];
          const uint2 cw = *reinterpret_cast<const uint2 *>(sm.cur + (ty * 8 + lane8) * CP + tx * 8);
#pragma unroll
          for (int x = 0; x < 8; x++) {
            const int v16 = (int)(int16_t)(acc[x] >> 6);       // int16 store of ipol-generic.c:285 et al.
            const int pix = clampi((v16 + 32) >> 6, 0, 255);
            d[x] = (int)(((x < 4 ? cw.x : cw.y) >> (8 * (x & 3))) & 0xFF) - pix;
          }
          const uint32_t s = satd8_rows(d, lane8);
          if (on && lane8 == 0) atomicAdd(&sm.cost4[k], s);
        }
        g.sync();
        // lane k of every warp prices candidate k, then the four lanes are read by all
        uint32_t costs[4], bits[4];
        bool ok[4];
        {
          const int k = g.tid & 3;
          const int cxk = mvx + sqx[i + k], cyk = mvy + sqy[i + k];
          const bool okk = free_mv || KVZ_REFINE_WT(f, j, cxk * (1 << mv_shift), cyk * (1 << mv_shift));
          uint32_t ck = sm.cost4[k], bk = 0;
          if (okk) {
            bk = mvd_cost_bits(j, cxk, cyk, mv_shift);
            ck += bk * (uint32_t)j.lambda;
          }
#pragma unroll
          for (int q = 0; q < 4; q++) {
            costs[q] = __shfl_sync(0xffffffffu, ck, q);
            bits[q] = __shfl_sync(0xffffffffu, bk, q);
            ok[q] = __shfl_sync(0xffffffffu, (int)okk, q) != 0;
          }
        }
#pragma unroll
        for (int k = 0; k < 4; k++)
          if (ok[k] && costs[k] < fbest_cost) { fbest_cost = costs[k]; fbest_bits = bits[k]; best_index = i + k; }
        i += 4;
        if (step == 1 || step == f.fme_level - 1) {
          mvx += sqx[best_index]; mvy += sqy[best_index];
          if (step == min(f.fme_level - 1, 1)) {
            mvx *= 2; mvy *= 2;
            hx = sqx[best_index]; hy = sqy[best_index];
            best_index = 0;
            i = 1;
          }
        }
      }
      best_mvx = mvx; best_mvy = mvy;
      best_cost = fbest_cost; best_bits = fbest_bits;
    }
  }
  if (SPEC) {
    if (g.tid == 0 && do_fme) {
      both->fme_mv[0] = (int16_t)best_mvx; both->fme_mv[1] = (int16_t)best_mvy;
      both->fme_cost = best_cost; both->fme_bits = best_bits;
    }
    return;
  }

  // ---- merge match, AMVP candidate choice, carry of *inter_cost (search_inter.c:1258-1297) ----
  int merged = 0, merge_idx = 0;
  {
    const uint32_t q = pack_mv(best_mvx, best_mvy);
    for (merge_idx = 0; merge_idx < pu.num_merge_cand; merge_idx++)
      if (((j.mvalid >> merge_idx) & 1u) && j.mkey[merge_idx] == q) { merged = 1; break; }
  }
  int cu_mv_cand = 0;
  if (!merged) {
    const bool same = j.mv_cand[0][0] == j.mv_cand[1][0] && j.mv_cand[0][1] == j.mv_cand[1][1];
    if (!same) {
      const uint32_t c0 = mvd_bits(best_mvx - j.mv_cand[0][0], best_mvy - j.mv_cand[0][1]);
      const uint32_t c1 = mvd_bits(best_mvx - j.mv_cand[1][0], best_mvy - j.mv_cand[1][1]);
      cu_mv_cand = c1 < c0 ? 1 : 0;
    }
  }
  const bool took = best_cost < inter_cost;
  if (took) inter_cost = best_cost;
  g.sync();                                   // everyone has read the parked winner
  if (g.tid == 0) {
    kvz_cuda_me_result r;
    r.mv[0] = (int16_t)best_mvx; r.mv[1] = (int16_t)best_mvy;
    r.cost = best_cost; r.bitcost = best_bits;
    r.merged = (uint8_t)merged; r.merge_idx = (uint8_t)merge_idx;
    r.mv_cand_idx = (uint8_t)cu_mv_cand; r.took = took ? 1 : 0;
    *out = r;
  }
}

// One thread: which of a job's two outcomes the carry selects, then the merge match and the AMVP candidate choice
// (search_inter.c:1241-1297)
__device__ void refine_finish(const MeFrame &f, const kvz_cuda_me_pu &pu, const kvz_cuda_merge_cand *merge,
                              const kvz_cuda_me_job &job, const RefineBoth &b, uint32_t &inter_cost, kvz_cuda_me_result *out)
{
  JobCtx j;
  make_ctx(j, f, pu, merge, job);
  int mvx = b.int_mv[0], mvy = b.int_mv[1];
  uint32_t cost = b.int_cost, bits = b.int_bits;
  if (f.fme_level > 0 && b.int_cost < inter_cost) { mvx = b.fme_mv[0]; mvy = b.fme_mv[1]; cost = b.fme_cost; bits = b.fme_bits; }
  else if (b.int_cost < 0xFFFFFFFFu) cost = b.satd_cost;
  int merged = 0, merge_idx = 0;
  {
    const uint32_t q = pack_mv(mvx, mvy);
    for (merge_idx = 0; merge_idx < pu.num_merge_cand; merge_idx++)
      if (((j.mvalid >> merge_idx) & 1u) && j.mkey[merge_idx] == q) { merged = 1; break; }
  }
  int cu_mv_cand = 0;
  if (!merged) {
    const bool same = j.mv_cand[0][0] == j.mv_cand[1][0] && j.mv_cand[0][1] == j.mv_cand[1][1];
    if (!same) {
      const uint32_t c0 = mvd_bits(mvx - j.mv_cand[0][0], mvy - j.mv_cand[0][1]);
      const uint32_t c1 = mvd_bits(mvx - j.mv_cand[1][0], mvy - j.mv_cand[1][1]);
      cu_mv_cand = c1 < c0 ? 1 : 0;
    }
  }
  const bool took = cost < inter_cost;
  if (took) inter_cost = cost;
  kvz_cuda_me_result r;
  r.mv[0] = (int16_t)mvx; r.mv[1] = (int16_t)mvy;
  r.cost = cost; r.bitcost = bits;
  r.merged = (uint8_t)merged; r.merge_idx = (uint8_t)merge_idx;
  r.mv_cand_idx = (uint8_t)cu_mv_cand; r.took = took ? 1 : 0;
  *out = r;
}

template <int G, int S, int T>
__global__ void __launch_bounds__(T)
me_refine_kernel(const __grid_constant__ MeFrame f, const kvz_cuda_me_pu *__restrict__ pus,
                 const kvz_cuda_me_job *__restrict__ jobs, kvz_cuda_me_result *results, int n_pus)
{
  constexpr int kGroups = T / G;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  RefineSmem<G, S> *all = reinterpret_cast<RefineSmem<G, S> *>(smem_raw);
  Group<G> g;
  g.tid = threadIdx.x % G;
  g.id = threadIdx.x / G;
  RefineSmem<G, S> &sm = all[g.id];
  const int pi = blockIdx.x * kGroups + g.id;
  if (pi >= n_pus) return;
  const kvz_cuda_me_pu &pu = pus[pi];
  const bool small = pu.w <= 16 && pu.h <= 16;
  if (small != (S == 16)) return;            // the other instantiation owns this PU
  if ((pu.w | pu.h) & 7) {                   // sides 4 / 12 (SMP / AMP): the 4-sample SATD strips are not implemented
    for (int q = g.tid; q < pu.num_jobs; q += G) {
      kvz_cuda_me_result r = {};
      r.cost = 0xFFFFFFFFu; r.took = 0xFE;
      results[pu.first_job + q] = r;
    }
    return;
  }

  // stage the PU's source block (pitch 64) and its merge candidates
  for (int i = g.tid; i < pu.h * (pu.w >> 2); i += G) {
    const int r = i / (pu.w >> 2), c = i % (pu.w >> 2);
    const uint8_t *src = f.cur + (size_t)(f.tile_y + pu.y + r) * f.cur_pitch + f.tile_x + pu.x + c * 4;
    *reinterpret_cast<uint32_t *>(&sm.cur[r * RefineSmem<G, S>::kCurPitch + c * 4]) = *reinterpret_cast<const uint32_t *>(src);
  }
  if (g.tid < 5) sm.merge[g.tid] = pu.merge_cand[g.tid];
  g.sync();
  uint32_t inter_cost = 0x7FFFFFFFu;         // *inter_cost = MAX_INT (search_inter.c:1460)
  for (int q = 0; q < pu.num_jobs; q++)
    refine_job<G, S>(g, sm, f, pu, jobs[pu.first_job + q], inter_cost, results + pu.first_job + q);
}


// ======================================================================================
// one PU per CTA, one launch: the form an encoder host calls while it walks its CTUs
// (kvz_cuda_enc_search_pu, enc_session.cu).  Phase 1: the jobs whose windows all lie inside a box map
// set, one warp each (scan_job<true>).  Phase 2: every other job, by the whole CTA: a window that lies
// inside a box is read from the maps, any other window is evaluated from a copy of the reference region
// in shared memory.  Phase 3: the refinement of the PU's jobs in order (refine_job).
// ======================================================================================
constexpr int kPuThreads = 512;

// KVZ_CUDA_STATS: nanoseconds (globaltimer) the CTAs of me_pu_kernel spent per phase, by PU size class
// [size 8/16/32/64][0 box scan, 1 CTA scan, 2 refinement, 3 calls]
__device__ unsigned long long g_pu_phase_ns[4][4];
__device__ __forceinline__ unsigned long long gtime()
{
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

struct CoopSmem {
  ScanSmemT<false> tab;                       // bit tables and merge overrides of the current window
  unsigned long long wmin[kPuThreads / 32];
  alignas(16) uint8_t cur[64 * 64];           // the PU's source rows, w bytes apart
  alignas(16) uint32_t region[1];             // (2R + h) rows of `pitch_w` words: the reference round the window
};

__host__ __device__ inline int coop_region_pitch_words(int R, int w) { return (((2 * R + w + 3 + 3) >> 2) + 1) | 1; }
__host__ __device__ inline size_t coop_smem_bytes(int R)
{
  return sizeof(CoopSmem) + (size_t)(2 * R + 64) * coop_region_pitch_words(R, 64) * 4;
}

// search_mv_full (search_inter.c:886-956) for one job by the whole CTA; the winner is parked in *out like scan_job does
__device__ void coop_scan_job(CoopSmem &sm, const MeFrame &f, const kvz_cuda_me_pu &pu, const kvz_cuda_merge_cand *merge,
                              const kvz_cuda_me_job &job, kvz_cuda_me_result *out)
{
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  JobCtx j;
  make_ctx(j, f, pu, merge, job);
  const MeRef &ref = f.refs[job.ref_idx];
  const int R = f.range, K = 2 * R + 1;
  const bool free_mv = !(f.owf && f.wpp) && f.mv_constraint == 0;
  const int fx0 = f.tile_x + j.x, fy0 = f.tile_y + j.y;

  int wcx[7], wcy[7];
  const int nwin = build_windows(pu, merge, job, wcx, wcy);
  int level = -1;
  if (j.w == j.h && (j.w == 8 || j.w == 16 || j.w == 32 || j.w == 64) && (j.x % j.w) == 0 && (j.y % j.h) == 0)
    level = j.w == 8 ? 0 : j.w == 16 ? 1 : j.w == 32 ? 2 : 3;

  // the PU's source block
  __syncthreads();
  for (int i = tid; i < j.h * (j.w >> 2); i += kPuThreads) {
    const int r = i / (j.w >> 2), c = i - r * (j.w >> 2);
    reinterpret_cast<uint32_t *>(sm.cur)[i] =
        *reinterpret_cast<const uint32_t *>(f.cur + (size_t)(fy0 + r) * f.cur_pitch + fx0 + c * 4);
  }

  unsigned long long best = ~0ull;
  for (int wi = 0; wi < nwin; wi++) {
    const int cx = wcx[wi], cy = wcy[wi];
    // a window wholly covered by earlier ones adds nothing (equal costs lose the tie to the earlier window)
    bool dup = false;
    for (int q = 0; q < wi; q++) dup |= wcx[q] == cx && wcy[q] == cy;
    if (dup) continue;
    __syncthreads();
    for (int i = tid; i < K; i += kPuThreads) {
      const int qx = (cx - R + i) * 4, qy = (cy - R + i) * 4;
      const uint32_t lam = (uint32_t)j.lambda;
      sm.tab.bt.gx[i] = make_uint2(lam * golomb_bits(abs(qx - j.mv_cand[0][0])), lam * golomb_bits(abs(qx - j.mv_cand[1][0])));
      sm.tab.bt.gy[i] = make_uint2(lam * golomb_bits(abs(qy - j.mv_cand[0][1])), lam * golomb_bits(abs(qy - j.mv_cand[1][1])));
    }
    if (tid == 0) {
      int n = 0;
#pragma unroll
      for (int i = 0; i < 5; i++) {
        if (!((j.mvalid >> i) & 1u)) continue;
        const int mvx = (int16_t)(j.mkey[i] >> 16), mvy = (int16_t)(j.mkey[i] & 0xFFFFu);
        if ((mvx & 3) || (mvy & 3)) continue;
        const int col = (mvx >> 2) - (cx - R), row = (mvy >> 2) - (cy - R);
        if (col < 0 || col >= K || row < 0 || row >= K) continue;
        const int pos = row * K + col;
        bool seen = false;
        for (int q = 0; q < n; q++) seen |= sm.tab.hit_pos[q] == pos;
        if (!seen) { sm.tab.hit_pos[n] = pos; sm.tab.hit_bits[n] = i; n++; }
      }
      sm.tab.n_hit = n;
    }
    // a box map set that holds the whole window?
    int bi = -1;
    if (level >= 0) {
      for (int m = 0; m < ref.n_sets; m++) {
        const MeMapSet &ms = ref.sets[m];
        const void *mp = level == 0 ? (const void *)ms.sad8 : level == 1 ? (const void *)ms.sad16
                       : level == 2 ? (const void *)ms.sad32 : (const void *)ms.sad64;
        if (ms.n_dx > 0 && mp != nullptr && cx - R >= ms.cx && cx + R < ms.cx + ms.n_dx && cy - R >= ms.cy &&
            cy + R < ms.cy + ms.n_dy) { bi = m; break; }
      }
    }
    int off = 0, pitch_w = 0;
    if (bi < 0) {
      // ---- copy of the reference round the window: rows cy-R .. cy+R+h-1, columns cx-R .. cx+R+w-1 of the PU's
      // position, every coordinate clamped into the bordered plane (== the per-sample clamp of image.c:320-452);
      // the copy starts at a 4-byte boundary of the plane, `off` columns to the left ----
      const int rx0 = fx0 + cx - R + ref.pad, ry0 = fy0 + cy - R + ref.pad;      // plane coordinates, may be outside
      off = rx0 & 3;
      const int ax0 = rx0 - off;
      pitch_w = coop_region_pitch_words(R, j.w);
      const int rows = 2 * R + j.h, words = (2 * R + j.w + off + 3) >> 2;
      const int pw = f.width + 2 * ref.pad, ph = f.height + 2 * ref.pad;
      const bool inside_x = ax0 >= 0 && ax0 + words * 4 <= pw;
      for (int i = tid; i < rows * words; i += kPuThreads) {
        const int r = i / words, c = i - r * words;
        const int py = clampi(ry0 + r, 0, ph - 1);
        const uint8_t *rowp = ref.plane + (size_t)py * ref.pitch;
        uint32_t v;
        if (inside_x) {
          v = __ldg(reinterpret_cast<const uint32_t *>(rowp + ax0 + c * 4));
        } else {
          v = 0;
#pragma unroll
          for (int b = 0; b < 4; b++) v |= (uint32_t)rowp[clampi(ax0 + c * 4 + b, 0, pw - 1)] << (8 * b);
        }
        sm.region[r * pitch_w + c] = v;
      }
    }
    __syncthreads();
    WinCtx wc;
    wc.f = &f; wc.j = &j; wc.sm = &sm.tab; wc.cx = cx; wc.cy = cy; wc.R = R; wc.K = K; wc.n_hit = sm.tab.n_hit; wc.free_mv = free_mv;
    auto consider = [&](int row, int col, uint32_t sad) {
      const int mx = cx - R + col, my = cy - R + row;
      for (int q = 0; q < wi; q++)
        if (mx >= wcx[q] - R && mx <= wcx[q] + R && my >= wcy[q] - R && my <= wcy[q] + R) return;
      const uint32_t cost = exact_cost(wc, row, col, sad);
      if (cost == 0xFFFFFFFFu) return;
      const unsigned long long key = ((unsigned long long)cost << 32) | (unsigned)((wi << 16) | (row * K + col));
      best = key < best ? key : best;
    };
    if (bi >= 0) {
      const MeMapSet &ms = ref.sets[bi];
      const int n = 8 >> level;
      const int ctu = (j.y >> 6) * f.ctus_x + (j.x >> 6);
      const size_t prow = (size_t)ctu * n * n + ((j.y & 63) / j.h) * n + (j.x & 63) / j.w;
      const size_t pus_level = (size_t)f.ctus_x * ((f.tile_h + 63) >> 6) * n * n;
      const int nch = (ms.n_dx + 15) >> 4;
      const bool half_last = (ms.n_dx & 8) != 0;
      for (int p = tid; p < K * K; p += kPuThreads) {
        const int row = p / K, col = p - row * K;
        const int bx = cx - R + col - ms.cx, by = cy - R + row - ms.cy;      // entry of the box
        const int c = bx >> 4, wc_ = (c == nch - 1 && half_last) ? 8 : 16;
        const size_t at = (size_t)16 * c * pus_level * ms.n_dy + (prow * ms.n_dy + by) * wc_ + (bx - 16 * c);
        const uint32_t sad = level < 2 ? (uint32_t)__ldg((level == 0 ? ms.sad8 : ms.sad16) + at)
                                       : __ldg((level == 2 ? ms.sad32 : ms.sad64) + at);
        consider(row, col, sad);
      }
    } else {
      // a unit = 4 neighbouring columns of the copy (one word) in one window row: four candidates that share
      // every word they read; candidate e of the unit starts e bytes into the word
      const int qpr = (K + off + 3) >> 2;                       // units per window row
      const int wpr = j.w >> 2;                                 // words per PU row
      const uint32_t *cur32 = reinterpret_cast<const uint32_t *>(sm.cur);
      for (int u = tid; u < K * qpr; u += kPuThreads) {
        const int row = u / qpr, q = u - row * qpr;
        uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        for (int y = 0; y < j.h; y++) {
          const uint32_t *rw = sm.region + (row + y) * pitch_w + q;
          const uint32_t *cw = cur32 + y * wpr;
          uint32_t w0 = rw[0];
          for (int k = 0; k < wpr; k++) {
            const uint32_t w1 = rw[k + 1], c = cw[k];
            a0 = sad4(c, w0, a0);
            a1 = sad4(c, __funnelshift_r(w0, w1, 8), a1);
            a2 = sad4(c, __funnelshift_r(w0, w1, 16), a2);
            a3 = sad4(c, __funnelshift_r(w0, w1, 24), a3);
            w0 = w1;
          }
        }
        const uint32_t acc[4] = { a0, a1, a2, a3 };
#pragma unroll
        for (int e = 0; e < 4; e++) {
          const int col = q * 4 + e - off;
          if (col >= 0 && col < K) consider(row, col, acc[e]);
        }
      }
    }
  }
  best = warp_min(best);
  __syncthreads();
  if (lane == 0) sm.wmin[warp] = best;
  __syncthreads();
  if (tid == 0) {
    for (int q = 1; q < kPuThreads / 32; q++) best = sm.wmin[q] < best ? sm.wmin[q] : best;
    park_winner(best, f, pu, merge, job, out);
  }
}

// a PU and its jobs as kernel arguments: the single-PU call of an encoder host has them in the launch itself
// instead of behind a pointer the CTA would first have to chase (to mapped host memory: a PCIe round trip each)
struct PuArgs {
  kvz_cuda_me_pu pu;
  kvz_cuda_me_job jobs[KVZ_CUDA_MAX_REFS];
};
struct PuArgsN { PuArgs a[kMePuArgsMax]; };     // the PUs of one launch that travel as kernel arguments (CTA i takes a[i])

template <bool ARGS>
__global__ void __launch_bounds__(kPuThreads)
me_pu_kernel(const __grid_constant__ MeFrame f, const __grid_constant__ PuArgsN argsn, const kvz_cuda_me_pu *__restrict__ pus,
             const kvz_cuda_me_job *__restrict__ jobs, kvz_cuda_me_result *results, int n_pus, int box_dx, int box_dy,
             int timing, volatile uint32_t *done_flag, uint32_t done_seq)
{
  unsigned long long t0 = 0, t1 = 0, t2 = 0;
  if (timing && threadIdx.x == 0) t0 = gtime();
  extern __shared__ __align__(16) unsigned char pu_dyn[];
  __shared__ kvz_cuda_merge_cand merge[5];
  __shared__ uint8_t handled[KVZ_CUDA_MAX_REFS];
  __shared__ unsigned long long part[kPuThreads / 32];
  // the PU, its jobs and its results live in shared memory while the CTA works: `pus`, `jobs` and `results` may be
  // mapped HOST memory (kvz_cuda_enc_search_pu), read once and written once
  __shared__ kvz_cuda_me_pu spu;
  __shared__ kvz_cuda_me_job sjobs[KVZ_CUDA_MAX_REFS];
  __shared__ __align__(16) kvz_cuda_me_result sres[KVZ_CUDA_MAX_REFS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (ARGS) {
    const PuArgs &args = argsn.a[blockIdx.x];
    if (threadIdx.x < sizeof(kvz_cuda_me_pu) / 4)
      reinterpret_cast<uint32_t *>(&spu)[threadIdx.x] = reinterpret_cast<const uint32_t *>(&args.pu)[threadIdx.x];
    if (threadIdx.x < sizeof(sjobs) / 4)
      reinterpret_cast<uint32_t *>(sjobs)[threadIdx.x] = reinterpret_cast<const uint32_t *>(args.jobs)[threadIdx.x];
    if (threadIdx.x < KVZ_CUDA_MAX_REFS) handled[threadIdx.x] = 0;
    __syncthreads();
    if (threadIdx.x < 5) merge[threadIdx.x] = spu.merge_cand[threadIdx.x];
    __syncthreads();
  } else {
    const uint32_t *src = reinterpret_cast<const uint32_t *>(pus + blockIdx.x);
    if (threadIdx.x < sizeof(kvz_cuda_me_pu) / 4) reinterpret_cast<uint32_t *>(&spu)[threadIdx.x] = src[threadIdx.x];
    if (threadIdx.x < KVZ_CUDA_MAX_REFS) handled[threadIdx.x] = 0;
    __syncthreads();
    const int nj = min((int)spu.num_jobs, KVZ_CUDA_MAX_REFS);
    const uint32_t *jsrc = reinterpret_cast<const uint32_t *>(jobs + spu.first_job);
    if (threadIdx.x < nj * (int)(sizeof(kvz_cuda_me_job) / 4)) reinterpret_cast<uint32_t *>(sjobs)[threadIdx.x] = jsrc[threadIdx.x];
    if (threadIdx.x < 5) merge[threadIdx.x] = spu.merge_cand[threadIdx.x];
    __syncthreads();
  }
  const kvz_cuda_me_pu &pu = spu;
  const kvz_cuda_me_job *pj = sjobs;
  kvz_cuda_me_result *pr = sres;
  kvz_cuda_me_result *const gres = results + (ARGS ? blockIdx.x * (unsigned)KVZ_CUDA_MAX_REFS : spu.first_job);

  // ---- phase 1: integer search ----
  if (f.ime_search != KVZ_CUDA_IME_FULL) {
    for (int q = warp; q < pu.num_jobs; q += kPuThreads / 32) heuristic_job(lane, f, pu, merge, pj[q], pr + q);
  } else {
    // the warps of the CTA are dealt out to the jobs: 4 warps per job up to 4 jobs, 2 up to 8, else 1
    const int nsub = pu.num_jobs <= 4 ? 4 : pu.num_jobs <= 8 ? 2 : 1;
    const int q = warp / nsub, sub = warp - q * nsub;
    if (box_dx > 0 && q < pu.num_jobs) {
      unsigned char *base = pu_dyn + warp * box_scan_warp_bytes(box_dx, box_dy);
      uint4 *ring = reinterpret_cast<uint4 *>(base);
      uint4 *gxv = ring + kScanRing * 32;
      const int nv = box_dx / 8;
      uint2 *gy = reinterpret_cast<uint2 *>(gxv + 4 * nv);
      uint32_t *rowinfo = reinterpret_cast<uint32_t *>(gy + box_dy);
      int *hits = reinterpret_cast<int *>(reinterpret_cast<unsigned char *>(rowinfo) + (((size_t)box_dy * 4 + 15) & ~(size_t)15));
      ScanSmemT<true> sm = { ring, gxv, nv, gy, rowinfo, hits, hits + 5, hits[10] };
      const bool ok = scan_job<true>(lane, sm, f, pu, merge, pj[q], pr + q, false, sub, nsub, &part[warp]);
      if (lane == 0 && sub == 0 && ok) handled[q] = 1;
    }
    __syncthreads();
    if (box_dx > 0 && q < pu.num_jobs && sub == 0 && lane == 0 && handled[q]) {
      unsigned long long best = part[warp];
      for (int k = 1; k < nsub; k++) best = part[warp + k] < best ? part[warp + k] : best;
      park_winner(best, f, pu, merge, pj[q], pr + q);
    }
    // ---- phase 2: the jobs with a window outside every box ----
    __syncthreads();
    if (timing && threadIdx.x == 0) t1 = gtime();
    CoopSmem &cs = *reinterpret_cast<CoopSmem *>(pu_dyn);
    for (int q = 0; q < pu.num_jobs; q++)
      if (!handled[q]) coop_scan_job(cs, f, pu, merge, pj[q], pr + q);
  }
  __syncthreads();                         // the parked winners (shared memory) are visible to the refinement
  if (timing && threadIdx.x == 0) { t2 = gtime(); if (!t1) t1 = t2; }

  // ---- phase 3: refinement.  The references of the PU are refined SIDE BY SIDE, each by its own group of threads,
  // with both outcomes of the `best_cost < *inter_cost` gate computed (refine_job<SPEC>); thread 0 then walks the
  // references in order and lets the carry pick (refine_finish).  A 64-wide PU needs the whole CTA per reference. ----
  __shared__ RefineBoth both[KVZ_CUDA_MAX_REFS];
  const int msz = max((int)pu.w, (int)pu.h);
  if ((pu.w | pu.h) & 7) {                   // sides 4 / 12 (SMP / AMP): the 4-sample SATD strips are not implemented
    if (threadIdx.x < pu.num_jobs) {
      kvz_cuda_me_result r = {};
      r.cost = 0xFFFFFFFFu; r.took = 0xFE;
      gres[threadIdx.x] = r;
    }
    return;
  }
  auto stage_cur = [&](uint8_t *dst, int pitch, int tid, int nthreads) {
    for (int i = tid; i < pu.h * (pu.w >> 2); i += nthreads) {
      const int r = i / (pu.w >> 2), c = i % (pu.w >> 2);
      const uint8_t *src = f.cur + (size_t)(f.tile_y + pu.y + r) * f.cur_pitch + f.tile_x + pu.x + c * 4;
      *reinterpret_cast<uint32_t *>(&dst[r * pitch + c * 4]) = *reinterpret_cast<const uint32_t *>(src);
    }
  };
  uint32_t dummy = 0;
  if (msz <= 16) {
    using SM = RefineSmem<32, 16>;
    constexpr int kGroups = 8;
    if (warp < kGroups && warp < pu.num_jobs) {
      SM &sm = reinterpret_cast<SM *>(pu_dyn)[warp];
      Group<32> g;
      g.tid = lane; g.id = warp;
      stage_cur(sm.cur, SM::kCurPitch, lane, 32);
      if (lane < 5) sm.merge[lane] = pu.merge_cand[lane];
      g.sync();
      for (int q = warp; q < pu.num_jobs; q += kGroups) refine_job<32, 16, true>(g, sm, f, pu, pj[q], dummy, pr + q, &both[q]);
    }
  } else if (msz <= 32) {
    using SM = RefineSmem<128, 32>;
    constexpr int kGroups = kPuThreads / 128;
    const int gi = threadIdx.x >> 7;
    if (gi < pu.num_jobs) {
      SM &sm = reinterpret_cast<SM *>(pu_dyn)[gi];
      Group<128> g;
      g.tid = threadIdx.x & 127; g.id = gi; g.bar = 1 + gi;
      stage_cur(sm.cur, SM::kCurPitch, g.tid, 128);
      if (g.tid < 5) sm.merge[g.tid] = pu.merge_cand[g.tid];
      g.sync();
      for (int q = gi; q < pu.num_jobs; q += kGroups) refine_job<128, 32, true>(g, sm, f, pu, pj[q], dummy, pr + q, &both[q]);
    }
  } else {
    using SM = RefineSmem<kPuThreads, 64>;
    SM &sm = *reinterpret_cast<SM *>(pu_dyn);
    Group<kPuThreads> g;
    g.tid = threadIdx.x; g.id = 0;
    stage_cur(sm.cur, SM::kCurPitch, g.tid, kPuThreads);
    if (g.tid < 5) sm.merge[g.tid] = pu.merge_cand[g.tid];
    g.sync();
    for (int q = 0; q < pu.num_jobs; q++) refine_job<kPuThreads, 64, true>(g, sm, f, pu, pj[q], dummy, pr + q, &both[q]);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t inter_cost = 0x7FFFFFFFu;         // *inter_cost = MAX_INT (search_inter.c:1460)
    for (int q = 0; q < pu.num_jobs; q++) refine_finish(f, pu, merge, pj[q], both[q], inter_cost, pr + q);
  }
  __syncthreads();
  static_assert(sizeof(kvz_cuda_me_result) == 16, "one 16-byte store per result");
  if (threadIdx.x < pu.num_jobs && threadIdx.x < 32)
    reinterpret_cast<uint4 *>(gres)[threadIdx.x] = reinterpret_cast<const uint4 *>(sres)[threadIdx.x];
  // A host that polls instead of synchronising the stream (kvz_cuda_enc_search_pu) names a word of mapped memory:
  // it receives this call's sequence number after the results, behind a system-scope fence.  Without one the host
  // reads the results after the kernel has completed, which orders them.
  if (done_flag != nullptr && threadIdx.x < 32) {
    __syncwarp();
    if (threadIdx.x == 0) { __threadfence_system(); done_flag[ARGS ? blockIdx.x : 0] = done_seq; }
  }
  if (timing && threadIdx.x == 0) {
    const unsigned long long t3 = gtime();
    const int m = max((int)pu.w, (int)pu.h), sz = m > 32 ? 3 : m > 16 ? 2 : m > 8 ? 1 : 0;
    atomicAdd(&g_pu_phase_ns[sz][0], t1 - t0);
    atomicAdd(&g_pu_phase_ns[sz][1], t2 - t1);
    atomicAdd(&g_pu_phase_ns[sz][2], t3 - t2);
    atomicAdd(&g_pu_phase_ns[sz][3], 1ull);
  }
}

}  // namespace

int launch_me_search(const MeFrame &f, const kvz_cuda_me_pu *pus, const kvz_cuda_me_job *jobs,
                     kvz_cuda_me_result *results, int n_pus, cudaStream_t stream)
{
  if (n_pus <= 0) return 1;
  int box_dx = 0, box_dy = 0;                   // the largest box sides among the map sets
  for (int r = 0; r < f.n_refs; r++)
    for (int m = 0; m < f.refs[r].n_sets; m++) {
      if (f.refs[r].sets[m].n_dx > box_dx) box_dx = f.refs[r].sets[m].n_dx;
      if (f.refs[r].sets[m].n_dy > box_dy) box_dy = f.refs[r].sets[m].n_dy;
    }
  if (f.ime_search != KVZ_CUDA_IME_FULL) box_dx = box_dy = 0;      // heuristic searches evaluate their probes on demand
  if (box_dx > 0) {
    const size_t smem = kScanWarps * box_scan_warp_bytes(box_dx, box_dy);
    auto kern = me_scan_kernel<true>;
    KVZ_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<n_pus, kScanWarps * 32, smem, stream>>>(f, pus, jobs, results, n_pus, box_dx, box_dy);
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
  }
  me_scan_kernel<false><<<n_pus, kScanWarps * 32, 0, stream>>>(f, pus, jobs, results, n_pus, box_dx, 0);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  {
    using SM = RefineSmem<32, 16>;
    const size_t smem = sizeof(SM) * 4;
    auto kern = me_refine_kernel<32, 16, kCtaThreads>;
    KVZ_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(n_pus + 3) / 4, kCtaThreads, smem, stream>>>(f, pus, jobs, results, n_pus);
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
  }
  {
    using SM = RefineSmem<kBigGroup, 64>;
    const size_t smem = sizeof(SM);
    auto kern = me_refine_kernel<kBigGroup, 64, kBigGroup>;
    KVZ_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<n_pus, kBigGroup, smem, stream>>>(f, pus, jobs, results, n_pus);
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
  }
  return 1;
}

// device time of me_pu_kernel's phases (KVZ_CUDA_STATS): out[size class][box scan ns, CTA scan ns, refinement ns, calls]
int me_pu_phase_times(unsigned long long out[4][4])
{
  KVZ_CUDA_TRY(cudaMemcpyFromSymbol(out, g_pu_phase_ns, sizeof(g_pu_phase_ns)));
  unsigned long long bs[4];
  KVZ_CUDA_TRY(cudaMemcpyFromSymbol(bs, g_box_scan_ns, sizeof(bs)));
  if (bs[3]) fprintf(stderr, "kvz_cuda_enc: box scan per job: tables %.1f us, seed %.1f us, stream %.1f us (%llu jobs)\n",
                     bs[0] / 1000.0 / bs[3], bs[1] / 1000.0 / bs[3], bs[2] / 1000.0 / bs[3], bs[3]);
  return 1;
}

// every PU by one CTA in ONE launch (me_pu_kernel).  host_pu / host_jobs non-null (n_pus <= kMePuArgsMax): the PUs and
// their jobs (KVZ_CUDA_MAX_REFS job records per PU) travel as kernel arguments and `pus` / `jobs` are not read; the
// results of PU i go to results[i * KVZ_CUDA_MAX_REFS ..], done_flag[i] receives done_seq after them.
int launch_me_search_pu(const MeFrame &f, const kvz_cuda_me_pu *pus, const kvz_cuda_me_job *jobs,
                        kvz_cuda_me_result *results, int n_pus, cudaStream_t stream, const kvz_cuda_me_pu *host_pu,
                        const kvz_cuda_me_job *host_jobs, uint32_t *done_flag, uint32_t done_seq)
{
  if (n_pus <= 0) return 1;
  int box_dx = 0, box_dy = 0;
  for (int r = 0; r < f.n_refs; r++)
    for (int m = 0; m < f.refs[r].n_sets; m++) {
      if (f.refs[r].sets[m].n_dx > box_dx) box_dx = f.refs[r].sets[m].n_dx;
      if (f.refs[r].sets[m].n_dy > box_dy) box_dy = f.refs[r].sets[m].n_dy;
    }
  if (f.ime_search != KVZ_CUDA_IME_FULL) box_dx = box_dy = 0;
  size_t smem = sizeof(RefineSmem<kPuThreads, 64>);
  if (8 * sizeof(RefineSmem<32, 16>) > smem) smem = 8 * sizeof(RefineSmem<32, 16>);
  if ((kPuThreads / 128) * sizeof(RefineSmem<128, 32>) > smem) smem = (kPuThreads / 128) * sizeof(RefineSmem<128, 32>);
  if (coop_smem_bytes(f.range) > smem) smem = coop_smem_bytes(f.range);
  if (box_dx > 0 && (kPuThreads / 32) * box_scan_warp_bytes(box_dx, box_dy) > smem) smem = (kPuThreads / 32) * box_scan_warp_bytes(box_dx, box_dy);
  static size_t opted_args = 0, opted_batch = 0;  // opt in once per size (one device per process; any host thread)
  static std::mutex opt_lock;
  KVZ_CUDA_TRY(opt_in_dynamic_smem(me_pu_kernel<true>, smem, opted_args, opt_lock));
  KVZ_CUDA_TRY(opt_in_dynamic_smem(me_pu_kernel<false>, smem, opted_batch, opt_lock));
  static const int timing = getenv("KVZ_CUDA_STATS") != nullptr;
  static bool timing_set = false;
  if (timing && !timing_set) { KVZ_CUDA_TRY(cudaMemcpyToSymbol(g_box_scan_timing, &timing, sizeof(int))); timing_set = true; }
  static bool seed_set = false;
  if (!seed_set) {
    const int use = getenv("KVZ_CUDA_PU_SEED") ? atoi(getenv("KVZ_CUDA_PU_SEED")) : 1;
    if (!use) KVZ_CUDA_TRY(cudaMemcpyToSymbol(g_pu_use_seed, &use, sizeof(int)));
    seed_set = true;
  }
  if (host_pu != nullptr && host_jobs != nullptr) {
    if (n_pus > kMePuArgsMax) return fail_msg("launch_me_search_pu: too many PUs for the argument form");
    PuArgsN a;
    for (int i = 0; i < n_pus; i++) {
      a.a[i].pu = host_pu[i];
      a.a[i].pu.first_job = 0;
      const int nj = host_pu[i].num_jobs < KVZ_CUDA_MAX_REFS ? host_pu[i].num_jobs : KVZ_CUDA_MAX_REFS;
      memcpy(a.a[i].jobs, host_jobs + (size_t)i * KVZ_CUDA_MAX_REFS, (size_t)nj * sizeof(kvz_cuda_me_job));
      if (nj < KVZ_CUDA_MAX_REFS) memset(a.a[i].jobs + nj, 0, (size_t)(KVZ_CUDA_MAX_REFS - nj) * sizeof(kvz_cuda_me_job));
    }
    me_pu_kernel<true><<<n_pus, kPuThreads, smem, stream>>>(f, a, nullptr, nullptr, results, n_pus, box_dx, box_dy, timing, done_flag, done_seq);
  } else {
    static const PuArgsN none = {};
    me_pu_kernel<false><<<n_pus, kPuThreads, smem, stream>>>(f, none, pus, jobs, results, n_pus, box_dx, box_dy, timing, nullptr, 0u);
  }
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
