// kvazaar_b200/csrc/me_search.cu
//
// Per-PU motion decision on device: for every PU of a batch and every reference
// picture it is searched against, reproduce search_pu_inter_ref()
// (search_inter.c:1134-1297) minus the candidate derivation (the AMVP / merge
// candidates are inputs):
//
//   1. integer search = search_mv_full (search_inter.c:886-956): the zero
//      window is read from the dense SAD maps of ime_sad.cu; the windows round
//      the previous-frame vector and the merge candidates are evaluated on
//      demand from the border-replicated reference plane.  Every candidate costs
//      SAD + calc_mvd_cost (search_inter.c:373-412) and the winner is the FIRST
//      candidate in the reference's visiting order with the smallest cost
//      (check_mv_cost accepts on strict '<', search_inter.c:195-232), which is a
//      lexicographic arg-min over (cost, window, row, column);
//   2. fractional refinement = search_frac (search_inter.c:965-1128): the 8-tap
//      filters of kvz_filter_{hpel,qpel}_blocks_*_luma (ipol-generic.c:192-658,
//      int16 intermediates included) and kvz_satd_any_size_quad, four candidates
//      per step, gated by `best_cost < *inter_cost` carried across the reference
//      loop of search_pu_inter (search_inter.c:1505-1511); otherwise SATD at the
//      integer vector (search_inter.c:1244-1256);
//   3. merge-index match and select_mv_cand (search_inter.c:1258-1297).
//
// One thread GROUP works on one PU: a warp for PUs up to 16x16 (four PUs per
// CTA), the whole 128-thread CTA for larger ones.
#include "common.cuh"
#include "me_search.cuh"
#include "satd.cuh"

namespace kvz {
namespace {

constexpr int kCtaThreads = 128;

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t d;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

__device__ __constant__ int8_t c_luma_fir[4][8] = {       // HEVC luma DCT-IF (filter.c:54-60)
  { 0, 0, 0, 64, 0, 0, 0, 0 }, { -1, 4, -10, 58, 17, -5, 1, 0 },
  { -1, 4, -11, 40, 40, -11, 4, -1 }, { 0, 1, -5, 17, 58, -10, 4, -1 } };

// 2*floor(log2(v+2)): get_ep_ex_golomb_bitcost (search_inter.c:235-253)
__device__ __forceinline__ uint32_t golomb_bits(uint32_t v) { return 2u * (31u - __clz(v + 2u)); }

__device__ __forceinline__ uint32_t mvd_bits(int dx, int dy) { return golomb_bits(abs(dx)) + golomb_bits(abs(dy)); }

// ---- thread group: a warp or the whole CTA ----
template <int G>
struct Group {
  int tid;        // 0..G-1 inside the group
  int id;         // group index inside the CTA
  __device__ __forceinline__ void sync() const
  {
    if (G == 32) __syncwarp();
    else __syncthreads();
  }
};

struct JobCtx {
  // PU
  int x, y, w, h;            // tile-relative origin, size
  int ref_idx;
  int16_t mv_cand[2][2];
  int n_merge;
  const kvz_cuda_merge_cand *merge;
  int lambda;
};

// fracmv_within_tile (search_inter.c:87-172); x,y quarter-pel
__device__ bool within_tile(const MeFrame &f, const JobCtx &j, int x, int y)
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

// calc_mvd_cost (search_inter.c:373-412): bits for vector (x,y) << mv_shift
__device__ uint32_t mvd_cost_bits(const MeFrame &f, const JobCtx &j, int x, int y, int mv_shift)
{
  x *= 1 << mv_shift;
  y *= 1 << mv_shift;
  for (int i = 0; i < j.n_merge; i++) {
    const kvz_cuda_merge_cand &m = j.merge[i];
    if (m.dir == 3) continue;
    const int l = m.dir - 1;
    if (m.mv[l][0] == x && m.mv[l][1] == y && f.ref_LX[l][m.ref[l]] == j.ref_idx) return (uint32_t)i;
  }
  const uint32_t c0 = mvd_bits(x - j.mv_cand[0][0], y - j.mv_cand[0][1]);
  const uint32_t c1 = mvd_bits(x - j.mv_cand[1][0], y - j.mv_cand[1][1]);
  return min(c0, c1);
}

// SAD of the PU (staged in smem, pitch 64) against the reference block whose top-left
// sample is (rx, ry) in frame coordinates: kvz_image_calc_sad (image.c:455-479).  The
// block origin is clamped into the replicated border, which reads the same samples as
// the per-pixel coordinate clamp as long as pad >= PU size.
__device__ uint32_t sad_on_demand(const MeFrame &f, const MeRef &r, const uint8_t *cur, int w, int h, int rx, int ry)
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
    const uint32_t *cw = reinterpret_cast<const uint32_t *>(cur + yy * 64);
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
      acc = sad4(cw[c], v, acc);
    }
    rw += wpitch;
  }
  return acc;
}

// sample of the replicated plane with the coordinate clamped into it (== the per-pixel
// clamp of kvz_get_extended_block, ipol-generic.c:761-781)
__device__ __forceinline__ int ref_px(const MeFrame &f, const MeRef &r, int x, int y)
{
  x = clampi(x, -r.pad, f.width + r.pad - 1);
  y = clampi(y, -r.pad, f.height + r.pad - 1);
  return r.plane[(size_t)(y + r.pad) * r.pitch + (x + r.pad)];
}

template <int G>
__device__ __forceinline__ unsigned long long group_min(const Group<G> &g, unsigned long long v, unsigned long long *scratch)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = __shfl_xor_sync(0xffffffffu, v, o);
    v = t < v ? t : v;
  }
  if (G == 32) return v;
  const int warp = g.tid >> 5;
  g.sync();
  if ((g.tid & 31) == 0) scratch[warp] = v;
  g.sync();
  unsigned long long m = scratch[0];
#pragma unroll
  for (int i = 1; i < G / 32; i++) m = scratch[i] < m ? scratch[i] : m;
  return m;
}

// per-window separable tables of the AMVP bit cost: bits = min_c(gx[c][col] + gy[c][row])
struct BitTables {
  alignas(16) uint2 gx[132];     // lambda * bits of column i for AMVP candidate 0 (.x) and 1 (.y)
  uint2 gy[132];                 // same for row i
};

template <int G, int S>
struct GroupSmem {
  static constexpr int kWinPitch = S + 16;
  alignas(16) uint8_t cur[S * 64 > 0 ? S * 64 : 1];            // PU rows, pitch 64
  alignas(16) uint8_t win[(S + 9) * kWinPitch];                // FME source window
  alignas(16) int16_t hor[4][(S + 7) * S];                     // horizontally filtered candidates
  BitTables bt;
  unsigned long long red[4];
  uint32_t cost4[4];
  int hit_pos[5];
  int hit_bits[5];
  int n_hit;
};

template <int G, int S>
__device__ void search_one_ref(const Group<G> &g, GroupSmem<G, S> &sm, const MeFrame &f, const kvz_cuda_me_pu &pu,
                               const kvz_cuda_me_job &job, uint32_t &inter_cost, kvz_cuda_me_result *out)
{
  JobCtx j;
  j.x = pu.x; j.y = pu.y; j.w = pu.w; j.h = pu.h;
  j.ref_idx = job.ref_idx;
  j.mv_cand[0][0] = job.mv_cand[0][0]; j.mv_cand[0][1] = job.mv_cand[0][1];
  j.mv_cand[1][0] = job.mv_cand[1][0]; j.mv_cand[1][1] = job.mv_cand[1][1];
  j.n_merge = pu.num_merge_cand;
  j.merge = pu.merge_cand;
  j.lambda = f.lambda_sqrt_i;
  const MeRef &ref = f.refs[job.ref_idx];
  const int R = f.range, K = 2 * R + 1;
  const bool free_mv = !(f.owf && f.wpp) && f.mv_constraint == 0;
  const int fx0 = f.tile_x + j.x, fy0 = f.tile_y + j.y;       // frame-absolute PU origin

  // ---- window list of search_mv_full ----
  int wcx[7], wcy[7], nwin = 1;
  wcx[0] = 0; wcy[0] = 0;
  {
    const int ex = job.extra_mv[0] >> 2, ey = job.extra_mv[1] >> 2;
    bool in_merge = false;                       // mv_in_merge (search_inter.c:260-273)
    for (int i = 0; i < j.n_merge; i++) {
      const kvz_cuda_merge_cand &m = j.merge[i];
      if (m.dir == 3) continue;
      if (((m.mv[m.dir - 1][0] + 2) >> 2) == ex && ((m.mv[m.dir - 1][1] + 2) >> 2) == ey) in_merge = true;
    }
    if (!in_merge) { wcx[nwin] = ex; wcy[nwin] = ey; nwin++; }
    for (int i = 0; i < j.n_merge; i++) {
      const kvz_cuda_merge_cand &m = j.merge[i];
      if (m.dir == 3) continue;
      const int mx = m.mv[m.dir - 1][0] >> 2, my = m.mv[m.dir - 1][1] >> 2;
      if (mx == 0 && my == 0) continue;
      wcx[nwin] = mx; wcy[nwin] = my; nwin++;
    }
  }

  // level of the dense maps that holds this PU, if any
  int level = -1;
  if (j.w == j.h && (j.w == 8 || j.w == 16 || j.w == 32 || j.w == 64) && (j.x % j.w) == 0 && (j.y % j.h) == 0)
    level = j.w == 8 ? 0 : j.w == 16 ? 1 : j.w == 32 ? 2 : 3;

  unsigned long long best = ~0ull;

  for (int wi = 0; wi < nwin; wi++) {
    const int cx = wcx[wi], cy = wcy[wi];
    // separable AMVP bit tables and the merge overrides of this window
    g.sync();
    for (int i = g.tid; i < K; i += G) {
      const int qx = (cx - R + i) * 4, qy = (cy - R + i) * 4;
      const uint32_t lam = (uint32_t)j.lambda;
      sm.bt.gx[i] = make_uint2(lam * golomb_bits(abs(qx - j.mv_cand[0][0])), lam * golomb_bits(abs(qx - j.mv_cand[1][0])));
      sm.bt.gy[i] = make_uint2(lam * golomb_bits(abs(qy - j.mv_cand[0][1])), lam * golomb_bits(abs(qy - j.mv_cand[1][1])));
    }
    if (g.tid == 0) {
      int n = 0;
      for (int i = 0; i < j.n_merge; i++) {
        const kvz_cuda_merge_cand &m = j.merge[i];
        if (m.dir == 3) continue;
        const int l = m.dir - 1;
        if ((m.mv[l][0] & 3) || (m.mv[l][1] & 3) || f.ref_LX[l][m.ref[l]] != j.ref_idx) continue;
        const int col = (m.mv[l][0] >> 2) - (cx - R), row = (m.mv[l][1] >> 2) - (cy - R);
        if (col < 0 || col >= K || row < 0 || row >= K) continue;
        const int pos = row * K + col;
        bool dup = false;
        for (int q = 0; q < n; q++) dup |= sm.hit_pos[q] == pos;     // first matching candidate wins
        if (!dup) { sm.hit_pos[n] = pos; sm.hit_bits[n] = i; n++; }
      }
      sm.n_hit = n;
    }
    g.sync();
    const int n_hit = sm.n_hit;

    // exact per-candidate path: tile restriction and merge override included
    auto consider = [&](int row, int col, uint32_t sad) {
      const int mx = cx - R + col, my = cy - R + row;
      if (!free_mv && !within_tile(f, j, mx * 4, my * 4)) return;
      const uint2 bx = sm.bt.gx[col], by = sm.bt.gy[row];
      uint32_t mvd = min(bx.x + by.x, bx.y + by.y);
      const int pos = row * K + col;
      for (int q = 0; q < n_hit; q++)
        if (sm.hit_pos[q] == pos) mvd = (uint32_t)sm.hit_bits[q] * (uint32_t)j.lambda;
      const uint32_t cost = sad + mvd;
      const unsigned long long key = ((unsigned long long)cost << 32) | (unsigned)((wi << 16) | pos);
      best = key < best ? key : best;
    };

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
      // Window straight from the dense maps.  K = 8m + 1: m full vectors of 8 entries per row,
      // then the last column.  A thread meets its candidates in visiting order, so within the
      // window a strict '<' keeps the first of equal costs.
      const int n = 8 >> level;
      const int ctu = (j.y >> 6) * f.ctus_x + (j.x >> 6);
      const size_t prow = (size_t)ctu * n * n + ((j.y & 63) / j.h) * n + (j.x & 63) / j.w;
      const int vpr = K >> 3;
      uint32_t wbest = 0xFFFFFFFFu;
      int wpos = 0;
      // vectors holding a merge-override position take the exact path
      int hv[5];
#pragma unroll
      for (int q = 0; q < 5; q++)
        hv[q] = (q < n_hit && (sm.hit_pos[q] % K) < K - 1) ? (sm.hit_pos[q] / K) * vpr + ((sm.hit_pos[q] % K) >> 3) : -1;
      const bool fast = free_mv;
      const uint16_t *map16 = (level == 0 ? ms.sad8 : ms.sad16) + prow * (K + 1) * f.map_pitch;
      const uint32_t *map32 = (level == 2 ? ms.sad32 : ms.sad64) + prow * (K + 1) * f.map_pitch;
      for (int i = g.tid; i < K * vpr; i += G) {
        const int row = i / vpr, c0 = (i - row * vpr) * 8;
        uint32_t sad[8];
        if (level < 2) {
          const uint4 v = __ldg(reinterpret_cast<const uint4 *>(map16 + (size_t)row * f.map_pitch + c0));
          sad[0] = v.x & 0xFFFFu; sad[1] = v.x >> 16; sad[2] = v.y & 0xFFFFu; sad[3] = v.y >> 16;
          sad[4] = v.z & 0xFFFFu; sad[5] = v.z >> 16; sad[6] = v.w & 0xFFFFu; sad[7] = v.w >> 16;
        } else {
          const uint4 v0 = __ldg(reinterpret_cast<const uint4 *>(map32 + (size_t)row * f.map_pitch + c0));
          const uint4 v1 = __ldg(reinterpret_cast<const uint4 *>(map32 + (size_t)row * f.map_pitch + c0 + 4));
          sad[0] = v0.x; sad[1] = v0.y; sad[2] = v0.z; sad[3] = v0.w;
          sad[4] = v1.x; sad[5] = v1.y; sad[6] = v1.z; sad[7] = v1.w;
        }
        const bool special = !fast || i == hv[0] || i == hv[1] || i == hv[2] || i == hv[3] || i == hv[4];
        if (special) {
#pragma unroll
          for (int e = 0; e < 8; e++) consider(row, c0 + e, sad[e]);
        } else {
          const uint2 by = sm.bt.gy[row];
          const uint4 *bxp = reinterpret_cast<const uint4 *>(&sm.bt.gx[c0]);
#pragma unroll
          for (int e2 = 0; e2 < 4; e2++) {
            const uint4 bx = bxp[e2];                 // columns c0+2*e2 and c0+2*e2+1
            const uint32_t ca = sad[2 * e2] + min(bx.x + by.x, bx.y + by.y);
            const uint32_t cb = sad[2 * e2 + 1] + min(bx.z + by.x, bx.w + by.y);
            if (ca < wbest) { wbest = ca; wpos = row * K + c0 + 2 * e2; }
            if (cb < wbest) { wbest = cb; wpos = row * K + c0 + 2 * e2 + 1; }
          }
        }
      }
      // last column
      for (int row = g.tid; row < K; row += G) {
        const uint32_t sad = level < 2 ? (uint32_t)__ldg(map16 + (size_t)K * f.map_pitch + row)   // row K = column dx = +R
                                       : __ldg(map32 + (size_t)K * f.map_pitch + row);
        consider(row, K - 1, sad);
      }
      if (wbest != 0xFFFFFFFFu) {
        const unsigned long long key = ((unsigned long long)wbest << 32) | (unsigned)((wi << 16) | wpos);
        best = key < best ? key : best;
      }
    } else {
      // on demand, skipping what an earlier window of this list already covered
      for (int p = g.tid; p < K * K; p += G) {
        const int row = p / K, col = p % K;
        const int mx = cx - R + col, my = cy - R + row;
        bool covered = false;
        for (int q = 0; q < wi; q++)
          covered |= (mx >= wcx[q] - R && mx <= wcx[q] + R && my >= wcy[q] - R && my <= wcy[q] + R);
        if (covered) continue;
        if (!free_mv && !within_tile(f, j, mx * 4, my * 4)) continue;
        consider(row, col, sad_on_demand(f, ref, sm.cur, j.w, j.h, fx0 + mx, fy0 + my));
      }
    }
  }
  best = group_min<G>(g, best, sm.red);

  // ---- winner of the integer search ----
  int best_mvx = 0, best_mvy = 0;            // quarter-pel (info->best_mv starts at 0,0)
  uint32_t best_cost = 0xFFFFFFFFu, best_bits = 0;
  if (best != ~0ull) {
    const int wi = (int)((best >> 16) & 0xFFFF), pos = (int)(best & 0xFFFF);
    best_mvx = (wcx[wi] - R + pos % K) * 4;
    best_mvy = (wcy[wi] - R + pos / K) * 4;
    best_cost = (uint32_t)(best >> 32);
    best_bits = mvd_cost_bits(f, j, best_mvx >> 2, best_mvy >> 2, 2);
  }

  const bool do_fme = f.fme_level > 0 && best_cost < inter_cost;
  const bool do_satd = !do_fme && best_cost < 0xFFFFFFFFu;
  if (do_fme || do_satd) {
    // ---- source window round the integer vector: columns -4..w+3, rows -4..h+4 ----
    const int imx = best_mvx >> 2, imy = best_mvy >> 2;
    const int bx = fx0 + imx, by = fy0 + imy;
    constexpr int WP = GroupSmem<G, S>::kWinPitch;
    g.sync();
    for (int i = g.tid; i < (j.h + 9) * (j.w + 8); i += G) {
      const int r = i / (j.w + 8), c = i % (j.w + 8);
      sm.win[r * WP + c] = (uint8_t)ref_px(f, ref, bx - 4 + c, by - 4 + r);
    }
    g.sync();

    const int tiles_x = j.w >> 3, tiles = tiles_x * (j.h >> 3);
    const int lane8 = g.tid & 7;

    // SATD at the integer position: kvz_satd_any_size (strategies-picture.h:62-100), 8x8 tiles
    if (g.tid < 4) sm.cost4[g.tid] = 0;
    g.sync();
    for (int u0 = 0; u0 < tiles * 8; u0 += G) {
      const int u = u0 + g.tid;
      const bool on = u < tiles * 8;
      const int t = on ? (u >> 3) : 0;
      const int ty = t / tiles_x, tx = t % tiles_x;
      int d[8];
      const uint8_t *c = sm.cur + (ty * 8 + lane8) * 64 + tx * 8;
      const uint8_t *p = sm.win + (ty * 8 + lane8 + 4) * WP + tx * 8 + 4;
#pragma unroll
      for (int x = 0; x < 8; x++) d[x] = (int)c[x] - (int)p[x];
      const uint32_t s = satd8_rows(d, lane8);
      if (on && lane8 == 0) atomicAdd(&sm.cost4[0], s);
    }
    g.sync();
    const uint32_t satd_int = sm.cost4[0];

    if (do_satd) {
      // search_inter.c:1244-1256
      best_cost = satd_int + best_bits * (uint32_t)j.lambda;
    } else {
      // ---- search_frac ----
      const int sqx[9] = { 0, -1, 1, 0, 0, -1, 1, -1, 1 };
      const int sqy[9] = { 0, 0, 0, -1, 1, -1, -1, 1, 1 };
      int mvx = imx, mvy = imy;
      uint32_t bits0 = mvd_cost_bits(f, j, mvx, mvy, 2);
      uint32_t fbest_cost = satd_int + bits0 * (uint32_t)j.lambda;
      uint32_t fbest_bits = bits0;
      int best_index = 0, hx = 0, hy = 0, i = 1;
      mvx *= 2; mvy *= 2;
      for (int step = 0; step < f.fme_level; step++) {
        const int mv_shift = step < 2 ? 1 : 0;
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
        // horizontal pass: hor[k][r][x], r = 0..h+6 (ipol-generic.c:233-262 et al.)
        const int hw = j.w >> 2;                              // 4 outputs per item
        const int items = 4 * (j.h + 7) * hw;
        for (int it = g.tid; it < items; it += G) {
          const int k = it / ((j.h + 7) * hw);
          const int rem = it % ((j.h + 7) * hw);
          const int r = rem / hw, x = (rem % hw) * 4;
          const int sx = qx[k] >> 2, fx = qx[k] & 3, sy = qy[k] >> 2;
          const uint8_t *src = sm.win + (r + sy + 1) * WP + (x + sx + 1);
          int px[11];
#pragma unroll
          for (int q = 0; q < 11; q++) px[q] = src[q];
#pragma unroll
          for (int o = 0; o < 4; o++) {
            int s = 0;
#pragma unroll
            for (int t = 0; t < 8; t++) s += c_luma_fir[fx][t] * px[o + t];
            sm.hor[k][r * j.w + x + o] = (int16_t)s;          // >> (BD-8) == >> 0
          }
        }
        g.sync();
        // vertical pass + SATD: one 8x8 tile row per lane, 8 lanes per tile
        const int units = 4 * tiles * 8;
        for (int u0 = 0; u0 < units; u0 += G) {
          const int u = u0 + g.tid;
          const bool on = u < units;
          const int k = on ? u / (tiles * 8) : 0;
          const int t = on ? (u % (tiles * 8)) >> 3 : 0;
          const int ty = t / tiles_x, tx = t % tiles_x;
          const int fy = qy[k] & 3;
          int acc[8];
#pragma unroll
          for (int x = 0; x < 8; x++) acc[x] = 0;
          const int16_t *hp = sm.hor[k] + (ty * 8 + lane8) * j.w + tx * 8;
#pragma unroll
          for (int t8 = 0; t8 < 8; t8++) {
            const int cf = c_luma_fir[fy][t8];
            const uint4 v = *reinterpret_cast<const uint4 *>(hp + t8 * j.w);
            const uint32_t wv[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
            for (int x = 0; x < 8; x++) acc[x] += cf * (int)(int16_t)((wv[x >> 1] >> ((x & 1) * 16)) & 0xFFFF);
          }
          int d[8];
          const uint8_t *c = sm.cur + (ty * 8 + lane8) * 64 + tx * 8;
#pragma unroll
          for (int x = 0; x < 8; x++) {
            const int v16 = (int)(int16_t)(acc[x] >> 6);       // int16 store of ipol-generic.c:285 et al.
            const int pix = clampi((v16 + 32) >> 6, 0, 255);
            d[x] = (int)c[x] - pix;
          }
          const uint32_t s = satd8_rows(d, lane8);
          if (on && lane8 == 0) atomicAdd(&sm.cost4[k], s);
        }
        g.sync();
        uint32_t costs[4], bits[4];
        bool ok[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int cxk = mvx + sqx[i + k], cyk = mvy + sqy[i + k];
          ok[k] = free_mv || within_tile(f, j, cxk * (1 << mv_shift), cyk * (1 << mv_shift));
          costs[k] = sm.cost4[k];
          bits[k] = 0;
          if (ok[k]) {
            bits[k] = mvd_cost_bits(f, j, cxk, cyk, mv_shift);
            costs[k] += bits[k] * (uint32_t)j.lambda;
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

  // ---- merge match, AMVP candidate choice, carry of *inter_cost (search_inter.c:1258-1297) ----
  int merged = 0, merge_idx = 0;
  for (merge_idx = 0; merge_idx < j.n_merge; merge_idx++) {
    const kvz_cuda_merge_cand &m = j.merge[merge_idx];
    if (m.dir != 3 && m.mv[m.dir - 1][0] == best_mvx && m.mv[m.dir - 1][1] == best_mvy &&
        f.ref_LX[m.dir - 1][m.ref[m.dir - 1]] == j.ref_idx) { merged = 1; break; }
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
  if (g.tid == 0) {
    kvz_cuda_me_result r;
    r.mv[0] = (int16_t)best_mvx; r.mv[1] = (int16_t)best_mvy;
    r.cost = best_cost; r.bitcost = best_bits;
    r.merged = (uint8_t)merged; r.merge_idx = (uint8_t)merge_idx;
    r.mv_cand_idx = (uint8_t)cu_mv_cand; r.took = took ? 1 : 0;
    *out = r;
  }
}

template <int G, int S>
__global__ void __launch_bounds__(kCtaThreads)
me_search_kernel(const __grid_constant__ MeFrame f, const kvz_cuda_me_pu *__restrict__ pus,
                 const kvz_cuda_me_job *__restrict__ jobs, kvz_cuda_me_result *__restrict__ results, int n_pus)
{
  constexpr int kGroups = kCtaThreads / G;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  GroupSmem<G, S> *all = reinterpret_cast<GroupSmem<G, S> *>(smem_raw);
  Group<G> g;
  g.tid = threadIdx.x % G;
  g.id = threadIdx.x / G;
  GroupSmem<G, S> &sm = all[g.id];
  const int pi = blockIdx.x * kGroups + g.id;
  if (pi >= n_pus) return;
  const kvz_cuda_me_pu &pu = pus[pi];
  const bool small = pu.w <= 16 && pu.h <= 16;
  if (small != (S == 16)) return;            // the other instantiation owns this PU

  // stage the PU's source block (pitch 64)
  for (int i = g.tid; i < pu.h * (pu.w >> 2); i += G) {
    const int r = i / (pu.w >> 2), c = i % (pu.w >> 2);
    const uint8_t *src = f.cur + (size_t)(f.tile_y + pu.y + r) * f.cur_pitch + f.tile_x + pu.x + c * 4;
    *reinterpret_cast<uint32_t *>(&sm.cur[r * 64 + c * 4]) = *reinterpret_cast<const uint32_t *>(src);
  }
  g.sync();
  uint32_t inter_cost = 0x7FFFFFFFu;         // *inter_cost = MAX_INT (search_inter.c:1460)
  for (int q = 0; q < pu.num_jobs; q++)
    search_one_ref<G, S>(g, sm, f, pu, jobs[pu.first_job + q], inter_cost, results + pu.first_job + q);
}

}  // namespace

int launch_me_search(const MeFrame &f, const kvz_cuda_me_pu *pus, const kvz_cuda_me_job *jobs,
                     kvz_cuda_me_result *results, int n_pus, cudaStream_t stream)
{
  if (n_pus <= 0) return 1;
  {
    using K = GroupSmem<32, 16>;
    const size_t smem = sizeof(K) * 4;
    auto kern = me_search_kernel<32, 16>;
    KVZ_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(n_pus + 3) / 4, kCtaThreads, smem, stream>>>(f, pus, jobs, results, n_pus);
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
  }
  {
    using K = GroupSmem<128, 64>;
    const size_t smem = sizeof(K);
    auto kern = me_search_kernel<128, 64>;
    KVZ_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<n_pus, kCtaThreads, smem, stream>>>(f, pus, jobs, results, n_pus);
    count_launch();
    KVZ_CUDA_LAUNCH_CHECK();
  }
  return 1;
}

}  // namespace kvz
