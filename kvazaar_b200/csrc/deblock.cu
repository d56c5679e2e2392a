// kvazaar_b200/csrc/deblock.cu
//
// The deblocking filter on a whole picture resident on the device: what
// kvz_filter_deblock_lcu (filter.c:770-779) leaves once it has run over every LCU.  The reference walks LCU by
// LCU (vertical edges of the LCU, the four postponed samples of the LCU to the left, horizontal edges) so that
// every horizontal-edge filter sees samples that all vertical-edge filters have already touched; the filters of
// one direction never overlap (edges lie on the 8x8 grid, a filter reads 4 and writes at most 3 samples either
// side; chroma: reads 2, writes 1 on the 8x8 chroma grid).  So the picture is two passes: ALL vertical edges,
// then ALL horizontal edges -- two launches, every 4-sample edge segment of a pass an independent thread:
//   * is_tu_boundary / is_pu_boundary (filter.c:190-244) on the SCU records,
//   * the boundary strength incl. the B-slice rules (filter.c:393-466), QP (get_qp_y_pred, filter.c:263-285),
//     beta / tc (HEVC tables 8-12),
//   * the on/off and strong/weak decisions from lines 0 and 3 and the luma filters (filter.c:83-153, :476-536),
//   * chroma: strength-2 edges on the 8x8 chroma grid (filter.c:554-615).
// Lanes of a warp run along x, so the rows a warp touches are contiguous.  In place.
#include "common.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kT = 128;

__device__ __constant__ uint8_t c_tc[54] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 1,
  2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 5, 5, 6, 6, 7, 8, 9, 10, 11, 13, 14, 16, 18, 20, 22, 24 };
__device__ __constant__ uint8_t c_beta[52] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 6, 7, 8, 9, 10, 11, 12, 13,
  14, 15, 16, 17, 18, 20, 22, 24, 26, 28, 30, 32, 34, 36, 38, 40, 42, 44, 46, 48, 50, 52, 54, 56, 58, 60, 62, 64 };
__device__ __constant__ uint8_t c_chroma_qp[58] = { 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19,
  20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 29, 30, 31, 32, 33, 33, 34, 34, 35, 35, 36, 36, 37, 37, 38, 39, 40, 41,
  42, 43, 44, 45, 46, 47, 48, 49, 50, 51 };
// cu.c:33-60: PUs per part mode and their origins in quarters of the CU width, packed x | y << 2 per PU
__device__ __constant__ uint8_t c_parts[8] = { 1, 2, 2, 4, 2, 2, 2, 2 };
__device__ __constant__ uint8_t c_part_ofs[8][4] = { { 0, 0, 0, 0 }, { 0, 8, 0, 0 }, { 0, 2, 0, 0 }, { 0, 2, 8, 10 },
                                                     { 0, 4, 0, 0 }, { 0, 12, 0, 0 }, { 0, 1, 0, 0 }, { 0, 3, 0, 0 } };

struct DbParams {
  kvz_cuda_deblock_cfg c;
  const kvz_cuda_scu *scus;
  uint8_t *pl[3];
  int pitch[3];
  int n_planes, bitdepth;
  int luma_items, chroma_items;       // of the pass being launched
};

__device__ __forceinline__ const kvz_cuda_scu &scu_at(const DbParams &p, int x, int y)
{
  return p.scus[(y >> 2) * (p.c.width >> 2) + (x >> 2)];
}

__device__ bool tu_edge(const DbParams &p, int x, int y, bool hor)
{
  const int tu = 64 >> scu_at(p, x, y).tr_depth;
  return ((hor ? y : x) & (tu - 1)) == 0;
}

__device__ bool pu_edge(const DbParams &p, int x, int y, bool hor)
{
  const int cw = 64 >> scu_at(p, x, y).depth;
  const int cx = x & ~(cw - 1), cy = y & ~(cw - 1);
  const int pm = scu_at(p, cx, cy).part_size & 7;
  bool hit = false;
  for (int i = 0; i < c_parts[pm]; i++) {
    const int o = c_part_ofs[pm][i];
    hit |= hor ? (cy + (o >> 2) * cw / 4 == y) : (cx + (o & 3) * cw / 4 == x);
  }
  return hit;
}

__device__ int edge_qp(const DbParams &p, int x, int y, bool hor)
{
  if (!p.c.cu_qp) return p.c.qp;
  const int qp_p = hor ? scu_at(p, x, y - 1).qp : scu_at(p, x - 1, y).qp;
  return (qp_p + scu_at(p, x, y).qp + 1) >> 1;
}

__device__ __forceinline__ bool far4(int a, int b) { return abs(a - b) >= 4; }

__device__ int strength(const DbParams &p, const kvz_cuda_scu &P, const kvz_cuda_scu &Q, bool tu)
{
  if (Q.type == 1 || P.type == 1) return 2;
  if (tu && (Q.cbf_y || P.cbf_y)) return 1;
  if (P.mv_dir != 3 && Q.mv_dir != 3) {
    const int lp = (P.mv_dir - 1) & 1, lq = (Q.mv_dir - 1) & 1;
    if (far4(Q.mv[lq][0], P.mv[lp][0]) || far4(Q.mv[lq][1], P.mv[lp][1])) return 1;
    if (Q.mv_ref[lq] != P.mv_ref[lp]) return 1;
  }
  if (!p.c.slice_b) return 0;
  int mp[2][2], mq[2][2], rp[2], rq[2];
#pragma unroll
  for (int l = 0; l < 2; l++) {
    const bool up = (P.mv_dir >> l) & 1, uq = (Q.mv_dir >> l) & 1;
    mp[l][0] = up ? P.mv[l][0] : 0; mp[l][1] = up ? P.mv[l][1] : 0;
    mq[l][0] = uq ? Q.mv[l][0] : 0; mq[l][1] = uq ? Q.mv[l][1] : 0;
    rp[l] = up ? p.c.ref_LX[l][P.mv_ref[l] & 15] : -1;
    rq[l] = uq ? p.c.ref_LX[l][Q.mv_ref[l] & 15] : -1;
  }
  if (!((rp[0] == rq[0] && rp[1] == rq[1]) || (rp[0] == rq[1] && rp[1] == rq[0]))) return 1;
  const bool straight = far4(mq[0][0], mp[0][0]) || far4(mq[0][1], mp[0][1]) || far4(mq[1][0], mp[1][0]) || far4(mq[1][1], mp[1][1]);
  const bool crossed = far4(mq[1][0], mp[0][0]) || far4(mq[1][1], mp[0][1]) || far4(mq[0][0], mp[1][0]) || far4(mq[0][1], mp[1][1]);
  if (rp[0] != rp[1]) return rp[0] == rq[0] ? straight : crossed;
  return straight && crossed;
}

__device__ __forceinline__ int clip3(int lo, int hi, int v) { return min(max(v, lo), hi); }

// one 4-sample luma segment whose first sample (q0 of line 0) is at (x, y); (ux, uy): where the reference
// evaluates the unit's TU / PU boundary test and QP
template <typename PIX, bool HOR>
__device__ void luma_segment(const DbParams &p, int x, int y, int ux, int uy)
{
  const bool tu = tu_edge(p, ux, uy, HOR);
  if (!tu && !pu_edge(p, ux, uy, HOR)) return;
  const kvz_cuda_scu &P = HOR ? scu_at(p, x, y - 1) : scu_at(p, x - 1, y);
  const int bs = strength(p, P, scu_at(p, x, y), tu);
  if (bs == 0) return;
  const int scale = 1 << (p.bitdepth - 8), maxv = (1 << p.bitdepth) - 1;
  const int qp = edge_qp(p, ux, uy, HOR);
  const int beta = c_beta[clip3(0, 51, qp + (p.c.beta_offset_div2 << 1))] * scale;
  const int tc = c_tc[clip3(0, 53, qp + 2 * (bs - 1) + (p.c.tc_offset_div2 << 1))] * scale;
  const int side = (beta + (beta >> 1)) >> 3;
  const int pitch = p.pitch[0] / (int)sizeof(PIX);
  PIX *e = reinterpret_cast<PIX *>(p.pl[0]) + (size_t)y * pitch + x;
  const int xs = HOR ? pitch : 1, ys = HOR ? 1 : pitch;
  int b[4][8];
#pragma unroll
  for (int l = 0; l < 4; l++)
#pragma unroll
    for (int i = 0; i < 8; i++) b[l][i] = e[(i - 4) * xs + l * ys];
  const int dp0 = abs(b[0][1] - 2 * b[0][2] + b[0][3]), dq0 = abs(b[0][4] - 2 * b[0][5] + b[0][6]);
  const int dp3 = abs(b[3][1] - 2 * b[3][2] + b[3][3]), dq3 = abs(b[3][4] - 2 * b[3][5] + b[3][6]);
  const int dp = dp0 + dp3, dq = dq0 + dq3;
  if (dp + dq >= beta) return;
  const bool strong = 2 * (dp0 + dq0) < (beta >> 2) && 2 * (dp3 + dq3) < (beta >> 2) &&
                      abs(b[0][3] - b[0][4]) < ((5 * tc + 1) >> 1) && abs(b[3][3] - b[3][4]) < ((5 * tc + 1) >> 1) &&
                      abs(b[0][0] - b[0][3]) + abs(b[0][4] - b[0][7]) < (beta >> 3) &&
                      abs(b[3][0] - b[3][3]) + abs(b[3][4] - b[3][7]) < (beta >> 3);
#pragma unroll
  for (int l = 0; l < 4; l++) {
    const int *m = b[l];
    PIX *o = e + l * ys;
    if (strong) {
      o[-3 * xs] = (PIX)clip3(m[1] - 2 * tc, m[1] + 2 * tc, (2 * m[0] + 3 * m[1] + m[2] + m[3] + m[4] + 4) >> 3);
      o[-2 * xs] = (PIX)clip3(m[2] - 2 * tc, m[2] + 2 * tc, (m[1] + m[2] + m[3] + m[4] + 2) >> 2);
      o[-1 * xs] = (PIX)clip3(m[3] - 2 * tc, m[3] + 2 * tc, (m[1] + 2 * m[2] + 2 * m[3] + 2 * m[4] + m[5] + 4) >> 3);
      o[0] = (PIX)clip3(m[4] - 2 * tc, m[4] + 2 * tc, (m[2] + 2 * m[3] + 2 * m[4] + 2 * m[5] + m[6] + 4) >> 3);
      o[1 * xs] = (PIX)clip3(m[5] - 2 * tc, m[5] + 2 * tc, (m[3] + m[4] + m[5] + m[6] + 2) >> 2);
      o[2 * xs] = (PIX)clip3(m[6] - 2 * tc, m[6] + 2 * tc, (m[3] + m[4] + m[5] + 3 * m[6] + 2 * m[7] + 4) >> 3);
    } else {
      int delta = (9 * (m[4] - m[3]) - 3 * (m[5] - m[2]) + 8) >> 4;
      if (abs(delta) < tc * 10) {
        const int tc2 = tc >> 1;
        delta = clip3(-tc, tc, delta);
        o[-1 * xs] = (PIX)clip3(0, maxv, m[3] + delta);
        o[0] = (PIX)clip3(0, maxv, m[4] - delta);
        if (dp < side) o[-2 * xs] = (PIX)clip3(0, maxv, m[2] + clip3(-tc2, tc2, (((m[1] + m[3] + 1) >> 1) - m[2] + delta) >> 1));
        if (dq < side) o[1 * xs] = (PIX)clip3(0, maxv, m[5] + clip3(-tc2, tc2, (((m[6] + m[4] + 1) >> 1) - m[5] - delta) >> 1));
      }
    }
  }
}

// one 4-sample chroma segment (both chroma planes) whose first sample is at chroma (x, y)
template <typename PIX, bool HOR>
__device__ void chroma_segment(const DbParams &p, int x, int y)
{
  const int lx = x << 1, ly = y << 1;
  if (!tu_edge(p, lx, ly, HOR) && !pu_edge(p, lx, ly, HOR)) return;
  const kvz_cuda_scu &P = HOR ? scu_at(p, lx, (y - 1) << 1) : scu_at(p, (x - 1) << 1, ly);
  if (scu_at(p, lx, ly).type != 1 && P.type != 1) return;
  const int scale = 1 << (p.bitdepth - 8), maxv = (1 << p.bitdepth) - 1;
  const int qp = c_chroma_qp[clip3(0, 57, edge_qp(p, lx, ly, HOR))];
  const int tc = c_tc[clip3(0, 53, qp + 2 + (p.c.tc_offset_div2 << 1))] * scale;
#pragma unroll
  for (int k = 1; k < 3; k++) {
    const int pitch = p.pitch[k] / (int)sizeof(PIX);
    const int off = HOR ? pitch : 1, step = HOR ? 1 : pitch;
    PIX *e = reinterpret_cast<PIX *>(p.pl[k]) + (size_t)y * pitch + x;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      PIX *s = e + i * step;
      const int m2 = s[-2 * off], m3 = s[-off], m4 = s[0], m5 = s[off];
      const int delta = clip3(-tc, tc, (((m4 - m3) * 4) + m2 - m5 + 4) >> 3);
      s[-off] = (PIX)clip3(0, maxv, m3 + delta);
      s[0] = (PIX)clip3(0, maxv, m4 - delta);
    }
  }
}

template <typename PIX, bool HOR>
__global__ void __launch_bounds__(kT) deblock_pass_kernel(const __grid_constant__ DbParams p)
{
  const int W = p.c.width, H = p.c.height;
  for (int it = blockIdx.x * kT + threadIdx.x; it < p.luma_items + p.chroma_items; it += gridDim.x * kT) {
    if (it < p.luma_items) {
      if (HOR) {
        // horizontal edges y = 8, 16, ...; segments x = 0, 4, ...
        const int per = W >> 2, x = (it % per) * 4, y = (it / per + 1) * 8;
        // the second half of the last unit of an LCU is filtered by filter_deblock_lcu_rightmost (filter.c:711-726),
        // which tests the boundary and takes the QP at its own first sample
        const bool postponed = (x & 63) == 60 && x + 4 != W;
        luma_segment<PIX, true>(p, x, y, postponed ? x : (x & ~7), y);
      } else {
        const int per = (W >> 3) - 1, x = (it % per + 1) * 8, y = (it / per) * 4;
        luma_segment<PIX, false>(p, x, y, x, y & ~7);
      }
    } else {
      const int ci = it - p.luma_items, CW = W >> 1;
      if (HOR) {
        const int per = CW >> 2, x = (ci % per) * 4, y = (ci / per + 1) * 8;
        chroma_segment<PIX, true>(p, x, y);
      } else {
        const int per = (CW >> 3) - (CW % 8 == 0 ? 1 : 0), x = (ci % per + 1) * 8, y = (ci / per) * 4;
        chroma_segment<PIX, false>(p, x, y);
      }
    }
  }
}

template <typename PIX, bool HOR>
int launch_pass(DbParams p, cudaStream_t stream)
{
  const int W = p.c.width, H = p.c.height, CW = W >> 1, CH = H >> 1;
  if (HOR) {
    p.luma_items = (W >> 2) * ((H >> 3) - 1);
    p.chroma_items = p.n_planes == 3 ? (CW >> 2) * ((CH + 7) / 8 - 1) : 0;
  } else {
    p.luma_items = ((W >> 3) - 1) * (H >> 2);
    p.chroma_items = p.n_planes == 3 ? ((CW >> 3) - (CW % 8 == 0 ? 1 : 0)) * (CH >> 2) : 0;
  }
  if (p.luma_items < 0) p.luma_items = 0;
  if (p.chroma_items < 0) p.chroma_items = 0;
  const int items = p.luma_items + p.chroma_items;
  if (items == 0) return 1;
  int grid = (items + kT - 1) / kT;
  if (grid > 148 * 16) grid = 148 * 16;
  deblock_pass_kernel<PIX, HOR><<<grid, kT, 0, stream>>>(p);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace

int launch_deblock_picture(const kvz_cuda_deblock_cfg &cfg, const kvz_cuda_scu *d_scus, const kvz_cuda_plane *planes,
                           int n_planes, int bitdepth, cudaStream_t stream)
{
  const char *who = "kvz_cuda_deblock_picture: bad plane";
  DbParams p;
  p.c = cfg; p.scus = d_scus; p.n_planes = n_planes; p.bitdepth = bitdepth;
  const int px = bitdepth > 8 ? 2 : 1;
  for (int k = 0; k < 3; k++) { p.pl[k] = nullptr; p.pitch[k] = 0; }
  for (int k = 0; k < n_planes; k++) {
    const int w = k ? cfg.width >> 1 : cfg.width, h = k ? cfg.height >> 1 : cfg.height;
    if (!planes[k].d_data || planes[k].width != w || planes[k].height != h || planes[k].pitch < w * px || (planes[k].pitch % px)) return fail_msg(who);
    p.pl[k] = static_cast<uint8_t *>(const_cast<void *>(planes[k].d_data));
    p.pitch[k] = planes[k].pitch;
  }
  if (bitdepth > 8) return launch_pass<uint16_t, false>(p, stream) && launch_pass<uint16_t, true>(p, stream);
  return launch_pass<uint8_t, false>(p, stream) && launch_pass<uint8_t, true>(p, stream);
}

}  // namespace kvz
