// kvazaar_b200/csrc/inter_cand.cu
//
// Inter candidate derivation on device: fills the records kvz_cuda_me_search consumes --
// the merge list of every PU (kvz_inter_get_merge_cand, inter.c:1314-1446) and, for every (PU, reference) job,
// the two AMVP predictors (kvz_inter_get_mv_cand, inter.c:1209-1228 -> get_mv_cand_from_candidates :1102-1194)
// and the colocated start vector of search_pu_inter_ref (search_inter.c:1188-1206) -- from frame-wide SCU record
// arrays: the current picture's (the device mirror of frame->cu_array / the working lcu_t) and the references'
// (frame->ref->cu_arrays).  Positions: spatial a0 a1 b0 b1 b2 (inter.c:799-878) with the z-order availability
// of a0 / b0 (:566-697), temporal H / C3 (:713-797); MV scaling by POC distance (:955-1009).
//
// One thread per PU: the spatial / temporal neighbours are fetched once and serve the merge list and all of the
// PU's jobs.  Pure integer index work on 20-byte records: latency-bound, a few hundred instructions per PU.
#include "common.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kT = 128;

struct Cands { const kvz_cuda_scu *a[2], *b[3], *c3, *h; };

__device__ __forceinline__ const kvz_cuda_scu *scu_at(const kvz_cuda_cand_frame &f, int x, int y)
{
  return &f.d_scus[(y >> 2) * (f.width >> 2) + (x >> 2)];
}

__device__ __forceinline__ int lowbit_min(int w, int h) { return min(w & -w, h & -h); }

__device__ bool a0_coded(int x, int y, int w, int h)
{
  int size = lowbit_min(w, h);
  if (h != size) y = y + h - size;
  while (size < 64) {
    const int parent = 2 * size;
    const int idx = (x % parent != 0) + 2 * (y % parent != 0);
    if (idx == 0) return true;
    if (idx != 2) return false;
    y -= size; size = parent;
  }
  return false;
}

__device__ bool b0_coded(int x, int y, int w, int h)
{
  int size = lowbit_min(w, h);
  if (w != size) x = x + w - size;
  while (size < 64) {
    const int parent = 2 * size;
    const int idx = (x % parent != 0) + 2 * (y % parent != 0);
    if (idx == 0 || idx == 2) return true;
    if (idx == 3) return false;
    x -= size; size = parent;
  }
  return true;
}

__device__ void spatial(const kvz_cuda_cand_frame &f, int x, int y, int w, int h, Cands &c)
{
  const int xl = x & 63, yl = y & 63;
  if (x != 0) {
    const kvz_cuda_scu *a1 = scu_at(f, x - 1, y + h - 1);
    if (a1->type == 2) c.a[1] = a1;
    if (yl + h < 64 && y + h < f.height) {
      const kvz_cuda_scu *a0 = scu_at(f, x - 1, y + h);
      if (a0->type == 2 && a0_coded(x, y, w, h)) c.a[0] = a0;
    }
  }
  if (y != 0) {
    if (x + w < f.width && (xl + w < 64 || yl == 0)) {
      const kvz_cuda_scu *b0 = scu_at(f, x + w, y - 1);
      if (b0->type == 2 && b0_coded(x, y, w, h)) c.b[0] = b0;
    }
    const kvz_cuda_scu *b1 = scu_at(f, x + w - 1, y - 1);
    if (b1->type == 2) c.b[1] = b1;
    if (x != 0) {
      const kvz_cuda_scu *b2 = scu_at(f, x - 1, y - 1);
      if (b2->type == 2) c.b[2] = b2;
    }
  }
}

__device__ void temporal(const kvz_cuda_cand_frame &f, int x, int y, int w, int h, Cands &c)
{
  c.c3 = c.h = nullptr;
  if (!f.n_refs || f.ref_LX_size[0] == 0) return;
  const kvz_cuda_scu *col = f.d_ref_scus[f.ref_LX[0][0] & 15];
  const int per_row = f.ref_scu_pitch;
  const unsigned xbr = x + w, ybr = y + h;
  if (xbr < (unsigned)f.in_width && ybr < (unsigned)f.in_height && ybr % 64 != 0) {
    const kvz_cuda_scu *s = &col[((xbr >> 4) << 2) + ((ybr >> 4) << 2) * per_row];
    if (s->type == 2) c.h = s;
  }
  const unsigned xc = x + w / 2, yc = y + h / 2;
  if (xc < (unsigned)f.in_width && yc < (unsigned)f.in_height) {
    const kvz_cuda_scu *s = &col[((xc >> 4) << 2) + ((yc >> 4) << 2) * per_row];
    if (s->type == 2) c.c3 = s;
  }
}

__device__ __forceinline__ int clip3i(int lo, int hi, int v) { return min(max(v, lo), hi); }

__device__ __forceinline__ int16_t scaled_mv(int mv, int scale)
{
  const int s = scale * mv;
  return (int16_t)clip3i(-32768, 32767, (s + 127 + (s < 0)) >> 8);
}

__device__ void scale_by_pocs(int cur_poc, int cur_ref_poc, int nb_poc, int nb_ref_poc, int16_t mv[2])
{
  int dc = cur_poc - cur_ref_poc, dn = nb_poc - nb_ref_poc;
  if (dc == dn) return;
  dc = clip3i(-128, 127, dc); dn = clip3i(-128, 127, dn);
  const int scale = clip3i(-4096, 4095, (dc * ((0x4000 + (abs(dn) >> 1)) / dn) + 32) >> 6);
  mv[0] = scaled_mv(mv[0], scale);
  mv[1] = scaled_mv(mv[1], scale);
}

__device__ bool temporal_mv(const kvz_cuda_cand_frame &f, int current_ref, const kvz_cuda_scu *col, int reflist, int16_t out[2])
{
  if (!col || f.ref_LX_size[0] == 0) return false;
  const int col_ref = f.ref_LX[0][0] & 15;
  int col_list = reflist;
  for (int i = 0; i < f.n_refs; i++)
    if (f.ref_pocs[i] > f.poc) { col_list = 1; break; }
  if ((col->mv_dir & (col_list + 1)) == 0) col_list = 1 - col_list;
  out[0] = col->mv[col_list][0]; out[1] = col->mv[col_list][1];
  scale_by_pocs(f.poc, f.ref_pocs[current_ref & 15], f.ref_pocs[col_ref], f.col_ref_pocs[col_list][col->mv_ref[col_list] & 15], out);
  return true;
}

__device__ bool mvp(const kvz_cuda_cand_frame &f, int cur_idx, const kvz_cuda_scu *cand, int reflist, bool scaling, int16_t out[2])
{
  if (!cand) return false;
  for (int i = 0; i < 2; i++) {
    const int cl = i == 0 ? reflist : !reflist;
    if ((cand->mv_dir & (1 << cl)) == 0) continue;
    const int cand_pic = f.ref_LX[cl][cand->mv_ref[cl] & 15] & 15, cur_pic = f.ref_LX[reflist][cur_idx & 15] & 15;
    if (scaling) {
      out[0] = cand->mv[cl][0]; out[1] = cand->mv[cl][1];
      scale_by_pocs(f.poc, f.ref_pocs[cur_pic], f.poc, f.ref_pocs[cand_pic], out);
      return true;
    }
    if (f.ref_LX[cl][cand->mv_ref[cl] & 15] == f.ref_LX[reflist][cur_idx & 15]) {
      out[0] = cand->mv[cl][0]; out[1] = cand->mv[cl][1];
      return true;
    }
  }
  return false;
}

__device__ void amvp(const kvz_cuda_cand_frame &f, const Cands &c, int reflist, int lx_idx, int16_t mv_cand[2][2])
{
  int n = 0, nb = 0;
  for (int i = 0; i < 2; i++) if (mvp(f, lx_idx, c.a[i], reflist, false, mv_cand[n])) { n++; break; }
  if (n == 0) for (int i = 0; i < 2; i++) if (mvp(f, lx_idx, c.a[i], reflist, true, mv_cand[n])) { n++; break; }
  for (int i = 0; i < 3; i++) if (mvp(f, lx_idx, c.b[i], reflist, false, mv_cand[n])) { nb++; break; }
  n += nb;
  if (c.a[0] || c.a[1]) nb = 1; else if (n != 2) nb = 0;
  if (!nb) for (int i = 0; i < 3; i++) if (mvp(f, lx_idx, c.b[i], reflist, true, mv_cand[n])) { n++; break; }
  if (n == 2 && mv_cand[0][0] == mv_cand[1][0] && mv_cand[0][1] == mv_cand[1][1]) n = 1;
  if (f.tmvp_enable && f.poc > 1 && f.n_refs && n < 2 && (c.h || c.c3) &&
      temporal_mv(f, f.ref_LX[reflist][lx_idx & 15], c.h ? c.h : c.c3, reflist, mv_cand[n])) n++;
  while (n < 2) { mv_cand[n][0] = 0; mv_cand[n][1] = 0; n++; }
}

__device__ bool same_motion(const kvz_cuda_scu *p, const kvz_cuda_scu *q)
{
  if (!q) return false;
  if (p->mv_dir != q->mv_dir) return false;
  for (int l = 0; l < 2; l++)
    if ((p->mv_dir & (1 << l)) && (p->mv[l][0] != q->mv[l][0] || p->mv[l][1] != q->mv[l][1] || p->mv_ref[l] != q->mv_ref[l])) return false;
  return true;
}

// the unused list reads as inter_clear_cu_unused (inter.c:546-555) leaves it: mv 0, ref 255
__device__ bool put_merge(const kvz_cuda_scu *cand, const kvz_cuda_scu *d1, const kvz_cuda_scu *d2, kvz_cuda_merge_cand &out)
{
  if (!cand || same_motion(cand, d1) || same_motion(cand, d2)) return false;
  for (int l = 0; l < 2; l++) {
    const bool used = (cand->mv_dir >> l) & 1;
    out.mv[l][0] = used ? cand->mv[l][0] : 0;
    out.mv[l][1] = used ? cand->mv[l][1] : 0;
    out.ref[l] = used ? cand->mv_ref[l] : 255;
  }
  out.dir = cand->mv_dir;
  return true;
}

__device__ void merge_list(const kvz_cuda_cand_frame &f, Cands c, int x, int y, int w, int h, bool use_a1, bool use_b1,
                           kvz_cuda_merge_cand out[5])
{
  for (int i = 0; i < 5; i++) { out[i].dir = 0; out[i].ref[0] = out[i].ref[1] = 0; out[i].reserved = 0;
                                out[i].mv[0][0] = out[i].mv[0][1] = out[i].mv[1][0] = out[i].mv[1][1] = 0; }
  if (!use_a1) c.a[1] = nullptr;
  if (!use_b1) c.b[1] = nullptr;
  int n = 0;
  if (put_merge(c.a[1], nullptr, nullptr, out[n])) n++;
  if (put_merge(c.b[1], c.a[1], nullptr, out[n])) n++;
  if (put_merge(c.b[0], c.b[1], nullptr, out[n])) n++;
  if (put_merge(c.a[0], c.a[1], nullptr, out[n])) n++;
  if (n < 4 && put_merge(c.b[2], c.a[1], c.b[1], out[n])) n++;

  if (f.tmvp_enable && n < 5 && f.n_refs) {
    out[n].dir = 0;
    for (int l = 0; l <= (f.slice_b ? 1 : 0); l++) {
      if (temporal_mv(f, f.ref_LX[l][0], c.h ? c.h : c.c3, l, out[n].mv[l])) {
        out[n].ref[l] = 0;
        out[n].dir |= (uint8_t)(1 << l);
      }
    }
    if (out[n].dir != 0) n++;
  }
  if (n < 5 && f.slice_b) {
    // combined bi-predictive candidates (inter.c:1386-1417); pairs packed i | j << 2
    const uint8_t pairs[12] = { 0 | 1 << 2, 1 | 0 << 2, 0 | 2 << 2, 2 | 0 << 2, 1 | 2 << 2, 2 | 1 << 2,
                                0 | 3 << 2, 3 | 0 << 2, 1 | 3 << 2, 3 | 1 << 2, 2 | 3 << 2, 3 | 2 << 2 };
    const int cutoff = n;
    for (int idx = 0; idx < cutoff * (cutoff - 1) && n != 5; idx++) {
      const int i = pairs[idx] & 3, j = pairs[idx] >> 2;
      if (i >= n || j >= n) break;
      if ((out[i].dir & 1) && (out[j].dir & 2)) {
        out[n].dir = 3;
        out[n].mv[0][0] = out[i].mv[0][0]; out[n].mv[0][1] = out[i].mv[0][1];
        out[n].mv[1][0] = out[j].mv[1][0]; out[n].mv[1][1] = out[j].mv[1][1];
        out[n].ref[0] = out[i].ref[0]; out[n].ref[1] = out[j].ref[1];
        if (!(f.ref_LX[0][out[i].ref[0] & 15] == f.ref_LX[1][out[j].ref[1] & 15] &&
              out[i].mv[0][0] == out[j].mv[1][0] && out[i].mv[0][1] == out[j].mv[1][1])) n++;
      }
    }
  }
  int num_ref = f.n_refs;
  if (n < 5 && f.slice_b) {
    int neg = 0, pos = 0;
    for (int j = 0; j < f.n_refs; j++) { if (f.ref_pocs[j] < f.poc) neg++; else pos++; }
    num_ref = min(neg, pos);
  }
  int zero_idx = 0;
  while (n != 5) {
    out[n].mv[0][0] = out[n].mv[0][1] = out[n].mv[1][0] = out[n].mv[1][1] = 0;
    out[n].ref[0] = (zero_idx >= num_ref - 1) ? 0 : (uint8_t)zero_idx;
    out[n].ref[1] = out[n].ref[0];
    out[n].dir = f.slice_b ? 3 : 1;
    zero_idx++;
    n++;
  }
}

__global__ void __launch_bounds__(kT)
inter_cand_kernel(const __grid_constant__ kvz_cuda_cand_frame f, kvz_cuda_me_pu *pus, int n_pus, const uint8_t *pu_flags,
                  kvz_cuda_me_job *jobs)
{
  const int i = blockIdx.x * kT + threadIdx.x;
  if (i >= n_pus) return;
  kvz_cuda_me_pu &pu = pus[i];
  const int x = pu.x, y = pu.y, w = pu.w, h = pu.h;
  Cands c = { { nullptr, nullptr }, { nullptr, nullptr, nullptr }, nullptr, nullptr };
  spatial(f, x, y, w, h, c);
  temporal(f, x, y, w, h, c);
  const int flags = pu_flags ? pu_flags[i] : 0;
  kvz_cuda_merge_cand mc[5];
  merge_list(f, c, x, y, w, h, !(flags & 1), !(flags & 2), mc);
#pragma unroll
  for (int k = 0; k < 5; k++) pu.merge_cand[k] = mc[k];
  pu.num_merge_cand = 5;                                      // MRG_MAX_NUM_CANDS: the list is always filled up
  for (int q = 0; q < pu.num_jobs; q++) {
    kvz_cuda_me_job &jb = jobs[pu.first_job + q];
    const int ref_idx = jb.ref_idx;
    // which list holds this picture, and at which index (search_inter.c:1142-1162)
    int reflist = -1, lx = 0;
    const int lmax = max((int)f.ref_LX_size[0], (int)f.ref_LX_size[1]);
    for (lx = 0; lx < lmax; lx++) {
      if (lx < f.ref_LX_size[0] && f.ref_LX[0][lx] == ref_idx) { reflist = 0; break; }
      if (lx < f.ref_LX_size[1] && f.ref_LX[1][lx] == ref_idx) { reflist = 1; break; }
    }
    int16_t mvc[2][2] = { { 0, 0 }, { 0, 0 } };
    if (reflist >= 0) amvp(f, c, reflist, lx, mvc);
    jb.mv_cand[0][0] = mvc[0][0]; jb.mv_cand[0][1] = mvc[0][1];
    jb.mv_cand[1][0] = mvc[1][0]; jb.mv_cand[1][1] = mvc[1][1];
    // colocated start vector (search_inter.c:1188-1206)
    int ex = 0, ey = 0;
    {
      const int mx = f.tile_offset_x + x + (w >> 1), my = f.tile_offset_y + y + (h >> 1);
      const kvz_cuda_scu &s = f.d_ref_scus[ref_idx & 15][(my >> 2) * f.ref_scu_pitch + (mx >> 2)];
      if (s.type == 2) {
        const int l = (s.mv_dir & 1) ? 0 : 1;
        ex = s.mv[l][0]; ey = s.mv[l][1];
      }
    }
    jb.extra_mv[0] = (int16_t)ex; jb.extra_mv[1] = (int16_t)ey;
  }
}

}  // namespace

int launch_inter_candidates(const kvz_cuda_cand_frame &f, kvz_cuda_me_pu *d_pus, int n_pus, const uint8_t *d_pu_flags,
                            kvz_cuda_me_job *d_jobs, cudaStream_t stream)
{
  if (n_pus == 0) return 1;
  inter_cand_kernel<<<(n_pus + kT - 1) / kT, kT, 0, stream>>>(f, d_pus, n_pus, d_pu_flags, d_jobs);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
