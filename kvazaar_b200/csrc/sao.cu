// kvazaar_b200/csrc/sao.cu
//
// Sample adaptive offset on whole pictures resident on the device:
//   * sao_stats_kernel: per CTU and colour plane, in ONE pass over the source and the deblocked reconstruction,
//     everything sao_search_best_mode (sao.c:465-575) asks the pixels for: the per-category sums / counts of
//     calc_sao_edge_dir (strategies/generic/sao-generic.c:90-117) for all four edge classes and the per-band sums /
//     counts of calc_sao_bands (sao.c:245-260).  kvz_sao_edge_ddistortion (sao-generic.c:51-81) and
//     kvz_sao_band_ddistortion (sao-generic.c:158-182) are exact functions of these statistics
//     (sum over categories of n * h^2 - 2 * h * E, the formula sao.c:381-388 itself uses), so the host's decision
//     (offsets, CABAC rate, merge candidates) needs no second pass over the pixels.
//   * sao_apply_kernel: kvz_sao_reconstruct (sao.c:277-337) + sao_reconstruct_color (sao-generic.c:120-155) for
//     every CTU of the picture: what encoder_sao_reconstruct (encoderstate.c:245-441) produces once every LCU has
//     been through it -- each sample filtered with the parameters of the CTU that contains it, neighbours taken
//     from the deblocked picture, an edge class leaving samples whose neighbour lies outside the picture alone.
// Both read / write 1.5 samples per pixel once: HBM-bound by construction, launch-latency-bound at 1080p.
#include "common.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kT = 256;
constexpr int kMaxPlanes = 3;

struct SaoPlane {
  const uint8_t *orig, *rec;    // stats: source and deblocked reconstruction; apply: rec = input
  uint8_t *out;                 // apply only
  int width, height, pitch_orig, pitch_rec, pitch_out;
  int ctu, ctus_x, n_ctus, first_cta;
};
struct SaoParams { SaoPlane pl[kMaxPlanes]; int n_planes, shift; };

// sao-generic.c:40-48: category of c against its neighbours a and b
__device__ __forceinline__ int eo_cat(int a, int b, int c)
{
  const int idx = 2 + ((c > a) - (c < a)) + ((c > b) - (c < b));
  return (0x43021 >> (4 * idx)) & 7;                   // { 1, 2, 0, 3, 4 }
}

template <typename PIX>
__global__ void __launch_bounds__(kT) sao_stats_kernel(const __grid_constant__ SaoParams p, kvz_cuda_sao_stats *out)
{
  __shared__ PIX s_rec[64 * 64];
  __shared__ int16_t s_diff[64 * 64];
  __shared__ int s_band[kT / 32][64];
  __shared__ int s_edge[40];
  int k = 0;
#pragma unroll
  for (int q = 1; q < kMaxPlanes; q++) k += (q < p.n_planes && (int)blockIdx.x >= p.pl[q].first_cta);
  const SaoPlane &pl = p.pl[k];
  const int ctu = blockIdx.x - pl.first_cta, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int x0 = (ctu % pl.ctus_x) * pl.ctu, y0 = (ctu / pl.ctus_x) * pl.ctu;
  const int bw = min(pl.ctu, pl.width - x0), bh = min(pl.ctu, pl.height - y0), n = bw * bh;

  for (int i = tid; i < (kT / 32) * 64; i += kT) (&s_band[0][0])[i] = 0;
  if (tid < 40) s_edge[tid] = 0;
  for (int i = tid; i < n; i += kT) {
    const int y = i / bw, x = i - y * bw;
    const int r = reinterpret_cast<const PIX *>(pl.rec + (size_t)(y0 + y) * pl.pitch_rec)[x0 + x];
    const int o = reinterpret_cast<const PIX *>(pl.orig + (size_t)(y0 + y) * pl.pitch_orig)[x0 + x];
    s_rec[i] = (PIX)r;
    s_diff[i] = (int16_t)(o - r);
  }
  __syncthreads();

  int sum[4][5], cnt[4][5];
#pragma unroll
  for (int e = 0; e < 4; e++)
#pragma unroll
    for (int c = 0; c < 5; c++) { sum[e][c] = 0; cnt[e][c] = 0; }
  for (int i = tid; i < n; i += kT) {
    const int y = i / bw, x = i - y * bw;
    const int c = s_rec[i], d = s_diff[i];
    const int band = c >> p.shift;
    atomicAdd(&s_band[warp][band], d);
    atomicAdd(&s_band[warp][32 + band], 1);
    if (x >= 1 && x < bw - 1 && y >= 1 && y < bh - 1) {       // the block's own border has no neighbours here
      const int cat[4] = { eo_cat(s_rec[i - 1], s_rec[i + 1], c), eo_cat(s_rec[i - bw], s_rec[i + bw], c),
                           eo_cat(s_rec[i - bw - 1], s_rec[i + bw + 1], c), eo_cat(s_rec[i - bw + 1], s_rec[i + bw - 1], c) };
#pragma unroll
      for (int e = 0; e < 4; e++)
#pragma unroll
        for (int q = 0; q < 5; q++) {
          const bool hit = cat[e] == q;
          sum[e][q] += hit ? d : 0;
          cnt[e][q] += hit ? 1 : 0;
        }
    }
  }
  kvz_cuda_sao_stats &dst = out[pl.first_cta + ctu];          // plane-major, CTUs in raster order
#pragma unroll
  for (int e = 0; e < 4; e++) {
#pragma unroll
    for (int q = 0; q < 5; q++) {
      int s = sum[e][q], c = cnt[e][q];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); c += __shfl_xor_sync(0xffffffffu, c, o); }
      if (lane == 0) {
        // ten values per class: [0][cat] sums, [1][cat] counts
        atomicAdd(&s_edge[e * 10 + q], s);
        atomicAdd(&s_edge[e * 10 + 5 + q], c);
      }
    }
  }
  __syncthreads();
  if (tid < 40) dst.edge[tid / 10][(tid % 10) / 5][tid % 5] = s_edge[tid];
  if (tid >= 64 && tid < 128) {
    const int b = tid - 64;
    int v = 0;
#pragma unroll
    for (int w = 0; w < kT / 32; w++) v += s_band[w][b];
    dst.band[b >> 5][b & 31] = v;
  }
}

template <typename PIX>
__global__ void __launch_bounds__(kT) sao_apply_kernel(const __grid_constant__ SaoParams p, const kvz_cuda_sao_params *luma,
                                                       const kvz_cuda_sao_params *chroma, int maxv)
{
  // CTA = one CTU-sized tile of one plane (64x64 luma, 32x32 chroma); thread = 4 samples of a row
  int k = 0;
#pragma unroll
  for (int q = 1; q < kMaxPlanes; q++) k += (q < p.n_planes && (int)blockIdx.x >= p.pl[q].first_cta);
  const SaoPlane &pl = p.pl[k];
  const int ctu = blockIdx.x - pl.first_cta;
  const int x0 = (ctu % pl.ctus_x) * pl.ctu, y0 = (ctu / pl.ctus_x) * pl.ctu;
  const int bw = min(pl.ctu, pl.width - x0), bh = min(pl.ctu, pl.height - y0);
  const kvz_cuda_sao_params sp = (k == 0 ? luma : chroma)[ctu];
  const int v = k == 2, type = sp.type;
  const int quads = (bw + 3) >> 2;
  int ax = 0, ay = 0, bx = 0, by = 0;
  if (type == 2) {
    const int e = sp.eo_class & 3;
    ax = e == 1 ? 0 : (e == 3 ? 1 : -1); ay = e == 0 ? 0 : -1;
    bx = -ax; by = -ay;
  }
  const int pos = sp.band_position[v];
  for (int i = threadIdx.x; i < quads * bh; i += kT) {
    const int yy = i / quads, xq = (i - yy * quads) * 4, y = y0 + yy;
    const PIX *row = reinterpret_cast<const PIX *>(pl.rec + (size_t)y * pl.pitch_rec);
    PIX *orow = reinterpret_cast<PIX *>(pl.out + (size_t)y * pl.pitch_out);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int x = x0 + xq + j;
      if (xq + j >= bw) break;
      const int c = row[x];
      int r = c;
      if (type == 1) {
        const int band = c >> p.shift;
        if (band >= pos && band < pos + 4) r = min(max(c + sp.offsets[band - pos + 1 + 5 * v], 0), maxv);
      } else if (type == 2) {
        const int xa = x + ax, ya = y + ay, xb = x + bx, yb = y + by;
        if (xa >= 0 && xa < pl.width && xb >= 0 && xb < pl.width && ya >= 0 && ya < pl.height && yb >= 0 && yb < pl.height) {
          const int a = reinterpret_cast<const PIX *>(pl.rec + (size_t)ya * pl.pitch_rec)[xa];
          const int b = reinterpret_cast<const PIX *>(pl.rec + (size_t)yb * pl.pitch_rec)[xb];
          r = min(max(c + sp.offsets[eo_cat(a, b, c) + 5 * v], 0), maxv);
        }
      }
      orow[x] = (PIX)r;
    }
  }
}

int fill(SaoParams &p, const kvz_cuda_plane *a, const kvz_cuda_plane *b, const kvz_cuda_plane *o, int n, int bitdepth,
         const char *who)
{
  if (!a || !b || n < 1 || n > kMaxPlanes) return fail_msg(who);
  const int px = bitdepth > 8 ? 2 : 1;
  p.n_planes = n; p.shift = bitdepth - 5;
  int ctas = 0;
  for (int k = 0; k < kMaxPlanes; k++) { p.pl[k] = SaoPlane{}; p.pl[k].first_cta = 0x7fffffff; }
  for (int k = 0; k < n; k++) {
    if (!a[k].d_data || !b[k].d_data || a[k].width <= 0 || a[k].height <= 0 || a[k].pitch < a[k].width * px ||
        b[k].width != a[k].width || b[k].height != a[k].height || b[k].pitch < b[k].width * px) return fail_msg(who);
    if (o && (!o[k].d_data || o[k].width != a[k].width || o[k].height != a[k].height || o[k].pitch < o[k].width * px ||
              o[k].d_data == b[k].d_data)) return fail_msg(who);
    SaoPlane &pl = p.pl[k];
    pl.orig = static_cast<const uint8_t *>(a[k].d_data); pl.rec = static_cast<const uint8_t *>(b[k].d_data);
    pl.out = o ? static_cast<uint8_t *>(const_cast<void *>(o[k].d_data)) : nullptr;
    pl.width = a[k].width; pl.height = a[k].height;
    pl.pitch_orig = a[k].pitch; pl.pitch_rec = b[k].pitch; pl.pitch_out = o ? o[k].pitch : 0;
    pl.ctu = k == 0 ? 64 : 32;                                // LCU_WIDTH, LCU_WIDTH_C (4:2:0)
    pl.ctus_x = (pl.width + pl.ctu - 1) / pl.ctu;
    pl.n_ctus = pl.ctus_x * ((pl.height + pl.ctu - 1) / pl.ctu);
    pl.first_cta = ctas;
    ctas += pl.n_ctus;
  }
  return ctas;
}

}  // namespace

int launch_sao_stats(const kvz_cuda_plane *orig, const kvz_cuda_plane *rec, int n_planes, int bitdepth,
                     kvz_cuda_sao_stats *d_stats, cudaStream_t stream)
{
  SaoParams p;
  const int ctas = fill(p, orig, rec, nullptr, n_planes, bitdepth, "kvz_cuda_sao_ctu_stats: bad plane");
  if (!ctas) return 0;
  if (bitdepth > 8) sao_stats_kernel<uint16_t><<<ctas, kT, 0, stream>>>(p, d_stats);
  else sao_stats_kernel<uint8_t><<<ctas, kT, 0, stream>>>(p, d_stats);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

int launch_sao_apply(const kvz_cuda_plane *rec_in, const kvz_cuda_plane *rec_out, int n_planes, int bitdepth,
                     const kvz_cuda_sao_params *d_luma, const kvz_cuda_sao_params *d_chroma, cudaStream_t stream)
{
  SaoParams p;
  const int ctas = fill(p, rec_in, rec_in, rec_out, n_planes, bitdepth, "kvz_cuda_sao_reconstruct: bad plane");
  if (!ctas) return 0;
  const int maxv = (1 << bitdepth) - 1;
  if (bitdepth > 8) sao_apply_kernel<uint16_t><<<ctas, kT, 0, stream>>>(p, d_luma, d_chroma, maxv);
  else sao_apply_kernel<uint8_t><<<ctas, kT, 0, stream>>>(p, d_luma, d_chroma, maxv);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
