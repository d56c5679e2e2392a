// kvazaar_b200/csrc/intra.cu
//
// Batched intra prediction and rough mode costs: for every PU of a batch, all 35 HEVC
// modes are predicted from the PU's reference samples and costed against the source
// block with SATD and SAD -- the arithmetic of search_intra_rough (search_intra.c:404-545)
// for every mode it may probe, so the host keeps only its coarse-to-fine ladder.
//   kvz_intra_predict          intra.c:281-332 (reference choice, DC, boundary filters)
//   intra_filter_reference     intra.c:163-193 ([1 2 1]/4 smoothing)
//   kvz_angular_pred           intra-generic.c:37-151
//   kvz_intra_pred_planar      intra-generic.c:153-189
//   satd_NxN / sad_NxN         picture-generic.c:189-196,343-346,460-486
// One CTA per PU; each of its 4 warps owns every 4th mode: it writes the prediction into
// its own shared-memory block and costs it with warp-level Hadamard butterflies.
#include "common.cuh"
#include "satd.cuh"
#include "../../include/kvz_cuda.h"

namespace kvz {
namespace {

constexpr int kWarps = 4;

// PIX = kvz_pixel of the build: uint8_t, or uint16_t for KVZ_BIT_DEPTH > 8
template <typename PIX>
struct Smem {
  PIX top[65], left[65], ftop[65], fleft[65];
  alignas(16) PIX orig[32 * 32];
  alignas(16) PIX pred[kWarps][32 * 32];
};

__device__ __constant__ int8_t c_disp[9] = { 0, 2, 5, 9, 13, 17, 21, 26, 32 };
__device__ __constant__ int16_t c_inv_disp[9] = { 0, 4096, 1638, 910, 630, 482, 390, 315, 256 };

// sample i (block coordinate, >= most negative index) of the main reference after the side
// reference has been projected onto its negative half (intra-generic.c:73-95)
template <typename PIX>
__device__ __forceinline__ int main_ref(const PIX *main_in, const PIX *side_in, int i, int inv)
{
  return i >= -1 ? main_in[i + 1] : side_in[(128 + inv * (-i - 1)) >> 8];
}

// the calling group strides the samples by `step`: 32 lanes of a warp, or 1 = a single thread predicts the block
template <typename PIX>
__device__ void predict_mode(const Smem<PIX> &sm, int lg, int mode, bool chroma, bool filter_boundary, PIX *dst, int lane, int pix_max,
                             int step = 32)
{
  const int w = 1 << lg;
  // which reference (intra.c:291-310)
  bool filtered = false;
  if (!chroma && mode != 1 && w != 4) {
    if (mode == 0) filtered = true;
    else {
      const int thres = lg == 3 ? 7 : (lg == 4 ? 1 : 0);
      filtered = min(abs(mode - 26), abs(mode - 10)) > thres;
    }
  }
  const PIX *top = filtered ? sm.ftop : sm.top, *left = filtered ? sm.fleft : sm.left;

  if (mode == 0) {
    const int tr = top[w + 1], bl = left[w + 1];
    for (int i = lane; i < w * w; i += step) {
      const int y = i >> lg, x = i & (w - 1);
      const int hor = (w - 1 - x) * left[y + 1] + (x + 1) * tr;
      const int ver = (w - 1 - y) * top[x + 1] + (y + 1) * bl;
      dst[i] = (PIX)((hor + ver + w) >> (lg + 1));
    }
  } else if (mode == 1) {
    int sum = 0;
    for (int i = lane; i < w; i += step) sum += top[i + 1] + left[i + 1];
    if (step == 32) sum = (int)warp_sum((uint32_t)sum);
    else if (step == 8) {                                     // the 8 lanes that share this mode (other groups of the warp are elsewhere)
      const unsigned gm = 0xFFu << ((threadIdx.x & 31) & ~7);
      sum += __shfl_xor_sync(gm, sum, 4); sum += __shfl_xor_sync(gm, sum, 2); sum += __shfl_xor_sync(gm, sum, 1);
    }
    const int dc = (sum + w) >> (lg + 1);
    const bool edge = !chroma && w < 32;                     // intra_pred_filtered_dc (intra.c:246-279)
    for (int i = lane; i < w * w; i += step) {
      const int y = i >> lg, x = i & (w - 1);
      int v = dc;
      if (edge) {
        if (x == 0 && y == 0) v = (left[1] + 2 * dc + top[1] + 2) / 4;
        else if (y == 0) v = (top[x + 1] + 3 * dc + 2) / 4;
        else if (x == 0) v = (left[y + 1] + 3 * dc + 2) / 4;
      }
      dst[i] = (PIX)v;
    }
  } else {
    const bool vertical = mode >= 18;
    const int md = vertical ? mode - 26 : 10 - mode;
    const int disp = (md < 0 ? -1 : 1) * c_disp[abs(md)];
    const int inv = c_inv_disp[abs(md)];
    const PIX *main_in = vertical ? top : left, *side_in = vertical ? left : top;
    const bool post = !chroma && w < 32 && filter_boundary && (mode == 10 || mode == 26);
    for (int i = lane; i < w * w; i += step) {
      // (x, y) in the orientation of a vertical mode; horizontal modes are its transpose
      const int y = i >> lg, x = i & (w - 1);
      const int pos = (y + 1) * disp, di = pos >> 5, fr = pos & 31;
      int v;
      if (disp < 0) {
        const int a = main_ref(main_in, side_in, x + di, inv);
        v = fr ? ((32 - fr) * a + fr * main_ref(main_in, side_in, x + di + 1, inv) + 16) >> 5 : a;
      } else {
        const int a = main_in[x + di + 1];
        v = fr ? ((32 - fr) * a + fr * main_in[x + di + 2] + 16) >> 5 : a;
      }
      // pure horizontal / vertical: first row / column follows the gradient of the other edge (intra.c:195-207)
      if (post && x == 0) v = clampi(v + (((int)side_in[y + 1] - (int)side_in[0]) >> 1), 0, pix_max);
      dst[vertical ? i : x * w + y] = (PIX)v;
    }
  }
}

template <typename PIX>
__global__ void __launch_bounds__(kWarps * 32)
intra_costs_kernel(const PIX *__restrict__ refs, const PIX *__restrict__ orig_base,
                   const kvz_cuda_intra_job *__restrict__ jobs, uint32_t *__restrict__ costs, PIX *__restrict__ preds, int pix_max,
                   int cost_shift)
{
  __shared__ Smem<PIX> sm;
  const kvz_cuda_intra_job jb = jobs[blockIdx.x];
  const int lg = jb.log2_width, w = 1 << lg, n = 2 * w + 1;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const PIX *rt = refs + jb.ref_offset, *rl = rt + n;
  for (int i = tid; i < n; i += kWarps * 32) { sm.top[i] = rt[i]; sm.left[i] = rl[i]; }
  if (jb.has_orig)
    for (int i = tid; i < w * w; i += kWarps * 32)
      sm.orig[i] = orig_base[jb.orig_offset + (size_t)(i >> lg) * jb.orig_stride + (i & (w - 1))];
  __syncthreads();
  for (int i = tid; i < n; i += kWarps * 32) {                // intra_filter_reference (intra.c:163-193)
    if (i == 0) sm.ftop[0] = sm.fleft[0] = (PIX)((sm.left[1] + 2 * sm.left[0] + sm.top[1] + 2) / 4);
    else if (i == n - 1) { sm.ftop[i] = sm.top[i]; sm.fleft[i] = sm.left[i]; }
    else {
      sm.ftop[i] = (PIX)((sm.top[i - 1] + 2 * sm.top[i] + sm.top[i + 1] + 2) / 4);
      sm.fleft[i] = (PIX)((sm.left[i - 1] + 2 * sm.left[i] + sm.left[i + 1] + 2) / 4);
    }
  }
  __syncthreads();
  if (lg == 2) {
    // 4x4: one THREAD per mode (a warp per mode would leave 31 lanes idle in the 4x4 Hadamard): 35 threads predict
    // their 16 samples, transform and sum on their own
    if (tid < 35) {
      const int mode = tid;
      PIX *mine = &sm.pred[0][0] + mode * 16;
      predict_mode<PIX>(sm, lg, mode, jb.is_chroma != 0, jb.filter_boundary != 0, mine, 0, pix_max, 1);
      if (preds)
        for (int i = 0; i < 16; i++) preds[(size_t)jb.pred_offset + (size_t)mode * 16 + i] = mine[i];
      if (jb.has_orig) {
        int d[16];
        uint32_t sad = 0;
#pragma unroll
        for (int i = 0; i < 16; i++) { d[i] = (int)mine[i] - (int)sm.orig[i]; sad += (uint32_t)abs(d[i]); }
        costs[((size_t)blockIdx.x * 35 + mode) * 2] = satd4_thread(d);          // satd_4x4: not shifted
        costs[((size_t)blockIdx.x * 35 + mode) * 2 + 1] = sad >> cost_shift;
      }
    }
    return;
  }
  if (lg == 3) {
    // 8x8: four modes per warp, EIGHT lanes per mode (one row of the block each: what the 8x8 Hadamard needs)
    const int sub = lane >> 3, lane8 = lane & 7;
    PIX *mine = sm.pred[warp] + sub * 64;
    for (int base = 0; base < 35; base += 4 * kWarps) {
      const int mode = base + warp * 4 + sub;
      const bool on = mode < 35;
      __syncwarp();
      if (on) predict_mode<PIX>(sm, lg, mode, jb.is_chroma != 0, jb.filter_boundary != 0, mine, lane8, pix_max, 8);
      __syncwarp();
      if (on && preds)
        for (int i = lane8; i < 64; i += 8) preds[(size_t)jb.pred_offset + (size_t)mode * 64 + i] = mine[i];
      if (jb.has_orig) {
        int d[8];
        uint32_t sad = 0;
#pragma unroll
        for (int x = 0; x < 8; x++) {
          d[x] = on ? (int)mine[lane8 * 8 + x] - (int)sm.orig[lane8 * 8 + x] : 0;
          sad += (uint32_t)abs(d[x]);
        }
        const uint32_t satd = satd8_rows(d, lane8);            // all 32 lanes take part
        sad += __shfl_xor_sync(0xffffffffu, sad, 1); sad += __shfl_xor_sync(0xffffffffu, sad, 2); sad += __shfl_xor_sync(0xffffffffu, sad, 4);
        if (on && lane8 == 0) {
          costs[((size_t)blockIdx.x * 35 + mode) * 2] = satd >> cost_shift;
          costs[((size_t)blockIdx.x * 35 + mode) * 2 + 1] = sad >> cost_shift;
        }
      }
    }
    return;
  }
  PIX *dst = sm.pred[warp];
  const StridedBlockT<PIX> a{ dst, w }, b{ sm.orig, w };
  for (int mode = warp; mode < 35; mode += kWarps) {
    __syncwarp();
    predict_mode<PIX>(sm, lg, mode, jb.is_chroma != 0, jb.filter_boundary != 0, dst, lane, pix_max);
    __syncwarp();
    if (preds)
      for (int i = lane; i < w * w; i += 32) preds[(size_t)jb.pred_offset + (size_t)mode * w * w + i] = dst[i];
    if (jb.has_orig) {
      // satd_NxN / sad_NxN shift their result by KVZ_BIT_DEPTH - 8, satd_4x4 does not (picture-generic.c:189-196, :322-340)
      const uint32_t satd = w == 4 ? warp_satd4_strip(a, b, 0, 0, 0, 0, 1, lane) : warp_satd8_area(a, b, 0, 0, w, w, lane) >> cost_shift;
      const uint32_t sad = warp_sad(a, b, w, w, lane) >> cost_shift;
      if (lane == 0) {
        costs[((size_t)blockIdx.x * 35 + mode) * 2] = satd;
        costs[((size_t)blockIdx.x * 35 + mode) * 2 + 1] = sad;
      }
    }
  }
}

}  // namespace

// offsets of the jobs are in SAMPLES (bytes at 8 bits)
int launch_intra_costs(const void *refs, const void *orig, const kvz_cuda_intra_job *jobs, int n_jobs,
                       uint32_t *costs, void *preds, cudaStream_t stream, int bitdepth)
{
  if (n_jobs <= 0) return 1;
  if (bitdepth > 8)
    intra_costs_kernel<uint16_t><<<n_jobs, kWarps * 32, 0, stream>>>((const uint16_t *)refs, (const uint16_t *)orig, jobs, costs,
                                                                     (uint16_t *)preds, (1 << bitdepth) - 1, bitdepth - 8);
  else
    intra_costs_kernel<uint8_t><<<n_jobs, kWarps * 32, 0, stream>>>((const uint8_t *)refs, (const uint8_t *)orig, jobs, costs,
                                                                    (uint8_t *)preds, 255, 0);
  count_launch();
  KVZ_CUDA_LAUNCH_CHECK();
  return 1;
}

}  // namespace kvz
