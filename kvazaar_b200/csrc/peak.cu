// kvazaar_b200/csrc/peak.cu -- integer-pipe micro-benchmark.
//
// SURVEY.md 8(d) / hard part 8: the dense SAD kernel is bound by the integer
// pipe, not by HBM, and MEASURED_PEAKS.json has no integer figure.  This
// kernel measures the denominator on the device the library runs on: a grid of
// (SM count x resident CTAs) threads issuing independent VABSDIFF4.U8.ACC
// chains from registers only -- the same SASS instruction the SAD kernels are
// made of -- so "fraction of integer-pipe peak" has a measured meaning.
#include "common.cuh"

namespace kvz {
namespace {

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c)
{
  uint32_t d;
  asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

constexpr int kChains = 16;
constexpr int kInner = 64;

__global__ void __launch_bounds__(256) vabsdiff4_peak_kernel(uint32_t *out, int iters, uint32_t seed)
{
  uint32_t acc[kChains];
  uint32_t a = seed * (threadIdx.x + 1), b = seed ^ (blockIdx.x * 0x9E3779B9u);
#pragma unroll
  for (int i = 0; i < kChains; i++) acc[i] = i;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < kInner; k++) {
#pragma unroll
      for (int i = 0; i < kChains; i++) acc[i] = sad4(a, b + i, acc[i]);
    }
    a += acc[0];
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kChains; i++) s += acc[i];
  if (s == 0xDEADBEEFu) out[0] = s;   // never true in practice; keeps the chains alive
}

}  // namespace

// returns pixel-absolute-differences per second (4 per lane-instruction)
int measure_vabsdiff4_peak(double *abs_diffs_per_s, cudaStream_t stream)
{
  int dev = 0, sms = 0;
  KVZ_CUDA_TRY(cudaGetDevice(&dev));
  KVZ_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  uint32_t *d_out = nullptr;
  KVZ_CUDA_TRY(cudaMalloc(&d_out, 4));
  cudaEvent_t e0, e1;
  KVZ_CUDA_TRY(cudaEventCreate(&e0));
  KVZ_CUDA_TRY(cudaEventCreate(&e1));
  const int ctas = sms * 8, iters = 200;
  double best = 0.0;
  for (int rep = 0; rep < 5; rep++) {
    KVZ_CUDA_TRY(cudaEventRecord(e0, stream));
    vabsdiff4_peak_kernel<<<ctas, 256, 0, stream>>>(d_out, iters, 12345u + rep);
    KVZ_CUDA_TRY(cudaEventRecord(e1, stream));
    KVZ_CUDA_TRY(cudaEventSynchronize(e1));
    count_launch();
    float ms = 0.f;
    KVZ_CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    const double ops = (double)ctas * 256.0 * iters * kInner * kChains * 4.0;
    const double rate = ops / (ms * 1e-3);
    if (rep > 0 && rate > best) best = rate;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d_out);
  KVZ_CUDA_LAUNCH_CHECK();
  *abs_diffs_per_s = best;
  return 1;
}

}  // namespace kvz
