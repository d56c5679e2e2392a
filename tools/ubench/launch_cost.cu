// tools/ubench/launch_cost.cu -- what one cudaLaunchKernel costs the calling thread when T host threads launch small
// kernels on streams of their own: parameter block of 256 B vs 6 KB, 1 vs 8 vs 16 threads.
// build: nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o launch_cost launch_cost.cu -lpthread
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>
#include <chrono>
#include <cuda_runtime.h>

struct Small { int v[64]; };
struct Big { int v[1536]; };
template <typename P> __global__ void k(const __grid_constant__ P p, int *out) { if (threadIdx.x == 0 && p.v[0] == 12345) *out = p.v[1]; }

template <typename P> double run(int threads, int iters, int smem)
{
  std::vector<std::thread> th;
  std::vector<double> us(threads);
  int *d; cudaMalloc(&d, 4);
  if (smem) cudaFuncSetAttribute(k<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int t = 0; t < threads; t++)
    th.emplace_back([&, t] {
      cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
      P p = {};
      for (int i = 0; i < 200; i++) k<P><<<1, 512, smem, s>>>(p, d);
      cudaStreamSynchronize(s);
      auto t0 = std::chrono::steady_clock::now();
      for (int i = 0; i < iters; i++) {
        k<P><<<1, 512, smem, s>>>(p, d);
        if ((i & 31) == 31) cudaStreamSynchronize(s);          // keep the queue shallow, as a host that waits for results does
      }
      cudaStreamSynchronize(s);
      us[t] = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / iters;
    });
  for (auto &x : th) x.join();
  double m = 0; for (double v : us) m += v;
  cudaFree(d);
  return m / threads;
}

int main()
{
  for (int smem : {0, 110 * 1024})
    for (int threads : {1, 8, 16})
      printf("smem %6d B, %2d threads: 256 B params %.2f us per launch (+ its share of the syncs), 6 KB params %.2f us\n", smem, threads,
             run<Small>(threads, 20000, smem), run<Big>(threads, 20000, smem));
  return 0;
}
