// development aid: issue-rate microbenchmark of integer ops on sm_100a
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/int_ops tools/ubench/int_ops.cu && /tmp/int_ops
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int OP>
__global__ void k(uint32_t *out, uint32_t a, uint32_t b, int iters)
{
  uint32_t r[8];
#pragma unroll
  for (int i = 0; i < 8; i++) r[i] = threadIdx.x + i;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 4; u++) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        if (OP == 0) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(a), "r"(b));
        if (OP == 1) asm volatile("dp2a.lo.s32.s32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(a), "r"(b));
        if (OP == 2) asm volatile("mad.lo.s32 %0, %1, %2, %0;" : "+r"(r[i]) : "r"(a), "r"(b));
        if (OP == 3) asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(r[i]) : "r"(a), "r"(b));
        if (OP == 4) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(a), "r"(b));
        if (OP == 5) asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(a), "r"(b));
        if (OP == 6) asm volatile("min.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(a + i));
        if (OP == 7) asm volatile("add.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(a));
        if (OP == 8) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(a), "r"(b));
        if (OP == 9) asm volatile("bfe.s32 %0, %0, 16, 16;" : "+r"(r[i]));
      }
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
void run(const char *name)
{
  uint32_t *d; cudaMalloc(&d, 148 * 8 * 1024 * 4);
  const int iters = 4096;
  k<OP><<<148 * 8, 1024>>>(d, 3, 5, 16);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<OP><<<148 * 8, 1024>>>(d, 3, 5, iters);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double ops = 148.0 * 8 * 1024 * iters * 32;
  printf("%-10s %.2f T lane-ops/s  = %.1f lanes/clk/SM at 1.965 GHz\n", name, ops / ms / 1e9, ops / ms / 1e9 * 1e12 / 148 / 1.965e9 / 1e0 / 1e0 * 1e-0 / 1.0);
  cudaFree(d);
}

int main()
{
  run<0>("dp4a"); run<1>("dp2a"); run<2>("mad.lo"); run<3>("vabsdiff4"); run<4>("prmt"); run<5>("shf");
  run<6>("min"); run<7>("add"); run<8>("lop3"); run<9>("bfe");
  return 0;
}
