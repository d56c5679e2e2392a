// kvazaar_b200/csrc/common.cuh -- shared device/host helpers for libkvz_cuda.so
//
// B200 (sm_100a) only.  No CPU fallback anywhere: every entry point returns 0
// and records an error string if CUDA is unavailable or a launch fails.
#pragma once
#include <mutex>

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

namespace kvz {

// ---- error reporting (C-ABI returns 1 ok / 0 fail like the reference's
// strategy registration, strategyselector.c:216-256) ----
extern thread_local char g_last_error[512];

inline int fail(const char *what, cudaError_t e, const char *file, int line)
{
  snprintf(g_last_error, sizeof(g_last_error), "%s: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
  return 0;
}

// Raises a kernel's dynamic shared-memory limit to `bytes` once per size, from any host thread: the sizes two threads
// ask for may differ (frames with boxes of different heights are in flight together), so the comparison and the
// attribute call are one critical section and the recorded size only grows.
template <typename K>
inline cudaError_t opt_in_dynamic_smem(K kernel, size_t bytes, size_t &recorded, std::mutex &lock)
{
  if (bytes <= __atomic_load_n(&recorded, __ATOMIC_ACQUIRE)) return cudaSuccess;
  std::lock_guard<std::mutex> guard(lock);
  if (bytes <= recorded) return cudaSuccess;
  const cudaError_t st = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (st == cudaSuccess) __atomic_store_n(&recorded, bytes, __ATOMIC_RELEASE);
  return st;
}

inline int fail_msg(const char *what)
{
  snprintf(g_last_error, sizeof(g_last_error), "%s", what);
  return 0;
}

#define KVZ_CUDA_TRY(expr)                                                  \
  do {                                                                      \
    cudaError_t e__ = (expr);                                               \
    if (e__ != cudaSuccess) return ::kvz::fail(#expr, e__, __FILE__, __LINE__); \
  } while (0)

#define KVZ_CUDA_LAUNCH_CHECK()                                             \
  do {                                                                      \
    cudaError_t e__ = cudaGetLastError();                                   \
    if (e__ != cudaSuccess) return ::kvz::fail("kernel launch", e__, __FILE__, __LINE__); \
  } while (0)

constexpr int kLcuWidth = 64;   // global.h:137: CTU is fixed at 64x64

extern int g_device;        // capi.cu: device selected by kvz_cuda_init(), -1 before

// launch counter: bench.py reports gpu_launches from this
extern unsigned long long g_launch_count;
inline void count_launch(int n = 1) { __atomic_fetch_add(&g_launch_count, (unsigned long long)n, __ATOMIC_RELAXED); }   // any host thread

// pixel traits: 8-bit build of the reference uses uint8_t, 10-bit uint16_t
// (kvazaar.h:72-77)
template <typename PIX> struct PixTraits;
template <> struct PixTraits<uint8_t>  { static constexpr int kBitDepth = 8; };
template <> struct PixTraits<uint16_t> { static constexpr int kBitDepth = 10; };

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

}  // namespace kvz
