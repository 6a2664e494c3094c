// common.cuh -- shared helpers for the sm_100a geometry kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/b200grasp.h"

// number of kernels this library has launched in this process (bench.py's gpu_launches)
extern std::atomic<unsigned long long> g_b200_launches;

#define B200_LAUNCH_CHECK()                         \
  do {                                              \
    cudaError_t e__ = cudaGetLastError();           \
    if (e__ != cudaSuccess) return (int)e__;        \
    ++g_b200_launches;                              \
  } while (0)

#define B200_CUDA(call)                             \
  do {                                              \
    cudaError_t e__ = (call);                       \
    if (e__ != cudaSuccess) return (int)e__;        \
  } while (0)

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

// SM count of the current device (cached per device; B200 = 148).
int b200_sm_count();

// ordinal of the current device (-1 if the runtime cannot tell)
int b200_device();

// Per-device one-time setup of a kernel (cudaFuncSetAttribute is a per-device setting): one static instance per
// kernel instantiation; devices beyond 63 are simply configured on every launch.
struct B200DeviceOnce {
  std::atomic<unsigned long long> bits{0};
  bool need(int dev) const { return dev < 0 || dev >= 64 || !((bits.load(std::memory_order_acquire) >> dev) & 1ull); }
  void mark(int dev) {
    if (dev >= 0 && dev < 64) bits.fetch_or(1ull << dev, std::memory_order_release);
  }
};

// Squared norm / distance with the association the legacy pointnet2 kernels
// compile to (nvcc --fmad=true on `a*a + b*b + c*c`): fma(c,c, fma(a,a, b*b)).
// Intrinsics are never re-contracted by the compiler, so this is exact by construction.
__device__ __forceinline__ float sq3_p2(float a, float b, float c) {
  return __fmaf_rn(c, c, __fmaf_rn(a, a, __fmul_rn(b, b)));
}

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }

__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
