// tmem_read.cu -- TMEM read throughput (tcgen05.ld 32x32b .x16/.x32/.x64, waits batched or not, 4 / 8 warps) and the cost
// of a 32-row max per column (redux.sync.max.f32 vs a shuffle transpose-reduce) on sm_100a.
#include <cstdio>
#include <cstdint>
#include "../../graspnet-baseline_b200/csrc/tc_common.cuh"

template <int NREG>
__device__ __forceinline__ void ld_n(uint32_t taddr, float *v);
template <> __device__ __forceinline__ void ld_n<16>(uint32_t taddr, float *v) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),
        "=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]) : "r"(taddr));
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
template <> __device__ __forceinline__ void ld_n<32>(uint32_t taddr, float *v) {
  uint32_t r[32];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),
        "=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]),"=r"(r[16]),"=r"(r[17]),"=r"(r[18]),"=r"(r[19]),
        "=r"(r[20]),"=r"(r[21]),"=r"(r[22]),"=r"(r[23]),"=r"(r[24]),"=r"(r[25]),"=r"(r[26]),"=r"(r[27]),"=r"(r[28]),"=r"(r[29]),
        "=r"(r[30]),"=r"(r[31]) : "r"(taddr));
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float wmax(float v) {
  float r; asm volatile("redux.sync.max.NaN.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v)); return r;
}

// MODE 0: x16 + wait each; 1: x32 + wait each; 2: 4 x (x16) then one wait (64 cols in flight); 3: x16 + wait + 16 redux;
// 4: x16 + wait + shuffle transpose-reduce (16 -> 1 value per lane pair ...); 5: x32 + wait + 32 redux
template <int MODE>
__global__ void __launch_bounds__(256, 1) read_kernel(int iters, long long *cycles, float *sink) {
  __shared__ uint32_t holder;
  if (threadIdx.x < 32) tc::tmem_alloc_512(tc::smem_u32(&holder));
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = holder;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t base = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 128;   // warps 4-7: upper 128 of 256 cols
  float acc = 0.f;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0 || MODE == 3 || MODE == 4) {
      for (int s = 0; s < 8; ++s) {
        float v[16];
        ld_n<16>(base + s * 16, v); ld_wait();
        if (MODE == 0) { for (int i = 0; i < 16; ++i) acc += v[i]; }
        if (MODE == 3) { for (int i = 0; i < 16; ++i) acc += wmax(v[i]); }
        if (MODE == 4) {
          // 16 values x 32 lanes -> max over lanes: halve the value count each stage
          float a[8];
          for (int i = 0; i < 8; ++i) {
            const float mine = (lane & 16) ? v[i + 8] : v[i], other = (lane & 16) ? v[i] : v[i + 8];
            a[i] = fmaxf(mine, __shfl_xor_sync(0xffffffffu, other, 16));
          }
          float b[4];
          for (int i = 0; i < 4; ++i) {
            const float mine = (lane & 8) ? a[i + 4] : a[i], other = (lane & 8) ? a[i] : a[i + 4];
            b[i] = fmaxf(mine, __shfl_xor_sync(0xffffffffu, other, 8));
          }
          float c[2];
          for (int i = 0; i < 2; ++i) {
            const float mine = (lane & 4) ? b[i + 2] : b[i], other = (lane & 4) ? b[i] : b[i + 2];
            c[i] = fmaxf(mine, __shfl_xor_sync(0xffffffffu, other, 4));
          }
          const float mine = (lane & 2) ? c[1] : c[0], other = (lane & 2) ? c[0] : c[1];
          float d = fmaxf(mine, __shfl_xor_sync(0xffffffffu, other, 2));
          d = fmaxf(d, __shfl_xor_sync(0xffffffffu, d, 1));
          acc += d;
        }
      }
    } else if (MODE == 1 || MODE == 5) {
      for (int s = 0; s < 4; ++s) {
        float v[32];
        ld_n<32>(base + s * 32, v); ld_wait();
        if (MODE == 1) { for (int i = 0; i < 32; ++i) acc += v[i]; }
        else { for (int i = 0; i < 32; ++i) acc += wmax(v[i]); }
      }
    } else {
      for (int s = 0; s < 2; ++s) {
        float v[64];
        for (int q = 0; q < 4; ++q) ld_n<16>(base + s * 64 + q * 16, v + 16 * q);
        ld_wait();
        for (int i = 0; i < 64; ++i) acc += v[i];
      }
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * 256 + threadIdx.x] = acc;
  tc::fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tc::tmem_dealloc_512(tmem);
}

template <int MODE>
void run(const char *name) {
  long long *cyc; float *sink;
  cudaMalloc(&cyc, 8 * 148); cudaMalloc(&sink, 4 * 256 * 148);
  const int iters = 256;
  read_kernel<MODE><<<148, 256>>>(iters, cyc, sink);
  read_kernel<MODE><<<148, 256>>>(iters, cyc, sink);
  cudaError_t e = cudaDeviceSynchronize();
  long long h = 0; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  // per iteration the CTA reads 128 lanes x 256 columns x 4 B = 128 KB (8 warps x 32 lanes x 128 cols)
  printf("%-44s %s  %.0f cycles per 128x256 fp32 tile read (%.1f B/cycle/SM)\n", name, cudaGetErrorString(e),
         (double)h / iters, 131072.0 * iters / (double)h);
  cudaFree(cyc); cudaFree(sink);
}

int main() {
  run<0>("x16 + wait, sum");
  run<1>("x32 + wait, sum");
  run<2>("4 x x16, one wait, sum");
  run<3>("x16 + wait + 16 redux.max");
  run<5>("x32 + wait + 32 redux.max");
  run<4>("x16 + wait + shuffle transpose-reduce max");
  return 0;
}
