// mma_rate.cu -- issue rate of tcgen05.mma (cta_group::1, M = 128) for kind::tf32 (K = 8) and kind::f16 (fp16, K = 16)
// at N = 128 / 256, operands in shared memory (K-major, no swizzle); also checks fp16 subnormal products.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu && ./mma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include "../../graspnet-baseline_b200/csrc/tc_common.cuh"

__host__ __device__ constexpr uint32_t idesc_f16_m128(int n) {      // D fp32, A/B fp16 (format 0), K-major
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
}
__device__ __forceinline__ void mma_f16(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

template <int KIND, int N>
__global__ void __launch_bounds__(128, 1) rate_kernel(int iters, long long *cycles, float *probe) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t holder;
  const uint32_t sm = tc::smem_u32(smem);
  // A: 128 rows x 32 bytes of K (2 chunks of 16 B): chunk-major -> 4 KB; B: N rows x 32 bytes -> N * 32 B
  for (int i = threadIdx.x; i < (128 * 32 + N * 32) / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
  if (KIND == 1 && threadIdx.x == 0) {
    // row 0 of A: k0 = 2^-20 (fp16 subnormal), k1 = 1.0; row 0 of B: k0 = 2^-4, k1 = 2^-24 (smallest subnormal)
    __half *a = reinterpret_cast<__half *>(smem), *b = reinterpret_cast<__half *>(smem + 128 * 32);
    a[0] = __float2half(9.5367431640625e-07f); a[1] = __float2half(1.0f);
    b[0] = __float2half(0.0625f); b[1] = __float2half(5.9604644775390625e-08f);
  }
  if (threadIdx.x == 0) { tc::mbar_init(tc::smem_u32(&bar), 1); tc::fence_barrier_init(); }
  if (threadIdx.x < 32) tc::tmem_alloc_512(tc::smem_u32(&holder));
  tc::fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = holder;
  long long t0 = 0, t1 = 0;
  if (threadIdx.x == 0) {
    const uint64_t ad = tc::smem_desc(sm, 128 * 16, 128), bd = tc::smem_desc(sm + 128 * 32, N * 16, 128);
    const uint32_t id = KIND == 0 ? tc::idesc_tf32_m128(N) : idesc_f16_m128(N);
    t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      if (KIND == 0) tc::mma_tf32(tmem, ad, bd, id, i > 0);
      else mma_f16(tmem, ad, bd, id, i > 0);
    }
    tc::mma_commit(tc::smem_u32(&bar));
  }
  tc::mbar_wait(tc::smem_u32(&bar), 0);
  tc::fence_after_sync();
  if (threadIdx.x == 0) { t1 = clock64(); cycles[blockIdx.x] = t1 - t0; }
  if (threadIdx.x < 32) {
    float v[16];
    tc::tmem_ld16(tmem, v);
    if (threadIdx.x == 0 && blockIdx.x == 0) probe[0] = v[0];
  }
  tc::fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) tc::tmem_dealloc_512(tmem);
}

template <int KIND, int N>
void run(const char *name, int grid) {
  long long *cyc; float *probe;
  cudaMalloc(&cyc, sizeof(long long) * grid); cudaMalloc(&probe, 4);
  const int smem = 128 * 32 + N * 32, iters = 4096;
  cudaFuncSetAttribute(rate_kernel<KIND, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int rep = 0; rep < 2; ++rep) rate_kernel<KIND, N><<<grid, 128, smem>>>(iters, cyc, probe);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  rate_kernel<KIND, N><<<grid, 128, smem>>>(iters, cyc, probe);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
  long long h[148]; float p = 0;
  cudaMemcpy(h, cyc, sizeof(long long) * (grid < 148 ? grid : 148), cudaMemcpyDeviceToHost);
  cudaMemcpy(&p, probe, 4, cudaMemcpyDeviceToHost);
  const double k = KIND == 0 ? 8 : 16, flop = 2.0 * 128 * N * k * iters * grid;
  printf("%-22s grid %3d: %s  %.1f cycles/MMA (cta 0), %.3f ms, %.1f TFLOP/s, probe %.9g\n", name, grid,
         cudaGetErrorString(err), (double)h[0] / iters, ms, flop / ms * 1e-9, p / iters);
  cudaFree(cyc); cudaFree(probe);
}

int main() {
  for (int grid : {1, 148}) {
    run<0, 256>("tf32 M128 N256 K8", grid);
    run<0, 128>("tf32 M128 N128 K8", grid);
    run<1, 256>("f16  M128 N256 K16", grid);
    run<1, 128>("f16  M128 N128 K16", grid);
  }
  // probe (f16): sum over iters of (2^-20 * 2^-4 + 1.0 * 2^-24) = 2^-24 + 2^-24 = 1.19209e-07 per MMA if fp16
  // subnormal inputs are honoured
  return 0;
}
