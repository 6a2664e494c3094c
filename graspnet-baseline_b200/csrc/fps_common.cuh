// fps_common.cuh -- pieces shared by the furthest-point-sampling kernels (fps.cu, fps_multi.cu):
// tie-priority encoding, cluster / mbarrier / st.async plumbing, packed fp32x2 arithmetic.
#pragma once
#include "common.cuh"

// fps_multi.cu: several samples per cluster exchange.  Returns B200_ECAPACITY when (C, NT, ppt) is not one of
// its instantiations; b200_furthest_point_sampling then runs the one-sample-per-step chain kernel of fps.cu.
// Where the speculative prefix check left its verdicts (0 = "the answer is the identity permutation"): up to three
// arrays, entry [cloud * stride]; ALL of them must say 0 for the shortcut.  The plain entry point keeps its single
// verdict in idx[cloud * m]; the *_spec entry point takes verdicts computed ahead of time (also for ancestors: a
// level's input is only the prefix the check looked at if the level above returned the identity too).
struct FpsVerdicts {
  const int32_t *p[3];
  int stride[3];
};
__device__ __forceinline__ bool fps_identity(const FpsVerdicts &v, int cloud) {
  bool ok = v.p[0] != nullptr;
#pragma unroll
  for (int k = 0; k < 3; ++k)
    if (v.p[k]) ok = ok && (v.p[k][(size_t)cloud * v.stride[k]] == 0);
  return ok;
}

int b200_fps_multi_launch(const float *xyz, int b, int n, int m, int C, int NT, int ppt, int bits, FpsVerdicts spec,
                          int32_t *idx, cudaStream_t st);

namespace {

constexpr int MAXC = 16;       // largest cluster supported (non-portable size, opt-in)
constexpr int MAXE = 256;      // published entries per step: C x warps-per-CTA
constexpr int ENT_BYTES = 20;  // key, prio, x, y (16 B) + z (4 B)

__device__ __forceinline__ unsigned prio_of(int k, int bits) {
  const unsigned r = bits ? (__brev((unsigned)k) >> (32 - bits)) : 0u;  // bitrev of the low `bits` bits
  return (r << 22) | ((unsigned)k >> bits);
}
__device__ __forceinline__ int index_of(unsigned prio, int bits) {
  const unsigned r = prio >> 22;
  const unsigned lo = bits ? (__brev(r) >> (32 - bits)) : 0u;
  return (int)(((prio & 0x3fffffu) << bits) | lo);
}

// ---- cluster / mbarrier / st.async plumbing (PTX; shared::cluster addresses are 32-bit)
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void st_async_v4(uint32_t addr, int a, unsigned b, float c, float d, uint32_t bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];"
               ::"r"(addr), "r"(a), "r"(b), "r"(__float_as_uint(c)), "r"(__float_as_uint(d)), "r"(bar) : "memory");
}
__device__ __forceinline__ void st_async_b32(uint32_t addr, float v, uint32_t bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];"
               ::"r"(addr), "r"(__float_as_uint(v)), "r"(bar) : "memory");
}

// order in which a thread visits its 512/NT residues (mod 512): ascending bit-reversed
// priority, i.e. residue tid + NT * brev(r) with brev over log2(R) bits
__host__ __device__ constexpr int brev_small(int r, int R) {
  int out = 0;
  for (int b = 1, o = R >> 1; b < R; b <<= 1, o >>= 1)
    if (r & b) out |= o;
  return out;
}

// ---- packed fp32x2 arithmetic (FADD2 / FMUL2 / FFMA2 on sm_100): two points per instruction,
// each half rounded exactly like the scalar op, so the legacy association is preserved
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

}  // namespace
