// fps_multi.cu -- furthest point sampling, several samples per cluster exchange (sm_100a).
//
// Same result as fps.cu's chain kernel (and the reference's furthest_point_sampling_kernel,
// pointnet2/_ext_src/src/sampling_gpu.cu:74-178); same data layout: a cluster of C CTAs per cloud, every
// thread keeps its points and their running min-distance in registers.  The chain kernel pays one cluster
// exchange (~600 cycles of reduce + DSMEM round trip) for every sample.  This kernel gets several samples
// out of one exchange:
//   * every warp publishes its best point {value, tie priority, x, y, z} as before, PLUS the value of its
//     second-best point.  B = the largest of those second-best values bounds every point that was NOT
//     published, now and later (min-distances only decrease).
//   * every warp then runs the same short chain over the published candidates (<= 128), which it holds one,
//     two or four per lane: take the best candidate w (max value, then min tie priority); the first one of a round is the
//     global maximum by construction; a later one is accepted only if its value is STRICTLY above B -- then it
//     beats every unpublished point whatever their tie priorities, and it beats the other candidates by
//     construction, so it is exactly the point the one-sample-per-step chain would pick next.  Accepting w
//     lowers the other candidates' values by their distance to w (the same fp32 expression the owning thread
//     evaluates), and the chain repeats until a candidate fails the test or ACC_MAX samples are accepted.
//   * every accepted sample is applied to the points (registers) WHILE the chain looks for the next one: the
//     chain is a string of dependent warp reductions with idle issue slots, the update is pure fp32 throughput.
//     A bookkeeping pass over the min-distances (best / runner-up per thread) opens the next round.
// FPS spreads its samples, so the runners-up are usually far from the winner and stay eligible: on a 20k-point
// table-top scene a round yields 5.4 samples on average (2047 samples in ~380 exchanges).
#include "common.cuh"
#include "fps_common.cuh"

namespace {

constexpr int ACC_MAX = 32;          // samples accepted per round at most (their priorities sit one per lane)
constexpr int MENT = 128;            // published entries per round: C x warps-per-CTA <= 32 * EL (EL per lane)
constexpr int MENT_BYTES = 24;       // key, prio, x, y (16 B) + z, bound (8 B)

__device__ __forceinline__ void st_async_v2(uint32_t addr, float a, int b, uint32_t bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];"
               ::"r"(addr), "r"(__float_as_uint(a)), "r"(b), "r"(bar) : "memory");
}

#ifdef FPS_PROFILE
// cycles of (cloud 0, rank 0, thread 0): compute, warp-reduce, publish, wait, chain, tail; [6] = rounds
__device__ unsigned long long g_fpsm_prof[8];
#define FPSM_TICK(i, dep)                                                                  \
  do {                                                                                     \
    long long t__ = tprev;                                                                 \
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %1, 0x7fc0dead;\n\t@q mov.u64 %0, %%clock64;\n\t}" \
                 : "+l"(t__) : "r"(dep) : "memory");                                       \
    if (prof_on) prof[i] += (unsigned long long)(t__ - tprev);                             \
    tprev = t__;                                                                           \
  } while (0)
#else
#define FPSM_TICK(i, dep) do { } while (0)
#endif

// PP point pairs per thread, NT threads per CTA, cluster of C CTAs with C * NT/32 <= 32 * EL.  Point layout and tie
// priorities exactly as in fps.cu (thread `tid` owns the residues tid + NT*b (mod 512) of every 512-block of
// its CTA's chunk, register slot order == tie-priority order); the per-CTA table holds x, y, z, prio.
template <int PP, int NT, int EL>
__global__ void __launch_bounds__(NT, 1)
fps_multi_kernel(const float *__restrict__ xyz, int n, int m, int C, int bits, FpsVerdicts spec, int32_t *idx) {
  constexpr int PPT = 2 * PP;
  constexpr int NW = NT / 32;
  constexpr int R = 512 / NT;
  constexpr int PPR = PPT / R;
  static_assert(PPT % R == 0 && PPR >= 1, "PPT must cover whole 512-blocks");
#define OFF(i) (((i) % PPR) * 512 + NT * brev_small((i) / PPR, R))
#define PRI(i) (((unsigned)((i) / PPR) << 22) + (unsigned)((i) % PPR))
  extern __shared__ float4 tab[];                  // x, y, z, prio of every point of the chunk
  __shared__ __align__(16) int4 ent[2][MENT];      // key, prio, x, y
  __shared__ __align__(8) int2 entz[2][MENT];      // z, bound
  __shared__ __align__(8) unsigned long long full[2];

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int rank = (int)cluster_ctarank();
  const int cloud = (int)(blockIdx.x / C);
  const float *p = xyz + (size_t)cloud * n * 3;
  int32_t *out = idx + (size_t)cloud * m;
  const int E = C * NW;
  const int base = rank * (PPT * NT);
  if (fps_identity(spec, cloud)) {                 // the prefix check found the identity (see fps.cu)
    if (rank == 0)
      for (int j = tid; j < m; j += NT) out[j] = j;
    return;
  }
#ifdef FPS_PROFILE
  unsigned long long prof[7] = {0, 0, 0, 0, 0, 0, 0};
  const bool prof_on = (blockIdx.x == 0 && tid == 0);
  long long tprev = clock64();
#endif
  const unsigned prio0 = prio_of(base + tid, bits);

  f32x2 X[PP], Y[PP], Z[PP];
  float md[PPT];
#pragma unroll
  for (int j = 0; j < PP; ++j) {
    float c[2][3];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int i = 2 * j + h;
      const int k = base + OFF(i) + tid;
      if (k < n) {
        c[h][0] = __ldg(p + 3 * (size_t)k);
        c[h][1] = __ldg(p + 3 * (size_t)k + 1);
        c[h][2] = __ldg(p + 3 * (size_t)k + 2);
        const float mag = sq3_p2(c[h][0], c[h][1], c[h][2]);
        md[i] = ((double)mag <= 1e-3) ? -1.0f : 1e10f;   // -1: never updated, never selected
      } else {
        c[h][0] = c[h][1] = c[h][2] = 0.f;
        md[i] = -1.0f;
      }
      tab[OFF(i) + tid] = make_float4(c[h][0], c[h][1], c[h][2], __uint_as_float(prio0 + PRI(i)));
    }
    X[j] = pack2(c[0][0], c[1][0]);
    Y[j] = pack2(c[0][1], c[1][1]);
    Z[j] = pack2(c[0][2], c[1][2]);
  }
  if (rank == 0 && tid == 0) out[0] = 0;

  // lane l < C publishes this warp's entry into CTA l; buffer 1 is a fixed byte offset after buffer 0
  const uint32_t bar_local = smem_u32(&full[0]);
  if (tid == 0) {
    mbar_init(bar_local, 1);
    mbar_init(bar_local + 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_arrive_expect_tx(bar_local, (uint32_t)E * MENT_BYTES);
    mbar_arrive_expect_tx(bar_local + 8, (uint32_t)E * MENT_BYTES);
  }
  const uint32_t peer = (uint32_t)(lane < C ? lane : 0);
  const int e_mine = rank * NW + warp;
  const uint32_t ent_remote = map_to_rank(smem_u32(&ent[0][e_mine]), peer);
  const uint32_t z_remote = map_to_rank(smem_u32(&entz[0][e_mine]), peer);
  const uint32_t bar_remote = map_to_rank(bar_local, peer);
  cluster_sync_all();                              // barriers initialised before any peer sends

  // apply one sample to pairs [J0, J1) of this thread's points: md = min(md, |p - s|^2), legacy association
#define APPLY(J0, J1, SX, SY, SZ)                                                          \
  do {                                                                                     \
    const f32x2 OX = pack2(SX, SX), OY = pack2(SY, SY), OZ = pack2(SZ, SZ);                \
    _Pragma("unroll") for (int j = (J0); j < (J1); ++j) {                                  \
      const f32x2 dx = sub2(X[j], OX), dy = sub2(Y[j], OY), dz = sub2(Z[j], OZ);           \
      const f32x2 d = fma2(dz, dz, fma2(dx, dx, mul2(dy, dy)));                            \
      float d0, d1;                                                                        \
      unpack2(d, d0, d1);                                                                  \
      md[2 * j] = fminf(d0, md[2 * j]);         /* md = -1 stays -1 (d >= 0 or NaN) */     \
      md[2 * j + 1] = fminf(d1, md[2 * j + 1]);                                            \
    }                                                                                      \
  } while (0)
  constexpr int J1 = PP / 3, J2 = (2 * PP) / 3;
  // `pend` = the last accepted sample.  It is applied to the points while the chain looks for the next one (the
  // chain is a sequence of dependent warp reductions: plenty of idle issue slots).  Applying a sample twice
  // changes nothing, so the first chain iteration of a round simply re-applies the previous round's last sample.
  float pendx = __ldg(p), pendy = __ldg(p + 1), pendz = __ldg(p + 2);
  APPLY(0, PP, pendx, pendy, pendz);               // point 0 is the first sample
  int s = 0;                                       // samples written so far = s + 1
  for (int r = 0; s + 1 < m; ++r) {
    const int par = r & 1;
    FPSM_TICK(5, r);
    // ---- registers: thread-local first strict maximum and runner-up value of the current min-distances
    float best = -1.0f, second = -1.0f;
    unsigned bc = 0;
#pragma unroll
    for (int i = 0; i < PPT; ++i) {
      const float v = md[i];
      second = fmaxf(second, fminf(best, v));
      if (v > best) { best = v; bc = (unsigned)OFF(i); }
    }
    FPSM_TICK(0, __float_as_uint(best));
    // ---- warp winner (max value, then min tie priority) and the warp's runner-up value
    const float4 t = tab[bc + tid];
    const int lkey0 = __float_as_int(best);        // >= 0 for real candidates, < 0 for none
    const int wkey = __reduce_max_sync(0xffffffffu, lkey0);
    const bool cand = (lkey0 == wkey) && (wkey >= 0);
    const unsigned lprio = cand ? __float_as_uint(t.w) : 0xffffffffu;
    const unsigned wprio = __reduce_min_sync(0xffffffffu, lprio);
    const bool own = cand && (lprio == wprio);
    const float sx = __uint_as_float(__reduce_max_sync(0xffffffffu, own ? __float_as_uint(t.x) : 0u));
    const float sy = __uint_as_float(__reduce_max_sync(0xffffffffu, own ? __float_as_uint(t.y) : 0u));
    const float sz = __uint_as_float(__reduce_max_sync(0xffffffffu, own ? __float_as_uint(t.z) : 0u));
    const int wbound = __reduce_max_sync(0xffffffffu, own ? __float_as_int(second) : lkey0);
    FPSM_TICK(1, __float_as_uint(sz));
    if (lane < C) {
      st_async_v4(ent_remote + par * (MENT * 16), wkey, wprio, sx, sy, bar_remote + par * 8);
      st_async_v2(z_remote + par * (MENT * 8), sz, wbound, bar_remote + par * 8);
    }
    FPSM_TICK(2, r);
    mbar_wait(bar_local + par * 8, (uint32_t)(r >> 1) & 1u);
    // re-arm this buffer for round r+2; peers cannot complete it before every warp of this CTA has
    // published round r+1, which happens after this point in program order
    if (tid == 0) mbar_arrive_expect_tx(bar_local + par * 8, (uint32_t)E * MENT_BYTES);
    FPSM_TICK(3, (unsigned)ent[par][0].x);
    // ---- chain over the published candidates, EL per lane (identical in every warp of the cluster)
    int ckey[EL], cbound = (int)0x80000000;
    unsigned cprio[EL];
    float cx[EL], cy[EL], cz[EL];
#pragma unroll
    for (int u = 0; u < EL; ++u) {
      const int e = u * 32 + lane;
      ckey[u] = (int)0x80000000; cprio[u] = 0xffffffffu; cx[u] = cy[u] = cz[u] = 0.f;
      if (e < E) {
        const int4 v = ent[par][e];
        const int2 w = entz[par][e];
        ckey[u] = v.x; cprio[u] = (unsigned)v.y; cx[u] = __int_as_float(v.z); cy[u] = __int_as_float(v.w);
        cz[u] = __int_as_float(w.x); cbound = max(cbound, w.y);
      }
    }
    const int B = __reduce_max_sync(0xffffffffu, cbound);        // non-negative floats order like their bits
    const int room = min(ACC_MAX, m - 1 - s);
    int aprio = -1, nacc = 0;
    for (;;) {
      int lkey = ckey[0];
#pragma unroll
      for (int u = 1; u < EL; ++u) lkey = max(lkey, ckey[u]);
      const int gkey = __reduce_max_sync(0xffffffffu, lkey);
      APPLY(0, J1, pendx, pendy, pendz);
      // this lane's best holder of the maximum (lowest tie priority among its EL candidates)
      unsigned lp = 0xffffffffu;
      float lx = 0.f, ly = 0.f, lz = 0.f;
#pragma unroll
      for (int u = 0; u < EL; ++u)
        if (ckey[u] == gkey && cprio[u] < lp) { lp = cprio[u]; lx = cx[u]; ly = cy[u]; lz = cz[u]; }
      // usually one lane holds the maximum: its coordinates come out of the same round of reductions as the
      // priority; only a tie on the value needs the priority first
      const bool hold = lp != 0xffffffffu;
      unsigned gp = __reduce_min_sync(0xffffffffu, lp);
      float wx = __uint_as_float(__reduce_max_sync(0xffffffffu, hold ? __float_as_uint(lx) : 0u));
      float wy = __uint_as_float(__reduce_max_sync(0xffffffffu, hold ? __float_as_uint(ly) : 0u));
      float wz = __uint_as_float(__reduce_max_sync(0xffffffffu, hold ? __float_as_uint(lz) : 0u));
      const unsigned holders = __ballot_sync(0xffffffffu, hold);
      APPLY(J1, J2, pendx, pendy, pendz);
      if (holders & (holders - 1)) {                             // tie: the lane with the lowest priority (unique)
        const bool win = hold && (lp == gp);
        wx = __uint_as_float(__reduce_max_sync(0xffffffffu, win ? __float_as_uint(lx) : 0u));
        wy = __uint_as_float(__reduce_max_sync(0xffffffffu, win ? __float_as_uint(ly) : 0u));
        wz = __uint_as_float(__reduce_max_sync(0xffffffffu, win ? __float_as_uint(lz) : 0u));
      }
      APPLY(J2, PP, pendx, pendy, pendz);
      // the first candidate of a round is the global maximum; a later one must beat every unpublished point
      if (nacc >= room || gkey < 0 || (nacc > 0 && gkey <= B)) break;
      if (lane == nacc) aprio = (int)gp;
      ++nacc;
      pendx = wx; pendy = wy; pendz = wz;
      // the other candidates' min-distances after this sample, as their owners will compute them
#pragma unroll
      for (int u = 0; u < EL; ++u) {
        const float d = sq3_p2(__fsub_rn(cx[u], wx), __fsub_rn(cy[u], wy), __fsub_rn(cz[u], wz));
        if (ckey[u] >= 0) ckey[u] = __float_as_int(fminf(d, __int_as_float(ckey[u])));
      }
    }
    FPSM_TICK(4, (unsigned)nacc);
    if (nacc == 0) {
      // no selectable point anywhere: the legacy kernel yields index 0 for all remaining samples
      if (rank == 0 && warp == 0)
        for (int j = s + 1 + lane; j < m; j += 32) out[j] = -1;
      break;
    }
    // the raw tie priorities are stored now, turned into point indices after the chain
    if (rank == 0 && warp == 0 && lane < nacc) out[s + 1 + lane] = aprio;
    s += nacc;
#ifdef FPS_PROFILE
    if (prof_on) prof[6] += 1;
#endif
  }
#undef APPLY
  if (rank == 0) {
    __syncthreads();
    for (int j = 1 + tid; j < m; j += NT) {
      const int v = out[j];
      out[j] = v == -1 ? 0 : index_of((unsigned)v, bits);
    }
  }
  cluster_sync_all();                              // no CTA leaves while peers may still address it
#ifdef FPS_PROFILE
  if (prof_on && cloud == 0 && rank == 0)
    for (int i = 0; i < 7; ++i) g_fpsm_prof[i] = prof[i];
#endif
#undef OFF
#undef PRI
}

template <int PP, int NT, int EL>
int launch_multi(const float *xyz, int b, int n, int m, int C, int bits, FpsVerdicts spec, int32_t *idx, cudaStream_t st) {
  constexpr int tab_bytes = 2 * PP * NT * 16;
  auto kern = fps_multi_kernel<PP, NT, EL>;
  static B200DeviceOnce configured;                // per instantiation, per device
  const int dev = b200_device();
  if (configured.need(dev)) {
    if (tab_bytes > 32 * 1024)
      B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tab_bytes));
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    configured.mark(dev);
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(b * C), 1, 1);
  cfg.blockDim = dim3(NT, 1, 1);
  cfg.dynamicSmemBytes = tab_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA(cudaLaunchKernelEx(&cfg, kern, xyz, n, m, C, bits, spec, idx));
  ++g_b200_launches;
  return 0;
}

}  // namespace

#ifdef FPS_PROFILE
extern "C" int b200_fps_multi_read_profile(unsigned long long *out7) {
  return (int)cudaMemcpyFromSymbol(out7, g_fpsm_prof, 7 * sizeof(unsigned long long));
}
#endif

// B200_ECAPACITY = this configuration is not covered here (the caller then runs the chain kernel of fps.cu)
int b200_fps_multi_launch(const float *xyz, int b, int n, int m, int C, int NT, int ppt, int bits, FpsVerdicts spec,
                          int32_t *idx, cudaStream_t st) {
  if (C < 2 || C > 16 || bits != 9) return B200_ECAPACITY;
  const int E = C * (NT / 32);
  // instantiated: 128 threads with any supported share (1 or 2 candidates per lane), 256 threads for the large
  // shares only (2 or 4 per lane)
#define FPSM_128(P)                                                                              \
  if (NT == 128 && ppt == P)                                                                     \
    return E <= 32 ? launch_multi<P / 2, 128, 1>(xyz, b, n, m, C, bits, spec, idx, st)           \
                   : launch_multi<P / 2, 128, 2>(xyz, b, n, m, C, bits, spec, idx, st);
#define FPSM_256(P)                                                                              \
  if (NT == 256 && ppt == P && E > 32)                                                           \
    return E <= 64 ? launch_multi<P / 2, 256, 2>(xyz, b, n, m, C, bits, spec, idx, st)           \
                   : launch_multi<P / 2, 256, 4>(xyz, b, n, m, C, bits, spec, idx, st);
  FPSM_128(4) FPSM_128(8) FPSM_128(12) FPSM_128(16) FPSM_128(20) FPSM_128(24) FPSM_128(32) FPSM_128(40) FPSM_128(48)
  FPSM_256(24) FPSM_256(32) FPSM_256(40) FPSM_256(48) FPSM_256(52)
#undef FPSM_128
#undef FPSM_256
  return B200_ECAPACITY;
}
