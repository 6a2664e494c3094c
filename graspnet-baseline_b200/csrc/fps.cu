// fps.cu -- furthest point sampling for sm_100a.
//
// Replaces furthest_point_sampling_kernel (reference pointnet2/_ext_src/src/sampling_gpu.cu:74-178).
//
// FPS is an m-step dependent chain (pick the point furthest from the chosen set, update
// every point's min-distance, repeat), so its cost is steps x per-step latency, not bytes.
// Design: one thread-block CLUSTER of C CTAs per cloud (C = 1 for small clouds).  Every
// thread keeps its points' x,y,z AND their running min-distance in registers for the whole
// chain: a step touches no global memory and no shared memory for point data (the legacy
// kernel re-reads 12 B/point from L2 and reads+writes a 4 B/point HBM temp every step).
// Per step there is exactly ONE exchange stage:
//   registers: d = |p - last|^2, md = min(md, d), thread-local first strict maximum
//   warp:      redux.sync max on the float bits (+ a min on the tie priority only when two
//              lanes hold the same maximum)
//   exchange:  every warp publishes its winner {key, prio, x, y, z} to ALL CTAs of the
//              cluster -- st.async into the peers' shared memory, completion counted by an
//              mbarrier (complete_tx), double buffered by step parity; with C = 1 it is a
//              plain shared-memory store and one __syncthreads
//   select:    every warp reduces the C x NW published entries and already holds the
//              winner's coordinates for the next step.
// No cluster-wide barrier, no second intra-CTA stage, no global round trip.
//
// Exactness: the winner is  max d2, then min bitrev(k mod BS) over log2(BS) bits, then
// min k, with BS = largest power of two <= n clamped to 512 -- the closed form of the
// legacy per-thread strided scan + stride-halving tree (oracle/geom_oracle.c:
// oracle_fps == oracle_fps_tree; tests/test_gpu_vs_legacy.py pins it to the legacy kernel
// itself).  Distances use the legacy FMA association; the skip test `(double)mag <= 1e-3`
// is done in fp64.
#include "common.cuh"
#include "fps_common.cuh"

namespace {

#ifdef FPS_PROFILE
// stage cycle counters of (cloud 0, rank 0, thread 0): compute, warp-reduce, publish, wait, select
__device__ unsigned long long g_fps_prof[8];
// the clock read is predicated on `dep`, so it cannot issue before that value exists
#define FPS_TICK(i, dep)                                                                   \
  do {                                                                                     \
    long long t__ = tprev;                                                                 \
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %1, 0x7fc0dead;\n\t@q mov.u64 %0, %%clock64;\n\t}" \
                 : "+l"(t__) : "r"(dep) : "memory");                                       \
    if (prof_on) prof[i] += (unsigned long long)(t__ - tprev);                             \
    tprev = t__;                                                                           \
  } while (0)
#else
#define FPS_TICK(i, dep) do { } while (0)
#endif

struct Entry { int key; unsigned prio; float x, y, z; };

__device__ __forceinline__ void better(Entry &a, const Entry &b) {   // a = the winner of (a, b)
  if (b.key > a.key || (b.key == a.key && b.prio < a.prio)) a = b;
}

constexpr int TAB_LIMIT = 200 * 1024;   // bytes of shared memory the per-CTA point table may take

// PP point PAIRS per thread (PPT = 2 PP points), NT threads per CTA.  A CTA owns `chunk` =
// PPT*NT consecutive points (a multiple of 512); thread `tid` owns the residues tid + NT*b
// (mod 512) of every 512-block of the chunk, visited so that register slot order ==
// tie-priority order.
// ET = 0: one CTA per cloud (entries exchanged through its own shared memory);
// ET > 0: cluster of C CTAs, C * NW <= 32 * ET entries, ET per lane in the select stage.
template <int PP, int NT, int ET>
__global__ void __launch_bounds__(NT, 1)
fps_kernel(const float *__restrict__ xyz, int n, int m, int C, int bits, FpsVerdicts spec, int32_t *idx) {
  constexpr bool CLUSTER = ET > 0;
  constexpr int PPT = 2 * PP;
  constexpr int NW = NT / 32;
  constexpr int R = 512 / NT;
  constexpr int PPR = PPT / R;
  constexpr bool TAB = PPT * NT * 16 <= TAB_LIMIT;
  constexpr bool SOLO = !CLUSTER && TAB;           // one CTA: winner by a shared-memory 64-bit max
  // cluster alternative: CTA winner as in SOLO, then ONE entry per CTA.  Measured on B200 it ties the flat
  // per-warp exchange (442 ns/step at C=16 vs 434 at C=8: the wait is straggler-bound, not message-bound)
  // while occupying twice the SMs, so it is compiled out.
  constexpr bool HIER = false && CLUSTER && TAB;
  constexpr bool WSLOT = SOLO || HIER;
  static_assert(PPT % R == 0 && PPR >= 1, "PPT must cover whole 512-blocks");
  // slot i of a thread: local point number OFF(i) + tid, tie-priority PRI(i) + prio0
#define OFF(i) (((i) % PPR) * 512 + NT * brev_small((i) / PPR, R))
#define PRI(i) (((unsigned)((i) / PPR) << 22) + (unsigned)((i) % PPR))
#define CODE(i) ((TAB && !WSLOT) ? (unsigned)OFF(i) : PRI(i))
  extern __shared__ float4 tab[];                  // TAB: x, y, z, prio of every point of the chunk
  __shared__ __align__(16) int4 ent[2][MAXE];      // key, prio, x, y
  __shared__ float entz[2][MAXE];
  __shared__ __align__(8) unsigned long long full[2];
  __shared__ __align__(16) int2 wslot[3][NW];      // SOLO: per warp {~prio, key} = one signed 64-bit value

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int rank = CLUSTER ? (int)cluster_ctarank() : 0;
  const int cloud = CLUSTER ? (int)(blockIdx.x / C) : (int)blockIdx.x;
  const float *p = xyz + (size_t)cloud * n * 3;
  int32_t *out = idx + (size_t)cloud * m;
  const int E = HIER ? C : C * NW;                 // entries published per step
  const int base = rank * (PPT * NT);
  // Speculative mode: the prefix-check kernels below left out[0] == 0 iff the answer is the
  // identity permutation (the input is already in furthest-point order); then nothing is left
  // to do but write it.  The test is uniform over the cluster, so every CTA leaves together.
  if (fps_identity(spec, cloud)) {
    if (rank == 0)
      for (int j = tid; j < m; j += NT) out[j] = j;
    return;
  }
#ifdef FPS_PROFILE
  unsigned long long prof[6] = {0, 0, 0, 0, 0, 0};
  const bool prof_on = (blockIdx.x == 0 && tid == 0);
  long long tprev = clock64();
#endif
  // tie priority of slot i = prio0 + PRI(i): brev9(tid + NT*brev(r)) == brev9(tid) + r for
  // r < R, and k >> 9 advances by one per 512-block.  (n < 512 runs with one valid slot.)
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
      if (TAB) tab[OFF(i) + tid] = make_float4(c[h][0], c[h][1], c[h][2], __uint_as_float(prio0 + PRI(i)));
    }
    X[j] = pack2(c[0][0], c[1][0]);
    Y[j] = pack2(c[0][1], c[1][1]);
    Z[j] = pack2(c[0][2], c[1][2]);
  }

  float ox = __ldg(p), oy = __ldg(p + 1), oz = __ldg(p + 2);
  if (rank == 0 && tid == 0) out[0] = 0;
  if (WSLOT) {
    if (tid < 3 * NW) wslot[tid / NW][tid % NW] = make_int2(0, (int)0x80000000);
    __syncthreads();                               // table and slots written
  }
  int cur = 0;                                     // SOLO: slot of this step (rotates over 3)

  // shared::cluster addresses of buffer 0 (buffer 1 is a fixed byte offset further on);
  // lane l < C publishes this warp's entry into CTA l
  uint32_t bar_local = 0, ent_remote = 0, z_remote = 0, bar_remote = 0;
  if (CLUSTER) {
    bar_local = smem_u32(&full[0]);
    if (tid == 0) {
      mbar_init(bar_local, 1);
      mbar_init(bar_local + 8, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      mbar_arrive_expect_tx(bar_local, (uint32_t)E * ENT_BYTES);
      mbar_arrive_expect_tx(bar_local + 8, (uint32_t)E * ENT_BYTES);
    }
    const uint32_t peer = (uint32_t)(lane < C ? lane : 0);
    const int e_mine = HIER ? rank : rank * NW + warp;
    ent_remote = map_to_rank(smem_u32(&ent[0][e_mine]), peer);
    z_remote = map_to_rank(smem_u32(&entz[0][e_mine]), peer);
    bar_remote = map_to_rank(bar_local, peer);
    cluster_sync_all();                            // barriers initialised before any peer sends
  }

  for (int s = 0; s + 1 < m; ++s) {
    const int par = s & 1;
    FPS_TICK(5, s);
    // ---- registers: update min-distances, thread-local first strict maximum
    const f32x2 OX = pack2(ox, ox), OY = pack2(oy, oy), OZ = pack2(oz, oz);
    float best = -1.0f;
    unsigned bc = 0;
#pragma unroll
    for (int j = 0; j < PP; ++j) {
      const f32x2 dx = sub2(X[j], OX), dy = sub2(Y[j], OY), dz = sub2(Z[j], OZ);
      const f32x2 d = fma2(dz, dz, fma2(dx, dx, mul2(dy, dy)));    // legacy association, per half
      float d0, d1;
      unpack2(d, d0, d1);
      const float m0 = fminf(d0, md[2 * j]);       // md = -1 stays -1 (d >= 0 or NaN)
      md[2 * j] = m0;
      if (m0 > best) { best = m0; bc = CODE(2 * j); }
      const float m1 = fminf(d1, md[2 * j + 1]);
      md[2 * j + 1] = m1;
      if (m1 > best) { best = m1; bc = CODE(2 * j + 1); }
    }
    FPS_TICK(0, __float_as_uint(best));
    if (WSLOT) {
      // ---- one CTA: per warp, lane 0 stores the warp's maximum key and the lanes holding it
      // post ~prio with a native 32-bit shared max (no unique owner needed); after the barrier
      // everybody takes the signed 64-bit max {key, ~prio} over the warps and looks the
      // winner's coordinates up in the point table.  The slot of step s+1 is cleared now: its
      // last readers (step s-2) are all past barrier s-1.
      const int lkey = __float_as_int(best);       // >= 0 for real candidates, < 0 for none
      const int wkey = __reduce_max_sync(0xffffffffu, lkey);
      const int nxt = cur == 2 ? 0 : cur + 1;
      if (lane == 0) {
        wslot[cur][warp].y = wkey;
        wslot[nxt][warp].x = 0;
      }
      if (lkey == wkey && wkey >= 0) atomicMax((unsigned *)&wslot[cur][warp].x, ~(prio0 + bc));
      FPS_TICK(1, (unsigned)wkey);
      __syncthreads();
      long long v;
      {
        long long e[NW];
        const longlong2 *row = reinterpret_cast<const longlong2 *>(&wslot[cur][0]);
#pragma unroll
        for (int t = 0; t < NW / 2; ++t) {
          const longlong2 q = row[t];
          e[2 * t] = q.x; e[2 * t + 1] = q.y;
        }
#pragma unroll
        for (int st = 1; st < NW; st *= 2)
#pragma unroll
          for (int t = 0; t + st < NW; t += 2 * st) e[t] = e[t] > e[t + st] ? e[t] : e[t + st];
        v = e[0];
      }
      cur = nxt;
      FPS_TICK(3, (unsigned)v);
      if (SOLO) {
        int sel = 0;
        if (v >= 0) {
          sel = index_of(~(unsigned)v, bits);
          const float4 t = tab[sel];
          ox = t.x; oy = t.y; oz = t.z;
        }  // else: every point is skipped -> legacy yields index 0 and nothing is ever updated
        FPS_TICK(4, __float_as_uint(ox));
        if (tid == 0) out[s + 1] = sel;
        continue;
      }
      // HIER: warp 0 looks the CTA winner's coordinates up and pushes ONE entry into every CTA
      if (warp == 0) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        unsigned cprio = 0xffffffffu;
        if (v >= 0) {
          cprio = ~(unsigned)v;
          t = tab[index_of(cprio, bits) - base];
        }
        if (lane < C) {
          st_async_v4(ent_remote + par * (MAXE * 16), (int)(v >> 32), cprio, t.x, t.y, bar_remote + par * 8);
          st_async_b32(z_remote + par * (MAXE * 4), t.z, bar_remote + par * 8);
        }
      }
    }
    if (!HIER) {
    // ---- this thread's candidate: coordinates and tie priority
    float cx = 0.f, cy = 0.f, cz = 0.f;
    unsigned cprio;
    if (TAB) {
      const float4 t = tab[bc + tid];
      cx = t.x; cy = t.y; cz = t.z; cprio = __float_as_uint(t.w);
    } else {
      cprio = prio0 + bc;
#pragma unroll
      for (int j = 0; j < PP; ++j) {
        float a0, a1, b0, b1, c0, c1;
        unpack2(X[j], a0, a1); unpack2(Y[j], b0, b1); unpack2(Z[j], c0, c1);
        if (bc == PRI(2 * j)) { cx = a0; cy = b0; cz = c0; }
        if (bc == PRI(2 * j + 1)) { cx = a1; cy = b1; cz = c1; }
      }
    }
    // ---- warp winner: max key, then min tie priority among the lanes holding it (unique);
    // its coordinates reach every lane as an unsigned max against zeros
    const int lkey = __float_as_int(best);         // >= 0 for real candidates, < 0 for none
    const int wkey = __reduce_max_sync(0xffffffffu, lkey);
    const bool cand = (lkey == wkey) && (wkey >= 0);
    const unsigned lprio = cand ? cprio : 0xffffffffu;
    const unsigned wprio = __reduce_min_sync(0xffffffffu, lprio);
    const bool own = cand && (lprio == wprio);
    const float sx = __uint_as_float(__reduce_max_sync(0xffffffffu, own ? __float_as_uint(cx) : 0u));
    const float sy = __uint_as_float(__reduce_max_sync(0xffffffffu, own ? __float_as_uint(cy) : 0u));
    const float sz = __uint_as_float(__reduce_max_sync(0xffffffffu, own ? __float_as_uint(cz) : 0u));
    FPS_TICK(1, __float_as_uint(sz));
    if (CLUSTER) {
      // push the entry into every CTA of the cluster (own CTA included); each store carries
      // its completion bytes to the destination CTA's mbarrier
      if (lane < C) {
        st_async_v4(ent_remote + par * (MAXE * 16), wkey, wprio, sx, sy, bar_remote + par * 8);
        st_async_b32(z_remote + par * (MAXE * 4), sz, bar_remote + par * 8);
      }
    } else {
      if (lane == 0) {
        ent[par][warp] = make_int4(wkey, (int)wprio, __float_as_int(sx), __float_as_int(sy));
        entz[par][warp] = sz;
      }
    }
    }  // !HIER
    if (CLUSTER) {
      FPS_TICK(2, s);
      mbar_wait(bar_local + par * 8, (uint32_t)(s >> 1) & 1u);
      // re-arm this buffer for step s+2; peers cannot complete it before every warp of this
      // CTA has published step s+1, which happens after this point in program order
      if (tid == 0 && s + 3 < m) mbar_arrive_expect_tx(bar_local + par * 8, (uint32_t)E * ENT_BYTES);
    } else {
      FPS_TICK(2, s);
      __syncthreads();
    }
    FPS_TICK(3, (unsigned)ent[par][0].x);
    // ---- select the global winner among the E published entries
    Entry w;
    constexpr bool SCAN = !CLUSTER && NW <= 8;     // few entries: private scan beats the redux rounds
    constexpr int ETL = CLUSTER ? ET : 1;
    if (SCAN) {
      // one entry per warp: every thread scans them all (broadcast reads), no shuffles
      Entry e[NW];
#pragma unroll
      for (int t = 0; t < NW; ++t) {
        const int4 v = ent[par][t];
        e[t].key = v.x; e[t].prio = (unsigned)v.y; e[t].x = __int_as_float(v.z); e[t].y = __int_as_float(v.w);
        e[t].z = entz[par][t];
      }
#pragma unroll
      for (int st = 1; st < NW; st *= 2)
#pragma unroll
        for (int t = 0; t + st < NW; t += 2 * st) better(e[t], e[t + st]);
      w = e[0];
    } else {
      // ETL entries per lane, then three redux rounds: key, tie priority, coordinates
      w.key = (int)0x80000000; w.prio = 0xffffffffu; w.x = w.y = w.z = 0.f;
#pragma unroll
      for (int t = 0; t < ETL; ++t) {
        const int e = t * 32 + lane;
        const int ee = e < E ? e : 0;
        const int4 v = ent[par][ee];
        Entry c;
        c.key = e < E ? v.x : (int)0x80000000; c.prio = (unsigned)v.y;
        c.x = __int_as_float(v.z); c.y = __int_as_float(v.w); c.z = entz[par][ee];
        better(w, c);
      }
      const int gkey = __reduce_max_sync(0xffffffffu, w.key);
      const unsigned cp = (w.key == gkey) ? w.prio : 0xffffffffu;
      const unsigned gprio = __reduce_min_sync(0xffffffffu, cp);
      const bool win = (cp == gprio) && (w.key == gkey) && (gkey >= 0);   // exactly one lane
      w.x = __uint_as_float(__reduce_max_sync(0xffffffffu, win ? __float_as_uint(w.x) : 0u));
      w.y = __uint_as_float(__reduce_max_sync(0xffffffffu, win ? __float_as_uint(w.y) : 0u));
      w.z = __uint_as_float(__reduce_max_sync(0xffffffffu, win ? __float_as_uint(w.z) : 0u));
      w.key = gkey; w.prio = gprio;
    }
    FPS_TICK(4, __float_as_uint(w.x));
    if (w.key >= 0) { ox = w.x; oy = w.y; oz = w.z; }
    // else: every point is skipped -> legacy yields index 0 and nothing is ever updated
    // the raw tie priority is stored now (one predicated store), turned into the point
    // index after the chain: keeps the bit-reversal off every step's critical path
    if (rank == 0 && tid == 0) out[s + 1] = w.key >= 0 ? (int)w.prio : -1;
  }
  if (!SOLO && rank == 0) {
    __syncthreads();
    for (int j = 1 + tid; j < m; j += NT) {
      const int v = out[j];
      out[j] = v == -1 ? 0 : index_of((unsigned)v, bits);
    }
  }
  if (CLUSTER) cluster_sync_all();                 // no CTA leaves while peers may still address it
#ifdef FPS_PROFILE
  if (prof_on && cloud == 0 && rank == 0)
    for (int i = 0; i < 6; ++i) g_fps_prof[i] = prof[i];
#endif
#undef OFF
#undef PRI
#undef CODE
}

template <int PP, int NT, int ET>
int launch(const float *xyz, int b, int n, int m, int C, int bits, FpsVerdicts spec, int32_t *idx, cudaStream_t st) {
  constexpr int tab_bytes = (2 * PP * NT * 16 <= TAB_LIMIT) ? 2 * PP * NT * 16 : 0;
  auto kern = fps_kernel<PP, NT, ET>;
  static B200DeviceOnce configured;                // per instantiation, per device
  const int dev = b200_device();
  if (configured.need(dev)) {
    if (tab_bytes > 32 * 1024)
      B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tab_bytes));
    if (ET > 0) B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    configured.mark(dev);
  }
  if (ET == 0) {
    kern<<<b, NT, tab_bytes, st>>>(xyz, n, m, 1, bits, spec, idx);
    B200_LAUNCH_CHECK();
    return 0;
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

template <int PP, int NT>
int launch_et(const float *xyz, int b, int n, int m, int C, int bits, FpsVerdicts spec, int32_t *idx, cudaStream_t st) {
  const int E = C * (NT / 32);
  if (C == 1) return launch<PP, NT, 0>(xyz, b, n, m, C, bits, spec, idx, st);
  if (E <= 32) return launch<PP, NT, 1>(xyz, b, n, m, C, bits, spec, idx, st);
  if (E <= 64) return launch<PP, NT, 2>(xyz, b, n, m, C, bits, spec, idx, st);
  if (E <= 128) return launch<PP, NT, 4>(xyz, b, n, m, C, bits, spec, idx, st);
  return B200_ECAPACITY;
}

// ---------------------------------------------------------------------------------------
// Speculative prefix check.  PointNet++ feeds SA2..SA4 the points SA1's FPS chose, in the
// order it chose them (reference models/backbone.py:110-119), and FPS over an FPS-ordered
// prefix returns 0,1,2,... -- unless a tie or the skip rule says otherwise.  Whether the
// identity IS the legacy answer can be decided without the serial chain: it is iff for
// every step j, point j beats every other point under (min-distance to points 0..j-1, tie
// priority).  Two flat kernels decide it (n*m + m*m/2 distance evaluations, no step-to-step
// synchronisation); the chain kernel then only runs when the check failed.  No scratch: the
// check keeps D_j (bits of point j's min-distance at step j) in idx[j] and its verdict in
// idx[0] (0 = identity holds, -1 = run the chain).
constexpr int SPEC_NT = 128;
constexpr int SPEC_SUB = 8;            // lanes cooperating on one point
constexpr int SPEC_PTS = SPEC_NT / SPEC_SUB;
constexpr int CHK_SUB = 32;            // ... in the check kernel (a whole warp per point: 4x shorter serial walks)
constexpr int CHK_PTS = SPEC_NT / CHK_SUB;
constexpr int SPEC_MAX_M = 2048;       // prefix length the check supports (32 KB of shared memory)
constexpr int SPEC_MAX_N = 4096;

// D_j = min(1e10, min_{i<j} |p_j - p_i|^2) for 1 <= j < m.  Eight lanes per j, each takes every
// eighth i; min is order-independent, so the result equals the chain's running minimum.
__global__ void __launch_bounds__(SPEC_NT)
fps_prefix_dist_kernel(const float *__restrict__ xyz, long long cstride, int n, int m, int32_t *__restrict__ idx) {
  __shared__ float4 sel[SPEC_MAX_M];
  const int cloud = blockIdx.y;
  const float *p = xyz + (size_t)cloud * cstride * 3;           // clouds are `cstride` points apart (>= n)
  int32_t *out = idx + (size_t)cloud * m;
  const int sub = threadIdx.x & (SPEC_SUB - 1);
  const int j = blockIdx.x * SPEC_PTS + (threadIdx.x >> 3);
  const int jhi = min(m, (int)(blockIdx.x + 1) * SPEC_PTS);     // this CTA needs points i < jhi - 1
  for (int i = threadIdx.x; i < jhi; i += SPEC_NT)
    sel[i] = make_float4(__ldg(p + 3 * (size_t)i), __ldg(p + 3 * (size_t)i + 1), __ldg(p + 3 * (size_t)i + 2), 0.f);
  __syncthreads();
  float md = 1e10f;
  if (j < m) {
    const float4 q = sel[j];
#pragma unroll 4
    for (int i = sub; i < j; i += SPEC_SUB) {
      const float4 c = sel[i];
      md = fminf(sq3_p2(__fsub_rn(q.x, c.x), __fsub_rn(q.y, c.y), __fsub_rn(q.z, c.z)), md);
    }
  }
#pragma unroll
  for (int o = SPEC_SUB / 2; o >= 1; o >>= 1) md = fminf(md, __shfl_xor_sync(0xffffffffu, md, o));
  if (j < m && sub == 0) out[j] = j ? __float_as_int(md) : 0;   // idx[0] = verdict, starts as "identity holds"
}

// CHK_SUB lanes per point k, each owning a contiguous range of steps j.  Pass 1: minimum of the
// distances in the lane's range; an exclusive prefix-min over the eight lanes gives the running
// min-distance at the start of each range; pass 2 walks the range again comparing with D_j.
// A violation (k would have been picked instead of j) or a skipped j clears the verdict.
__global__ void __launch_bounds__(SPEC_NT)
fps_prefix_check_kernel(const float *__restrict__ xyz, long long cstride, int n, int m, int bits, int32_t *idx) {
  __shared__ float4 sel[SPEC_MAX_M];                            // x, y, z of point j, D_j in .w
  const int cloud = blockIdx.y;
  const float *p = xyz + (size_t)cloud * cstride * 3;
  int32_t *out = idx + (size_t)cloud * m;
  bool bad = false;
  for (int j = threadIdx.x; j < m; j += SPEC_NT) {
    const float a = __ldg(p + 3 * (size_t)j), b = __ldg(p + 3 * (size_t)j + 1), c = __ldg(p + 3 * (size_t)j + 2);
    sel[j] = make_float4(a, b, c, j ? __int_as_float(out[j]) : 0.f);
    if (j && (double)sq3_p2(a, b, c) <= 1e-3) bad = true;       // a skipped point is never picked
  }
  __syncthreads();
  const int sub = threadIdx.x & (CHK_SUB - 1);
  const int k = blockIdx.x * CHK_PTS + threadIdx.x / CHK_SUB;
  // steps per lane, rounded to an odd count: the 8 lanes of a group then read 8 different 16-byte
  // bank groups (an even stride puts all of them on the same banks: 8-way conflict, measured 12x slower)
  const int len = ((m - 1 + CHK_SUB - 1) / CHK_SUB) | 1;
  const int j0 = 1 + sub * len, j1 = min(m, j0 + len);
  float x = 0.f, y = 0.f, z = 0.f;
  bool live = false;
  if (k < n) {
    x = __ldg(p + 3 * (size_t)k); y = __ldg(p + 3 * (size_t)k + 1); z = __ldg(p + 3 * (size_t)k + 2);
    live = !((double)sq3_p2(x, y, z) <= 1e-3);                  // skipped points never compete
  }
  float mn = 1e10f;
#pragma unroll 4
  for (int j = j0; j < j1; ++j) {
    const float4 c = sel[j - 1];
    mn = fminf(sq3_p2(__fsub_rn(x, c.x), __fsub_rn(y, c.y), __fsub_rn(z, c.z)), mn);
  }
  // exclusive prefix-min over the 8 lanes of the group
  float run = mn;
#pragma unroll
  for (int o = 1; o < CHK_SUB; o <<= 1) {
    const float t = __shfl_up_sync(0xffffffffu, run, o, CHK_SUB);
    if (sub >= o) run = fminf(run, t);
  }
  float md = __shfl_up_sync(0xffffffffu, run, 1, CHK_SUB);
  if (sub == 0) md = 1e10f;
  if (live && !bad) {
    const unsigned pk = prio_of(k, bits);
    // (no early exit: a violation is rare, and without the break the loop unrolls -- four independent distances in
    // flight under the dependent running minimum)
#pragma unroll 4
    for (int j = j0; j < j1; ++j) {
      const float4 c = sel[j - 1];
      md = fminf(sq3_p2(__fsub_rn(x, c.x), __fsub_rn(y, c.y), __fsub_rn(z, c.z)), md);
      const float dj = sel[j].w;
      if (md >= dj && k != j) {                                 // rare: decide with the tie priority
        if (md > dj || pk < prio_of(j, bits)) bad = true;
      }
    }
  }
  if (bad) out[0] = -1;
}

// smallest supported points-per-thread >= want (0 if none)
int round_ppt(int want, int nt) {
  static const int p512[] = {2, 4, 6, 8, 10, 12, 16, 20, 24};
  static const int p256[] = {2, 4, 6, 8, 10, 12, 16, 20, 24, 32, 40, 48, 52};
  static const int p128[] = {4, 8, 12, 16, 20, 24, 32, 40, 48};
  if (nt == 512) {
    for (int v : p512) if (v >= want) return v;
  } else if (nt == 256) {
    for (int v : p256) if (v >= want) return v;
  } else {
    for (int v : p128) if (v >= want) return v;
  }
  return 0;
}

bool spec_eligible(int b, int n, int m) {
  return m >= 2 && m <= n && m <= SPEC_MAX_M && n <= SPEC_MAX_N && b <= 65535;
}

int g_force_cluster = 0, g_force_threads = 0, g_spec_enabled = 1, g_multi_enabled = 1, g_max_cluster = 0;

// Largest cluster size the device will co-schedule for this kernel family (16 needs the
// non-portable opt-in and a GPC with 16 free SMs; fall back to the portable 8).
template <int NT>
int max_cluster() {
  static std::atomic<int> per_device[64];
  const int dev = b200_device();
  if (dev >= 0 && dev < 64 && per_device[dev].load(std::memory_order_relaxed)) return per_device[dev].load(std::memory_order_relaxed);
  auto kern = fps_kernel<(NT == 128 ? 2 : 1), NT, 2>;
  int cached = 8;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(MAXC, 1, 1);
    cfg.blockDim = dim3(NT, 1, 1);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = MAXC;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int nclusters = 0;
    if (cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg) == cudaSuccess && nclusters >= 1) cached = MAXC;
  }
  (void)cudaGetLastError();
  if (dev >= 0 && dev < 64) per_device[dev].store(cached, std::memory_order_relaxed);
  return cached;
}

}  // namespace

// Test / tuning hooks: force the cluster size and the CTA size (0 = automatic).
extern "C" void b200_fps_set_cluster(int c) { g_force_cluster = c; }
extern "C" void b200_fps_set_threads(int t) { g_force_threads = t; }
// Cap on the AUTOMATIC cluster size (0 = none).  A stream of scenes is SM-time-bound and a 16-CTA cluster holds 16 SMs'
// registers for the whole sampling chain: 8 CTAs take 395 us instead of 365 us per cloud but half the SMs (GraspNetPipeline
// captures its slots with 8: +3 % scenes/s); small clouds (one CTA) are not affected.
extern "C" void b200_fps_set_max_cluster(int c) { g_max_cluster = c; }
extern "C" void b200_fps_set_speculation(int on) { g_spec_enabled = on; }
extern "C" void b200_fps_set_multi(int on) { g_multi_enabled = on; }
#ifdef FPS_PROFILE
extern "C" int b200_fps_read_profile(unsigned long long *out5) {
  return (int)cudaMemcpyFromSymbol(out5, g_fps_prof, 6 * sizeof(unsigned long long));
}
#endif

// `ext` != nullptr: the caller ran b200_fps_prefix_check ahead of time (possibly for ancestor levels too) and the
// shortcut is decided by those verdicts; nullptr: the check is run here when the shape allows it
static int fps_impl(const float *xyz, int b, int n, int m, const FpsVerdicts *ext, int32_t *idx, void *stream) {
  if (b < 0 || m < 0 || n < 0) return B200_EINVAL;
  if (b == 0 || m == 0) return 0;
  if (n <= 0 || !xyz || !idx) return B200_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  int bs = 1, bits = 0;
  while (bs * 2 <= n && bs < 512) { bs *= 2; ++bits; }

  int C = 1, NT = 512;
  if (bits == 9) {
    // Per-step cost on B200 (tools/fps_prof.py): ~9 cycles per 128 points of a CTA's share plus
    // an exchange of ~340 cycles alone, ~520 in a cluster of 2-4, ~720 in a cluster of 8 and
    // >1000 in a cluster of 16 (stragglers).  So: one CTA up to ~3k points, 4 CTAs up to
    // ~12k, 8 beyond -- but never more CTAs than let every cloud of the batch be resident.
    const int sms = b200_sm_count();
    int want = n <= 3072 ? 1 : n <= 12288 ? 4 : 8;
    // fps_multi.cu pays the exchange once per ~6 samples, which moves the optimum to wider clusters
    // (tools/fps_multi_sweep.py: 8 CTAs from ~3k points, 16 beyond ~16k)
    if (g_multi_enabled) want = n <= 3072 ? 1 : (n <= 16384 || (long long)b * 16 > sms / 2) ? 8 : 16;
    while (want > 1 && (long long)b * want > sms) want /= 2;
    if (g_max_cluster > 0 && want > g_max_cluster) want = g_max_cluster;   // throughput setting: see b200_fps_set_max_cluster
    if (g_force_cluster > 0) want = g_force_cluster;
    NT = 128;
    if (g_force_threads == 128 || g_force_threads == 256 || g_force_threads == 512) NT = g_force_threads;
    const int cmax = NT == 128 ? max_cluster<128>() : NT == 256 ? max_cluster<256>() : max_cluster<512>();
    C = want > cmax ? cmax : want;
    if (g_force_threads == 0) {
      // automatic CTA size: the smallest that holds a CTA's share in registers; if the share does not
      // fit (very large clouds), widen the cluster until some CTA size does
      for (;; C *= 2) {
        const int share = ceil_div(ceil_div(n, C), 512) * 512;
        NT = share <= 128 * 48 ? 128 : share <= 256 * 52 ? 256 : 512;
        const bool fits = round_ppt(share / NT, NT) != 0 && C * (NT / 32) <= 128;
        if (fits || C * 2 > cmax) break;
      }
    } else {
      while (C > 1 && C * (NT / 32) > 128) C /= 2;             // at most 4 entries per lane
      while (C < cmax && C * 2 * (NT / 32) <= 128 &&
             round_ppt(ceil_div(ceil_div(n, C), 512) * 512 / NT, NT) == 0) C *= 2;
    }
  }
  const int blocks = ceil_div(ceil_div(n, C), 512);            // 512-blocks per CTA
  const int ppt = round_ppt(blocks * 512 / NT, NT);
  if (ppt == 0) return B200_ECAPACITY;
  // small clouds may be FPS-ordered already (SA2..SA4): decide that first, flat and parallel
  FpsVerdicts spec = {{nullptr, nullptr, nullptr}, {0, 0, 0}};
  if (ext) {
    spec = *ext;
  } else if (g_spec_enabled && spec_eligible(b, n, m)) {
    fps_prefix_dist_kernel<<<dim3((unsigned)ceil_div(m, SPEC_PTS), (unsigned)b), SPEC_NT, 0, st>>>(xyz, n, n, m, idx);
    B200_LAUNCH_CHECK();
    fps_prefix_check_kernel<<<dim3((unsigned)ceil_div(n, CHK_PTS), (unsigned)b), SPEC_NT, 0, st>>>(xyz, n, n, m, bits, idx);
    B200_LAUNCH_CHECK();
    spec.p[0] = idx;                                            // verdict of cloud c in idx[c * m]
    spec.stride[0] = m;
  }
  if (g_multi_enabled && C >= 2) {                              // several samples per exchange where covered
    const int rc = b200_fps_multi_launch(xyz, b, n, m, C, NT, ppt, bits, spec, idx, st);
    if (rc != B200_ECAPACITY) return rc;
  }
#define FPS_CASE(P, T) if (NT == T && ppt == P) return launch_et<P / 2, T>(xyz, b, n, m, C, bits, spec, idx, st);
  FPS_CASE(2, 512) FPS_CASE(4, 512) FPS_CASE(6, 512) FPS_CASE(8, 512) FPS_CASE(10, 512) FPS_CASE(12, 512)
  FPS_CASE(16, 512) FPS_CASE(20, 512) FPS_CASE(24, 512)
  FPS_CASE(2, 256) FPS_CASE(4, 256) FPS_CASE(6, 256) FPS_CASE(8, 256) FPS_CASE(10, 256) FPS_CASE(12, 256)
  FPS_CASE(16, 256) FPS_CASE(20, 256) FPS_CASE(24, 256) FPS_CASE(32, 256) FPS_CASE(40, 256) FPS_CASE(48, 256)
  FPS_CASE(52, 256)
  FPS_CASE(4, 128) FPS_CASE(8, 128) FPS_CASE(12, 128) FPS_CASE(16, 128) FPS_CASE(20, 128) FPS_CASE(24, 128)
  FPS_CASE(32, 128) FPS_CASE(40, 128) FPS_CASE(48, 128)
#undef FPS_CASE
  return B200_ECAPACITY;
}

extern "C" int b200_furthest_point_sampling(const float *xyz, int b, int n, int m, int32_t *idx,
                                            void *stream) {
  return fps_impl(xyz, b, n, m, nullptr, idx, stream);
}

extern "C" int b200_fps_prefix_check(const float *xyz, int b, int n, int m, long long cloud_stride,
                                     int32_t *verdict, void *stream) {
  if (b < 0 || n < 0 || m < 0 || cloud_stride < n) return B200_EINVAL;
  if (b == 0 || m == 0) return 0;
  if (!xyz || !verdict || n == 0) return B200_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (!g_spec_enabled || !spec_eligible(b, n, m)) {             // undecidable here: "run the chain" for every cloud
    B200_CUDA(cudaMemsetAsync(verdict, 0xff, sizeof(int32_t) * (size_t)b * m, st));
    return 0;
  }
  int bs = 1, bits = 0;
  while (bs * 2 <= n && bs < 512) { bs *= 2; ++bits; }
  fps_prefix_dist_kernel<<<dim3((unsigned)ceil_div(m, SPEC_PTS), (unsigned)b), SPEC_NT, 0, st>>>(xyz, cloud_stride, n, m,
                                                                                               verdict);
  B200_LAUNCH_CHECK();
  fps_prefix_check_kernel<<<dim3((unsigned)ceil_div(n, CHK_PTS), (unsigned)b), SPEC_NT, 0, st>>>(xyz, cloud_stride, n,
                                                                                                m, bits, verdict);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_furthest_point_sampling_spec(const float *xyz, int b, int n, int m, const int32_t *verdict0,
                                                 int stride0, const int32_t *verdict1, int stride1,
                                                 const int32_t *verdict2, int stride2, int32_t *idx, void *stream) {
  // the chain of verdicts must start at slot 0 (fps_identity treats an empty slot 0 as "no shortcut")
  FpsVerdicts v = {{verdict0, verdict0 ? verdict1 : nullptr, (verdict0 && verdict1) ? verdict2 : nullptr},
                   {stride0, stride1, stride2}};
  return fps_impl(xyz, b, n, m, &v, idx, stream);
}
