// fps.cu -- furthest point sampling for sm_100a.
//
// Replaces furthest_point_sampling_kernel (reference pointnet2/_ext_src/src/sampling_gpu.cu:74-178).
//
// Design: one thread-block CLUSTER of C CTAs per cloud (C = 1 for small clouds).
// Every thread keeps its points' x,y,z AND their running min-distance in
// registers for the whole m-step chain, so a step touches no global or shared
// memory for point data at all (the legacy kernel re-reads 12 B/point from L2 and
// reads+writes a 4 B/point HBM temp every step).  Per step:
//   registers: d = |p - last|^2, md = min(md, d), thread-local argmax
//   warp:      redux.sync max on the float bits, redux.sync min on the tie priority
//   CTA:       one __syncthreads over 16 per-warp slots (double buffered)
//   cluster:   the CTA winner {key, prio, x, y, z} is pushed into every peer's
//              shared memory (DSMEM), one cluster barrier, every CTA picks the
//              same global winner and already holds its coordinates.
//
// Exactness: the winner is  max d2, then min bitrev(k mod BS) over log2(BS) bits,
// then min k, with BS = largest power of two <= n clamped to 512 -- the
// closed form of the legacy per-thread strided scan + stride-halving tree
// (oracle/geom_oracle.c: oracle_fps == oracle_fps_tree).  Distances use the
// legacy FMA association; the skip test `(double)mag <= 1e-3` is done in fp64.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace {

constexpr int NT = 512;        // threads per CTA (16 warps)
constexpr int NW = NT / 32;
constexpr int MAXC = 16;       // largest cluster supported

struct __align__(16) Slot {
  int key;          // float bits of the candidate's min-distance (negative => none)
  unsigned prio;    // tie priority (smaller wins); encodes the point index
  float x, y;
  float z;
  int pad0, pad1, pad2;
};

__device__ __forceinline__ unsigned prio_of(int k, int bits) {
  const unsigned r = bits ? (__brev((unsigned)k) >> (32 - bits)) : 0u;  // bitrev of the low `bits` bits
  return (r << 22) | ((unsigned)k >> bits);
}
__device__ __forceinline__ int index_of(unsigned prio, int bits) {
  const unsigned r = prio >> 22;
  const unsigned lo = bits ? (__brev(r) >> (32 - bits)) : 0u;
  return (int)(((prio & 0x3fffffu) << bits) | lo);
}

template <int PPT, bool CLUSTER>
__global__ void __launch_bounds__(NT, 1)
fps_kernel(const float *__restrict__ xyz, int n, int m, int C, int chunk, int bits,
           int32_t *__restrict__ idx) {
  __shared__ int w_key[2][NW];
  __shared__ unsigned w_prio[2][NW];
  __shared__ Slot c_slot[2][MAXC];

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int rank = CLUSTER ? (int)(blockIdx.x % C) : 0;
  const int cloud = CLUSTER ? (int)(blockIdx.x / C) : (int)blockIdx.x;
  const float *p = xyz + (size_t)cloud * n * 3;
  int32_t *out = idx + (size_t)cloud * m;

  // chunk is a multiple of 512, so every point of one thread shares k mod BS and
  // a lower register slot i always has the smaller tie priority.
  const int base = rank * chunk;
  const int hi = min(n, base + chunk);

  float x[PPT], y[PPT], z[PPT], md[PPT];
#pragma unroll
  for (int i = 0; i < PPT; ++i) {
    const int k = base + i * NT + tid;
    if (k < hi) {
      x[i] = __ldg(p + 3 * (size_t)k);
      y[i] = __ldg(p + 3 * (size_t)k + 1);
      z[i] = __ldg(p + 3 * (size_t)k + 2);
      const float mag = sq3_p2(x[i], y[i], z[i]);
      md[i] = ((double)mag <= 1e-3) ? -1.0f : 1e10f;   // -1: never updated, never selected
    } else {
      x[i] = y[i] = z[i] = 0.f;
      md[i] = -1.0f;
    }
  }

  float ox = __ldg(p), oy = __ldg(p + 1), oz = __ldg(p + 2);
  if (rank == 0 && tid == 0) out[0] = 0;

  for (int j = 1; j < m; ++j) {
    const int par = j & 1;
    // ---- registers: update min-distances, thread-local first strict maximum
    float best = -1.0f;
    int bi = 0;
#pragma unroll
    for (int i = 0; i < PPT; ++i) {
      const float d = sq3_p2(__fsub_rn(x[i], ox), __fsub_rn(y[i], oy), __fsub_rn(z[i], oz));
      const float d2 = fminf(d, md[i]);     // md = -1 stays -1 (d >= 0 or NaN)
      md[i] = d2;
      if (d2 > best) { best = d2; bi = i; }
    }
    // ---- warp: max key, then min priority among the lanes holding it
    const int lkey = __float_as_int(best);            // >= 0 for real candidates, < 0 for none
    const int wkey = __reduce_max_sync(0xffffffffu, lkey);
    const unsigned lprio = (lkey == wkey && lkey >= 0) ? prio_of(base + bi * NT + tid, bits) : 0xffffffffu;
    const unsigned wprio = __reduce_min_sync(0xffffffffu, lprio);
    if (lane == 0) { w_key[par][warp] = wkey; w_prio[par][warp] = wprio; }
    __syncthreads();
    // ---- CTA: every warp reduces the 16 per-warp slots
    const int skey = lane < NW ? w_key[par][lane] : (int)0x80000000;
    const unsigned sprio = lane < NW ? w_prio[par][lane] : 0xffffffffu;
    const int ckey = __reduce_max_sync(0xffffffffu, skey);
    const unsigned cprio = __reduce_min_sync(0xffffffffu, skey == ckey ? sprio : 0xffffffffu);
    // ---- publish the CTA winner (with its coordinates) to every CTA of the cluster
    const bool owner = (ckey >= 0) ? (lprio == cprio && lkey == ckey) : (tid == 0);
    const unsigned own_mask = __ballot_sync(0xffffffffu, owner);
    if (own_mask) {                                   // warp-uniform: exactly one warp of the CTA
      float sx = 0.f, sy = 0.f, sz = 0.f;
      if (owner && ckey >= 0) {
#pragma unroll
        for (int i = 0; i < PPT; ++i)
          if (i == bi) { sx = x[i]; sy = y[i]; sz = z[i]; }
      }
      const int src = __ffs(own_mask) - 1;
      sx = __shfl_sync(0xffffffffu, sx, src);
      sy = __shfl_sync(0xffffffffu, sy, src);
      sz = __shfl_sync(0xffffffffu, sz, src);
      if (lane < C) {
        Slot s;
        s.key = ckey; s.prio = cprio; s.x = sx; s.y = sy; s.z = sz;
        s.pad0 = s.pad1 = s.pad2 = 0;
        Slot *dst = &c_slot[par][rank];
        if (CLUSTER) dst = cg::this_cluster().map_shared_rank(dst, lane);
        *dst = s;
      }
    }
    if (CLUSTER) cg::this_cluster().sync(); else __syncthreads();
    // ---- every CTA picks the same global winner
    const int gk = lane < C ? c_slot[par][lane].key : (int)0x80000000;
    const unsigned gp = lane < C ? c_slot[par][lane].prio : 0xffffffffu;
    const int gkey = __reduce_max_sync(0xffffffffu, gk);
    const unsigned gprio = __reduce_min_sync(0xffffffffu, gk == gkey ? gp : 0xffffffffu);
    const unsigned wmask = __ballot_sync(0xffffffffu, gk == gkey && gp == gprio);
    const int wr = __ffs(wmask) - 1;
    int sel = 0;
    if (gkey >= 0) {
      ox = c_slot[par][wr].x; oy = c_slot[par][wr].y; oz = c_slot[par][wr].z;
      sel = index_of(gprio, bits);
    }  // else: every point is skipped -> legacy yields index 0 and nothing is ever updated
    if (rank == 0 && tid == 0) out[j] = sel;
  }
}

template <int PPT>
int launch(const float *xyz, int b, int n, int m, int C, int chunk, int bits, int32_t *idx,
           cudaStream_t st) {
  if (C == 1) {
    fps_kernel<PPT, false><<<b, NT, 0, st>>>(xyz, n, m, 1, chunk, bits, idx);
    B200_LAUNCH_CHECK();
    return 0;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(b * C), 1, 1);
  cfg.blockDim = dim3(NT, 1, 1);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (C > 8)
    B200_CUDA(cudaFuncSetAttribute(fps_kernel<PPT, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  B200_CUDA(cudaLaunchKernelEx(&cfg, fps_kernel<PPT, true>, xyz, n, m, C, chunk, bits, idx));
  return 0;
}

}  // namespace

// Test / tuning hook: force the cluster size (0 = automatic).
static int g_force_cluster = 0;
extern "C" void b200_fps_set_cluster(int c) { g_force_cluster = c; }

extern "C" int b200_furthest_point_sampling(const float *xyz, int b, int n, int m, int32_t *idx,
                                            void *stream) {
  if (b < 0 || m < 0 || n < 0) return B200_EINVAL;
  if (b == 0 || m == 0) return 0;
  if (n <= 0 || !xyz || !idx) return B200_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  int bs = 1, bits = 0;
  while (bs * 2 <= n && bs < 512) { bs *= 2; ++bits; }

  // cluster size: as many CTAs per cloud as keep >= 2048 points per CTA, at most 8,
  // and no more than what lets every cloud of the batch be resident at once.
  int C = 1;
  const int sms = b200_sm_count();
  while (C < 8 && n / (C * 2) >= 2048 && (long long)b * (C * 2) <= sms) C *= 2;
  while (ceil_div(ceil_div(n, C), NT) > 24 && C < MAXC) C *= 2;   // capacity: 24 points/thread
  if (g_force_cluster > 0) C = g_force_cluster;
  int chunk = ceil_div(ceil_div(n, C), 512) * 512;
  const int ppt = chunk / NT;
  if (ppt > 24) return B200_ECAPACITY;
#define FPS_CASE(P) if (ppt <= P) return launch<P>(xyz, b, n, m, C, chunk, bits, idx, st);
  FPS_CASE(1) FPS_CASE(2) FPS_CASE(3) FPS_CASE(4) FPS_CASE(5) FPS_CASE(6) FPS_CASE(8)
  FPS_CASE(10) FPS_CASE(12) FPS_CASE(16) FPS_CASE(20) FPS_CASE(24)
#undef FPS_CASE
  return B200_ECAPACITY;
}
