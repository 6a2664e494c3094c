// query.cu -- ball query and (multi-depth) cylinder query for sm_100a.
//
// Replaces query_ball_point_kernel (reference pointnet2/_ext_src/src/ball_query_gpu.cu:14-49)
// and query_cylinder_point_kernel (cylinder_query_gpu.cu:9-57).
//
// The legacy kernels run ONE block per batch element with a thread per centre
// doing a private serial scan of all n points (stride-3 scalar loads).  Here:
// grid over (batch x centre-tiles); one WARP per centre; the point stream is
// staged through shared memory in tiles shared by the CTA's 8 centres (raw
// xyz floats, read back at stride 3 = conflict free); lanes test 32 consecutive
// points at once and __ballot_sync + popc of the lower lanes gives every hit its
// slot in index order -- the same "first nsample hits in index order, remaining
// slots = first hit, no hit = zeros" result (ball_query_gpu.cu:38-46 and the
// torch::zeros of ball_query.cpp:24-26).  A warp stops scanning once its lists
// are full; the CTA stops loading tiles once all its warps have.
#include "common.cuh"
#include "query_pred.cuh"

namespace {

constexpr int QT = 256;          // threads per CTA
constexpr int QW = QT / 32;      // centres per CTA
constexpr int TP = 1024;         // points per shared-memory tile (12 KB)

// NH = number of output lists (1 for ball query / single cylinder query).
// BALL: lists have no upper bound test.
template <class Pred, int NH, bool BALL>
__global__ void __launch_bounds__(QT)
query_kernel(const float *__restrict__ new_xyz, const float *__restrict__ xyz,
             const float *__restrict__ rot, int b, int n, int m, float radius, float hmin,
             HList hmax, int nsample, int32_t *__restrict__ idx) {
  __shared__ float tile[TP * 3];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tiles_per_batch = (m + QW - 1) / QW;
  const int bi = blockIdx.x / tiles_per_batch;
  const int j = (blockIdx.x % tiles_per_batch) * QW + warp;
  const bool active = j < m;
  const float *pts = xyz + (size_t)bi * n * 3;

  Pred pred;
  if (active) {
    pred.init(new_xyz + ((size_t)bi * m + j) * 3, BALL ? nullptr : rot + ((size_t)bi * m + j) * 9,
              radius, hmin);
  }
  int cnt[NH], first[NH];
#pragma unroll
  for (int h = 0; h < NH; ++h) { cnt[h] = 0; first[h] = 0; }
  const size_t list_stride = (size_t)b * m * nsample;
  int32_t *out = idx + ((size_t)bi * m + (active ? j : 0)) * nsample;
  bool done = !active;

  for (int t0 = 0; t0 < n; t0 += TP) {
    const int tn = min(TP, n - t0);
    const float *src = pts + (size_t)t0 * 3;
    for (int i = tid; i < tn * 3; i += QT) tile[i] = __ldg(src + i);
    __syncthreads();
    if (!done) {
      for (int i0 = 0; i0 < tn; i0 += 32) {
        const int k = i0 + lane;
        bool ok = false;
        float hv = 0.f;
        if (k < tn) ok = pred.test(tile[3 * k], tile[3 * k + 1], tile[3 * k + 2], hv);
        if (!__any_sync(0xffffffffu, ok)) continue;           // most 32-point groups hold no candidate at all
        bool all_full = true;
#pragma unroll
        for (int h = 0; h < NH; ++h) {
          if (cnt[h] < nsample) {                       // warp-uniform
            const bool hit = BALL ? ok : (ok && hv < hmax.v[h]);
            const unsigned bal = __ballot_sync(0xffffffffu, hit);
            if (bal) {
              if (cnt[h] == 0) first[h] = t0 + i0 + __ffs(bal) - 1;
              const int pos = cnt[h] + __popc(bal & lanemask_lt());
              if (hit && pos < nsample) out[h * list_stride + pos] = t0 + k;
              cnt[h] += __popc(bal);
            }
            if (cnt[h] < nsample) all_full = false;
          }
        }
        if (all_full) { done = true; break; }
      }
    }
    if (__syncthreads_and(done)) break;
  }
  if (active) {
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      const int fill = cnt[h] > 0 ? first[h] : 0;
      for (int s = min(cnt[h], nsample) + lane; s < nsample; s += 32) out[h * list_stride + s] = fill;
    }
  }
}

// ---------------------------------------------------------------------------------------
// Segmented variant for large clouds.  One warp per centre leaves a 1024-centre query with
// ~7 warps per SM, each walking the whole cloud alone (latency-bound).  Here a CTA handles
// CPB centres x SEG contiguous point segments (one warp each): every warp collects the
// first nsample hits of ITS segment into shared memory, then the lists are concatenated in
// segment order -- the same "first nsample hits in index order" result with SEG x the
// parallelism.  A warp stops when its own lists are full or when the segments before it
// have already filled them.
constexpr int SEG = 4;            // point segments per centre
constexpr int CPB = QW / SEG;     // centres per CTA
constexpr int TPS = 512;          // points per segment tile (6 KB each)
constexpr int MAXNS = 64;         // list length the shared-memory lists are sized for

template <class Pred, int NH, bool BALL>
__global__ void __launch_bounds__(QT)
query_seg_kernel(const float *__restrict__ new_xyz, const float *__restrict__ xyz,
                 const float *__restrict__ rot, int b, int n, int m, float radius, float hmin,
                 HList hmax, int nsample, int seglen, int32_t *__restrict__ idx) {
  __shared__ float tile[SEG][TPS * 3];
  __shared__ int hits[CPB][SEG][NH][MAXNS];
  __shared__ int cnts[CPB][SEG][NH];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int c = warp / SEG, seg = warp % SEG;
  const int groups_per_batch = (m + CPB - 1) / CPB;
  const int bi = blockIdx.x / groups_per_batch;
  const int j = (blockIdx.x % groups_per_batch) * CPB + c;
  const bool active = j < m;
  const float *pts = xyz + (size_t)bi * n * 3;
  if (tid < CPB * SEG * NH) (&cnts[0][0][0])[tid] = 0;

  Pred pred;
  if (active)
    pred.init(new_xyz + ((size_t)bi * m + j) * 3, BALL ? nullptr : rot + ((size_t)bi * m + j) * 9, radius, hmin);
  int cnt[NH];
#pragma unroll
  for (int h = 0; h < NH; ++h) cnt[h] = 0;
  const int s0 = seg * seglen, s1 = min(n, s0 + seglen);        // this warp's points
  bool done = !active || s0 >= s1;

  for (int t0 = 0; t0 < seglen; t0 += TPS) {
    __syncthreads();                                            // previous tile consumed, counts visible
    // segments before this one may already have filled every list
    if (!done) {
      bool need = false;
#pragma unroll
      for (int h = 0; h < NH; ++h) {
        int before = 0;
        for (int s = 0; s < seg; ++s) before += cnts[c][s][h];
        if (before < nsample && cnt[h] < nsample) need = true;
      }
      if (!need) done = true;
    }
#pragma unroll
    for (int s = 0; s < SEG; ++s) {
      const int lo = s * seglen + t0, hi = min(n, min((s + 1) * seglen, lo + TPS));
      const float *src = pts + (size_t)lo * 3;
      for (int i = tid; i < (hi - lo) * 3; i += QT) tile[s][i] = __ldg(src + i);
    }
    __syncthreads();
    if (!done) {
      const int base = s0 + t0;
      const int tn = min(TPS, s1 - base);
      for (int i0 = 0; i0 < tn; i0 += 32) {
        const int k = i0 + lane;
        bool ok = false;
        float hv = 0.f;
        if (k < tn) ok = pred.test(tile[seg][3 * k], tile[seg][3 * k + 1], tile[seg][3 * k + 2], hv);
        if (!__any_sync(0xffffffffu, ok)) continue;           // most 32-point groups hold no candidate at all
        bool all_full = true;
#pragma unroll
        for (int h = 0; h < NH; ++h) {
          if (cnt[h] < nsample) {                               // warp-uniform
            const bool hit = BALL ? ok : (ok && hv < hmax.v[h]);
            const unsigned bal = __ballot_sync(0xffffffffu, hit);
            if (bal) {
              const int pos = cnt[h] + __popc(bal & lanemask_lt());
              if (hit && pos < nsample) hits[c][seg][h][pos] = base + k;
              cnt[h] = min(nsample, cnt[h] + __popc(bal));
            }
            if (cnt[h] < nsample) all_full = false;
          }
        }
        if (all_full) { done = true; break; }
      }
      if (lane < NH) {
        int v = 0;
#pragma unroll
        for (int h = 0; h < NH; ++h) if (lane == h) v = cnt[h];
        cnts[c][seg][lane] = v;
      }
    }
    if (__syncthreads_and(done)) break;
  }
  __syncthreads();
  // concatenate the segment lists in order; warp (c, seg) writes lists seg, seg+SEG, ...
  if (active) {
    const size_t list_stride = (size_t)b * m * nsample;
    for (int h = seg; h < NH; h += SEG) {
      int32_t *out = idx + h * list_stride + ((size_t)bi * m + j) * nsample;
      int total = 0, first = 0;
      bool any = false;
      for (int s = 0; s < SEG && total < nsample; ++s) {
        const int cs = min(cnts[c][s][h], nsample - total);
        if (cs > 0 && !any) { any = true; first = hits[c][s][h][0]; }
        for (int i = lane; i < cs; i += 32) out[total + i] = hits[c][s][h][i];
        total += cs;
      }
      for (int i = total + lane; i < nsample; i += 32) out[i] = any ? first : 0;
    }
  }
}

template <class Pred, int NH, bool BALL>
int launch_query(const float *new_xyz, const float *xyz, const float *rot, int b, int n, int m,
                 float radius, float hmin, const HList &hmax, int nsample, int32_t *idx,
                 cudaStream_t st) {
  if (!BALL && n >= 4096 && nsample <= MAXNS) {   // (ball queries fill up early; segments would mostly be wasted)
    const long long grid = (long long)b * ceil_div(m, CPB);
    if (grid > 0x7fffffffLL) return B200_ECAPACITY;
    const int seglen = ceil_div(ceil_div(n, SEG), 32) * 32;
    query_seg_kernel<Pred, NH, BALL><<<(unsigned)grid, QT, 0, st>>>(new_xyz, xyz, rot, b, n, m, radius, hmin,
                                                                    hmax, nsample, seglen, idx);
    B200_LAUNCH_CHECK();
    return 0;
  }
  const long long grid = (long long)b * ceil_div(m, QW);
  if (grid > 0x7fffffffLL) return B200_ECAPACITY;
  query_kernel<Pred, NH, BALL><<<(unsigned)grid, QT, 0, st>>>(new_xyz, xyz, rot, b, n, m, radius,
                                                              hmin, hmax, nsample, idx);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // namespace

extern "C" int b200_ball_query(const float *new_xyz, const float *xyz, int b, int n, int m,
                               float radius, int nsample, int32_t *idx, void *stream) {
  if (b < 0 || n < 0 || m < 0 || nsample < 0) return B200_EINVAL;
  if (b == 0 || m == 0 || nsample == 0) return 0;
  if (!new_xyz || !idx || (n > 0 && !xyz)) return B200_EINVAL;
  HList h = {};
  return launch_query<BallPred, 1, true>(new_xyz, xyz, nullptr, b, n, m, radius, 0.f, h, nsample,
                                         idx, (cudaStream_t)stream);
}

extern "C" int b200_cylinder_query_multi(const float *new_xyz, const float *xyz, const float *rot,
                                         int b, int n, int m, float radius, float hmin,
                                         const float *hmax, int nh, int nsample, int32_t *idx,
                                         void *stream) {
  if (b < 0 || n < 0 || m < 0 || nsample < 0 || nh < 0) return B200_EINVAL;
  if (nh > MAXH) return B200_ECAPACITY;
  if (b == 0 || m == 0 || nsample == 0 || nh == 0) return 0;
  if (!new_xyz || !rot || !idx || !hmax || (n > 0 && !xyz)) return B200_EINVAL;
  HList h = {};
  for (int i = 0; i < nh; ++i) h.v[i] = hmax[i];
  cudaStream_t st = (cudaStream_t)stream;
#define CYL_CASE(N) case N: return launch_query<CylPred, N, false>(new_xyz, xyz, rot, b, n, m, radius, hmin, h, nsample, idx, st);
  switch (nh) {
    CYL_CASE(1) CYL_CASE(2) CYL_CASE(3) CYL_CASE(4) CYL_CASE(5) CYL_CASE(6) CYL_CASE(7) CYL_CASE(8)
  }
#undef CYL_CASE
  return B200_EINVAL;
}

extern "C" int b200_cylinder_query(const float *new_xyz, const float *xyz, const float *rot, int b,
                                   int n, int m, float radius, float hmin, float hmax, int nsample,
                                   int32_t *idx, void *stream) {
  return b200_cylinder_query_multi(new_xyz, xyz, rot, b, n, m, radius, hmin, &hmax, 1, nsample, idx,
                                   stream);
}
