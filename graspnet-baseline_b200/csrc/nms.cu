// nms.cu -- pairwise translation/rotation grasp NMS as a bitmask kernel (sm_100a).
//
// Replaces GraspGroup.nms -> grasp_nms.nms_grasp (graspnetAPI / grasp_nms, third
// party, not in the reference tree; call sites demo_pointcloud.py:249,
// grasp_service.py:276).  Semantics = oracle/geom_oracle.c:oracle_grasp_nms:
// stable descending score order; a kept grasp i suppresses every later j with
//   sqrt(|t_i - t_j|^2) < translation_thresh  and  (trace(R_i R_j^T) - 1)/2 > cos(rot_thresh)
// evaluated in fp64 with a fixed association, so the keep set is bit-exact.
//
// Three launches, no host round trip:
//   1. rank kernel   : rank[i] = #{j : s_j > s_i or (s_j == s_i and j < i)}  -> order[]
//   2. mask kernel   : 64x64 tiles of the upper triangle; thread = row, 64-bit word
//                      of "row suppresses column" bits (the torchvision-NMS layout)
//   3. sweep kernel  : one CTA; per 64-row block the diagonal word chain is resolved
//                      serially from shared memory by one thread, then the kept rows are
//                      OR-ed into the remaining removed-words (warp = column word, lane = row,
//                      one load per lane + a warp OR reduction); finally a popc prefix
//                      compacts the survivors.
#include "post_common.cuh"

namespace {

constexpr int G = 17;       // floats per grasp row (models/graspnet.py:131-133)
constexpr int RT = 256;

__global__ void __launch_bounds__(RT)
nms_rank_kernel(const float *__restrict__ grasps, int m, const int32_t *__restrict__ m_dev,
                int32_t *__restrict__ order) {
  __shared__ float sc[RT];
  if (m_dev) m = min(m, __ldg(m_dev));         // row count produced on the device (decode / compaction)
  if ((int)(blockIdx.x * RT) >= m) return;
  const int i = blockIdx.x * RT + threadIdx.x;
  // scores are ranked under a TOTAL order -- NaN counts as -inf, ties by row number -- so that the ranks are always a
  // permutation (with the plain comparisons NaN rows collided and left order[] slots unwritten: garbage row numbers)
  auto key = [](float s) { return nms_score_key(s); };
  const float si = i < m ? key(__ldg(grasps + (size_t)i * G)) : 0.f;
  int rank = 0;
  for (int j0 = 0; j0 < m; j0 += RT) {
    const int j = j0 + threadIdx.x;
    __syncthreads();
    sc[threadIdx.x] = j < m ? key(__ldg(grasps + (size_t)j * G)) : 0.f;
    __syncthreads();
    const int jn = min(RT, m - j0);
    for (int t = 0; t < jn; ++t) {
      const float sj = sc[t];
      rank += (sj > si) | ((sj == si) & (j0 + t < i));
    }
  }
  if (i < m) order[rank] = i;
}

// Does grasp a (translation at, rotation ar, object id ao) suppress grasp c?  The arithmetic of the upstream
// grasp_nms.nms_grasp is not available (PyPI package, no source in the reference tree), so every open choice of the
// restatement is a switch (b200_grasp_nms_set_options, include/b200grasp.h); flags = 0 is the documented default:
// both thresholds strict, fp64 arithmetic on the fp32 rows, cosine comparison, object ids ignored.
template <typename T>
__device__ __forceinline__ bool nms_pair(const T (&at)[3], const T (&ar)[9], const double *ct, const double *cr, float ao,
                                         float co, T tthr, T cthr, T athr, unsigned flags) {
  if ((flags & B200_NMS_SAME_OBJECT) && ao != co) return false;
  const T dx = at[0] - (T)ct[0], dy = at[1] - (T)ct[1], dz = at[2] - (T)ct[2];
  const T dist = sqrt((dx * dx + dy * dy) + dz * dz);
  if (!((flags & B200_NMS_TRANS_INCLUSIVE) ? dist <= tthr : dist < tthr)) return false;
  T tr = (T)0;
#pragma unroll
  for (int e = 0; e < 9; ++e) tr = tr + ar[e] * (T)cr[e];
  const T ca = (tr - (T)1) / (T)2;
  if (flags & B200_NMS_ACOS) {                                        // compare the angle itself
    const T ang = acos(ca < (T)-1 ? (T)-1 : (ca > (T)1 ? (T)1 : ca));
    return (flags & B200_NMS_ROT_INCLUSIVE) ? ang <= athr : ang < athr;
  }
  return (flags & B200_NMS_ROT_INCLUSIVE) ? ca >= cthr : ca > cthr;
}

unsigned g_nms_flags = 0u;

__global__ void __launch_bounds__(64)
nms_mask_kernel(const float *__restrict__ grasps, const int32_t *__restrict__ order, int m,
                const int32_t *__restrict__ m_dev, int W, double tthr, double cthr, unsigned flags,
                unsigned long long *__restrict__ mask) {
  const int rb = blockIdx.y, cb = blockIdx.x;
  if (m_dev) m = min(m, __ldg(m_dev));
  if (cb < rb || cb * 64 >= m) return;       // lower triangle never read by the sweep; W = row stride (max rows)
  __shared__ double ct[64][3];
  __shared__ double cr[64][9];
  __shared__ float cobj[64];
  const int t = threadIdx.x;
  const int c = cb * 64 + t;
  if (c < m) {
    const float *g = grasps + (size_t)__ldg(order + c) * G;
    cobj[t] = __ldg(g + 16);
#pragma unroll
    for (int e = 0; e < 3; ++e) ct[t][e] = (double)__ldg(g + 13 + e);
#pragma unroll
    for (int e = 0; e < 9; ++e) cr[t][e] = (double)__ldg(g + 4 + e);
  }
  __syncthreads();
  const int a = rb * 64 + t;
  if (a >= m) return;
  const float *g = grasps + (size_t)__ldg(order + a) * G;
  double at[3], ar[9];
#pragma unroll
  for (int e = 0; e < 3; ++e) at[e] = (double)__ldg(g + 13 + e);
#pragma unroll
  for (int e = 0; e < 9; ++e) ar[e] = (double)__ldg(g + 4 + e);
  const float aobj = __ldg(g + 16);
  unsigned long long bits = 0ull;
  const int cn = min(64, m - cb * 64);
  if (flags == 0u) {                                                  // the default: exact fp64, fixed association
    for (int k = 0; k < cn; ++k) {
      if (cb * 64 + k <= a) continue;          // only later (lower-score) grasps
      const double dx = __dsub_rn(at[0], ct[k][0]), dy = __dsub_rn(at[1], ct[k][1]),
                   dz = __dsub_rn(at[2], ct[k][2]);
      const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
      if (!(__dsqrt_rn(d2) < tthr)) continue;
      double tr = 0.0;
#pragma unroll
      for (int e = 0; e < 9; ++e) tr = __dadd_rn(tr, __dmul_rn(ar[e], cr[k][e]));
      const double ca = __ddiv_rn(__dsub_rn(tr, 1.0), 2.0);
      if (ca > cthr) bits |= 1ull << k;
    }
  } else if (flags & B200_NMS_FP32) {
    float atf[3], arf[9];
#pragma unroll
    for (int e = 0; e < 3; ++e) atf[e] = (float)at[e];
#pragma unroll
    for (int e = 0; e < 9; ++e) arf[e] = (float)ar[e];
    const float athr = acosf((float)cthr);
    for (int k = 0; k < cn; ++k)
      if (cb * 64 + k > a && nms_pair<float>(atf, arf, ct[k], cr[k], aobj, cobj[k], (float)tthr, (float)cthr, athr, flags))
        bits |= 1ull << k;
  } else {
    const double athr = acos(cthr);
    for (int k = 0; k < cn; ++k)
      if (cb * 64 + k > a && nms_pair<double>(at, ar, ct[k], cr[k], aobj, cobj[k], tthr, cthr, athr, flags))
        bits |= 1ull << k;
  }
  mask[(size_t)a * W + cb] = bits;
}

constexpr int ST = 256;
constexpr int MAXW = 512;                    // 32768 grasps / 64

// One CTA.  Per 64-row block: (a) the diagonal word chain is resolved serially by one thread from shared memory
// (the 64 words were prefetched during the previous block); (b) every kept row's remaining words are OR-ed into
// the removed set: a warp takes a column word, lane = row, ONE independent global load per lane and a warp OR
// reduction -- a single memory latency per block instead of one per kept row.
__global__ void __launch_bounds__(ST)
nms_sweep_kernel(const unsigned long long *__restrict__ mask, const int32_t *__restrict__ order, int m,
                 const int32_t *__restrict__ m_dev, int Ws, int32_t *__restrict__ keep,
                 int32_t *__restrict__ nkeep, const float *__restrict__ rows = nullptr,
                 float *__restrict__ out_rows = nullptr) {
  __shared__ unsigned long long removed[MAXW];
  __shared__ unsigned long long diag[64];
  __shared__ unsigned long long s_keep;
  __shared__ int s_prefix[ST];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (m_dev) m = min(m, __ldg(m_dev));
  const int W = (m + 63) / 64;                 // words in use; Ws = row stride of the mask
  for (int w = tid; w < W; w += ST) removed[w] = 0ull;
  unsigned long long dnext = 0ull;             // diagonal word of row (wb*64 + tid), fetched one block ahead
  if (tid < 64 && tid < m) dnext = mask[(size_t)tid * Ws];
  for (int wb = 0; wb < W; ++wb) {
    if (tid < 64) {
      diag[tid] = dnext;
      const int a = (wb + 1) * 64 + tid;
      dnext = (wb + 1 < W && a < m) ? mask[(size_t)a * Ws + wb + 1] : 0ull;
    }
    __syncthreads();
    if (tid == 0) {
      const int rn = min(64, m - wb * 64);
      unsigned long long rem = removed[wb] | (rn < 64 ? ~0ull << rn : 0ull), kb = 0ull;
#pragma unroll 16
      for (int bit = 0; bit < 64; ++bit) {
        const unsigned long long d = diag[bit];
        const unsigned long long alive = ((rem >> bit) & 1ull) ^ 1ull;
        kb |= alive << bit;
        rem |= d & (0ull - alive);
      }
      s_keep = kb;
      removed[wb] = ~kb;                     // final: bit set = not kept (padding bits too)
    }
    __syncthreads();
    const unsigned long long kb = s_keep;
    const unsigned kbl = (unsigned)kb, kbh = (unsigned)(kb >> 32);
    const bool k0 = (kbl >> lane) & 1u, k1 = (kbh >> lane) & 1u;
    const unsigned long long *r0 = mask + (size_t)(wb * 64 + lane) * Ws, *r1 = r0 + (size_t)32 * Ws;
    for (int w = wb + 1 + warp; w < W; w += ST / 32) {
      const unsigned long long v0 = k0 ? r0[w] : 0ull, v1 = k1 ? r1[w] : 0ull;
      const unsigned long long v = v0 | v1;
      const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)v);
      const unsigned hi = __reduce_or_sync(0xffffffffu, (unsigned)(v >> 32));
      if (lane == 0) removed[w] |= ((unsigned long long)hi << 32) | lo;
    }
    __syncthreads();
  }
  // compaction: survivors in sorted order; each thread owns a contiguous range of words
  const int per = (W + ST - 1) / ST;
  const int w0 = min(W, tid * per), w1 = min(W, w0 + per);
  int cnt = 0;
  for (int w = w0; w < w1; ++w) cnt += __popcll(~removed[w]);
  s_prefix[tid] = cnt;
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int t = 0; t < ST; ++t) { const int c = s_prefix[t]; s_prefix[t] = run; run += c; }
    *nkeep = run;
  }
  __syncthreads();
  int pos = s_prefix[tid];
  for (int w = w0; w < w1; ++w) {
    unsigned long long kb = ~removed[w];
    while (kb) {
      const int bit = __ffsll((long long)kb) - 1;
      kb &= kb - 1;
      keep[pos++] = __ldg(order + w * 64 + bit);
    }
  }
  if (out_rows) {                              // the fused chain: the kept rows themselves, in score order
    __syncthreads();                           // keep[] and s_prefix settled (same CTA: global writes are visible)
    __shared__ int s_total;
    if (tid == 0) s_total = *nkeep;
    __syncthreads();
    const int total = s_total * G;
    for (int t = tid; t < total; t += ST) {
      const int i = t / G, c = t - i * G;
      out_rows[t] = rows[(size_t)keep[i] * G + c];
    }
  }
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

extern "C" void b200_grasp_nms_set_options(unsigned flags) { g_nms_flags = flags; }

extern "C" size_t b200_grasp_nms_workspace(int m) {
  if (m <= 0) return 256;
  const size_t W = (size_t)(m + 63) / 64;
  return align_up(sizeof(int32_t) * (size_t)m, 256) + align_up(8 * W, 256) + align_up(8 * W * (size_t)m, 256);
}

static int nms_launch(const float *grasps, int m, const int32_t *m_dev, double translation_thresh,
                      double cos_rotation_thresh, int32_t *keep, int32_t *nkeep, void *workspace,
                      size_t workspace_bytes, cudaStream_t st) {
  if (m < 0 || !nkeep) return B200_EINVAL;
  if (m == 0) {
    B200_CUDA(cudaMemsetAsync(nkeep, 0, sizeof(int32_t), st));
    return 0;
  }
  if (!grasps || !keep || !workspace) return B200_EINVAL;
  if (m > 32768) return B200_ECAPACITY;
  if (workspace_bytes < b200_grasp_nms_workspace(m)) return B200_EWORKSPACE;
  const int W = (m + 63) / 64;
  char *ws = (char *)workspace;
  int32_t *order = (int32_t *)ws;
  ws += align_up(sizeof(int32_t) * (size_t)m, 256);
  ws += align_up(8 * (size_t)W, 256);          // (formerly the removed-words scratch; kept so the layout is stable)
  unsigned long long *mask = (unsigned long long *)ws;
  nms_rank_kernel<<<ceil_div(m, RT), RT, 0, st>>>(grasps, m, m_dev, order);
  B200_LAUNCH_CHECK();
  nms_mask_kernel<<<dim3((unsigned)W, (unsigned)W), 64, 0, st>>>(grasps, order, m, m_dev, W, translation_thresh,
                                                                cos_rotation_thresh, g_nms_flags, mask);
  B200_LAUNCH_CHECK();
  nms_sweep_kernel<<<1, ST, 0, st>>>(mask, order, m, m_dev, W, keep, nkeep);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_grasp_nms(const float *grasps, int m, double translation_thresh,
                              double cos_rotation_thresh, int32_t *keep, int32_t *nkeep,
                              void *workspace, size_t workspace_bytes, void *stream) {
  return nms_launch(grasps, m, nullptr, translation_thresh, cos_rotation_thresh, keep, nkeep, workspace,
                    workspace_bytes, (cudaStream_t)stream);
}

extern "C" int b200_grasp_nms_device(const float *grasps, const int32_t *m_dev, int m_max,
                                     double translation_thresh, double cos_rotation_thresh, int32_t *keep,
                                     int32_t *nkeep, void *workspace, size_t workspace_bytes, void *stream) {
  if (!m_dev) return B200_EINVAL;
  return nms_launch(grasps, m_max, m_dev, translation_thresh, cos_rotation_thresh, keep, nkeep, workspace,
                    workspace_bytes, (cudaStream_t)stream);
}

// mask + sweep for rows whose score order is already known (post_chain.cu); the sweep gathers the kept rows
int b200i_nms_ranked(const float *grasps, const int32_t *m_dev, int m_max, const int32_t *order, double translation_thresh,
                     double cos_rotation_thresh, int32_t *keep, int32_t *nkeep, float *out_rows,
                     unsigned long long *mask_ws, cudaStream_t st) {
  if (m_max < 0 || !nkeep || !m_dev) return B200_EINVAL;
  if (m_max == 0) {
    B200_CUDA(cudaMemsetAsync(nkeep, 0, sizeof(int32_t), st));
    return 0;
  }
  if (!grasps || !keep || !order || !mask_ws) return B200_EINVAL;
  if (m_max > 32768) return B200_ECAPACITY;
  const int W = (m_max + 63) / 64;
  nms_mask_kernel<<<dim3((unsigned)W, (unsigned)W), 64, 0, st>>>(grasps, order, m_max, m_dev, W, translation_thresh,
                                                                cos_rotation_thresh, g_nms_flags, mask_ws);
  B200_LAUNCH_CHECK();
  nms_sweep_kernel<<<1, ST, 0, st>>>(mask_ws, order, m_max, m_dev, W, keep, nkeep, grasps, out_rows);
  B200_LAUNCH_CHECK();
  return 0;
}

// ---- No-allocation variant (no counterpart in the reference) -----------
// For callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
// it): the same three launches with their scratch (order[m] | pad | bitmask m x ceil(m/64) words) in a file-scope
// device array instead of a caller workspace.  Every GPU that loads the library holds the array (532,736 bytes) until
// exit.  In HOST code `g_nms_static_scratch` is only a stand-in for the symbol, not a device address: the address is
// taken once per device, with that device current, through cudaGetSymbolAddress, and that pointer is what the
// kernels get.  Nothing in the scratch has to start at zero (rank, mask and sweep write every word they later read --
// the caller-workspace entry points are fed uninitialised memory too), so nothing is cleared; the array keeps the
// last call's contents.  ALL calls on a device share it: calls that overlap on the GPU (different streams, or graphs
// replayed concurrently) race -- serialise them on one stream.
namespace {
constexpr int NMS_STATIC_MAX_M = B200_NMS_STATIC_MAX_M;
constexpr size_t nms_static_align(size_t v) { return (v + 255) / 256 * 256; }
constexpr size_t NMS_STATIC_BYTES = nms_static_align(sizeof(int32_t) * (size_t)NMS_STATIC_MAX_M) +
                                    nms_static_align(8 * (size_t)((NMS_STATIC_MAX_M + 63) / 64)) +
                                    nms_static_align(8 * (size_t)((NMS_STATIC_MAX_M + 63) / 64) * NMS_STATIC_MAX_M);
}  // namespace

static __device__ __align__(256) unsigned char g_nms_static_scratch[NMS_STATIC_BYTES];

extern "C" int b200_grasp_nms_static(const float *grasps, int m, double translation_thresh,
                                     double cos_rotation_thresh, int32_t *keep, int32_t *nkeep, void *stream) {
  if (m < 0) return B200_EINVAL;
  if (m > NMS_STATIC_MAX_M) return B200_ESTATIC;
  static std::atomic<void *> per_device[64];             // device address of the array, per device
  const int dev = b200_device();
  void *ptr = (dev >= 0 && dev < 64) ? per_device[dev].load(std::memory_order_acquire) : nullptr;
  if (!ptr) {
    B200_CUDA(cudaGetSymbolAddress(&ptr, g_nms_static_scratch));    // the current device's copy
    if (!ptr) return B200_EINVAL;
    if (dev >= 0 && dev < 64) per_device[dev].store(ptr, std::memory_order_release);
  }
  return nms_launch(grasps, m, nullptr, translation_thresh, cos_rotation_thresh, keep, nkeep, ptr, NMS_STATIC_BYTES,
                    (cudaStream_t)stream);
}
