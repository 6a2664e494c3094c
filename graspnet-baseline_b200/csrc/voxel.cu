// voxel.cu -- voxel down-sample of the scene cloud behind ModelFreeCollisionDetector.__init__ (sm_100a).
//
// The reference calls open3d's PointCloud.voxel_down_sample (utils/collision_detector.py:32-35);
// open3d is a third-party dependency that is not in the reference tree, so this follows the
// restatement in oracle/post.py (PARITY UNPINNED):
//   origin = min_bound - voxel/2;  key = floor((p - origin) / voxel) per axis;
//   output = fp64 mean of the points of each occupied voxel, accumulated in POINT ORDER,
//   rows in ascending (kx, ky, kz) order.
// Everything stays on the device and is deterministic (no floating-point atomics):
//   1. bounds kernel  : per-axis minimum as an order-preserving 64-bit integer atomicMin
//   2. key kernel     : 3 x 21-bit voxel coordinates packed into one 63-bit key per point
//   3. stable radix sort of (key, point index)  [cub::DeviceRadixSort -- plain library sort]
//   4. head kernel    : flags the first element of every run of equal keys; exclusive scan
//                       [cub::DeviceScan] turns the flags into output row numbers
//   5. mean kernel    : the thread at each run head walks its run in sorted (= point) order,
//                       sequential fp64 adds, one divide; the last element publishes the row count.
// Clouds of up to SMALL_N points (one inference scene: 20000) take ONE launch instead of those seventeen:
// voxel_small_kernel does all five steps in a single CTA -- the keys of such a cloud use ~17 of the 63 bits, so the
// stable radix sort is three 8-bit passes (one per axis, over the bits the cloud's extent actually needs) instead of
// the library's eight over bits 0..63.  Same arithmetic, same order: bit-identical output (tests/test_gpu_postprocess.py).
#include <cub/cub.cuh>

#include "common.cuh"

namespace {

constexpr int VT = 256;

__device__ __forceinline__ long long order_key(double v) {      // monotone double -> int64
  const long long b = __double_as_longlong(v);
  return b ^ ((b >> 63) & 0x7fffffffffffffffLL);
}
__device__ __forceinline__ double order_val(long long k) {
  return __longlong_as_double(k ^ ((k >> 63) & 0x7fffffffffffffffLL));
}

__global__ void __launch_bounds__(VT)
voxel_bounds_kernel(const double *__restrict__ pts, int n, long long *__restrict__ mins) {
  long long m0 = 0x7fffffffffffffffLL, m1 = m0, m2 = m0;
  for (int i = blockIdx.x * VT + threadIdx.x; i < n; i += gridDim.x * VT) {
    const double x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
    if (!(isfinite(x) && isfinite(y) && isfinite(z))) continue;     // invalid points take no part (see voxel_key_kernel)
    m0 = min(m0, order_key(x));
    m1 = min(m1, order_key(y));
    m2 = min(m2, order_key(z));
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    m0 = min(m0, __shfl_xor_sync(0xffffffffu, m0, o));
    m1 = min(m1, __shfl_xor_sync(0xffffffffu, m1, o));
    m2 = min(m2, __shfl_xor_sync(0xffffffffu, m2, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(mins + 0, m0);
    atomicMin(mins + 1, m1);
    atomicMin(mins + 2, m2);
  }
}

__global__ void __launch_bounds__(VT)
voxel_key_kernel(const double *__restrict__ pts, int n, double voxel, const long long *__restrict__ mins,
                 unsigned long long *__restrict__ keys, int *__restrict__ order, int *__restrict__ overflow,
                 unsigned flags) {
  const int i = blockIdx.x * VT + threadIdx.x;
  if (i >= n) return;
  // grid origin: min_bound - voxel / 2 (Open3D's VoxelDownSample) unless B200_VOXEL_ORIGIN_AT_MIN
  const double half = (flags & B200_VOXEL_ORIGIN_AT_MIN) ? 0.0 : __dmul_rn(voxel, 0.5);
  // A point that is not finite, or falls outside 2^21 voxels per axis from the cloud's minimum corner (open3d raises
  // "voxel_size is too small" there), gets the key ~0: it sorts behind every voxel, is left out of the means and is
  // COUNTED in *overflow (b200_voxel_down_sample_checked hands the count out; the host API raises on it).
  unsigned long long key = 0;
  bool bad = false;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const double origin = __dsub_rn(order_val(mins[a]), half);
    const double q = floor(__ddiv_rn(__dsub_rn(pts[3 * (size_t)i + a], origin), voxel));
    bad |= !(q >= 0.0) || q >= 2097152.0;                               // 21 bits per axis; NaN / inf fail the first test
    key = (key << 21) | (unsigned long long)(bad ? 0 : (long long)q);
  }
  if (bad) {
    key = ~0ull;
    atomicAdd(overflow, 1);
  }
  keys[i] = key;
  order[i] = i;
}

__global__ void __launch_bounds__(VT)
voxel_head_kernel(const unsigned long long *__restrict__ keys, int n, int *__restrict__ head) {
  const int i = blockIdx.x * VT + threadIdx.x;
  if (i < n) head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

__global__ void __launch_bounds__(VT)
voxel_mean_kernel(const double *__restrict__ pts, const unsigned long long *__restrict__ keys,
                  const int *__restrict__ order, const int *__restrict__ head, const int *__restrict__ row,
                  int n, double *__restrict__ out, int *__restrict__ nvox, const int *__restrict__ overflow,
                  int *__restrict__ nbad) {
  const int i = blockIdx.x * VT + threadIdx.x;
  if (i >= n) return;
  if (i == n - 1) {                                    // row[] = exclusive scan of head, shifted by one below;
    *nvox = row[i] + 1 - (keys[i] == ~0ull ? 1 : 0);   // the run of excluded points, if any, is the last one
    if (nbad) *nbad = *overflow;
  }
  if (!head[i]) return;
  const unsigned long long k = keys[i];
  if (k == ~0ull) return;
  double sx = 0.0, sy = 0.0, sz = 0.0;
  int cnt = 0;
  for (int j = i; j < n && keys[j] == k; ++j) {        // sorted order is stable => ascending point index
    const double *p = pts + 3 * (size_t)order[j];
    sx = __dadd_rn(sx, p[0]); sy = __dadd_rn(sy, p[1]); sz = __dadd_rn(sz, p[2]);
    ++cnt;
  }
  const double c = (double)cnt;
  double *o = out + 3 * (size_t)row[i];
  o[0] = __ddiv_rn(sx, c); o[1] = __ddiv_rn(sy, c); o[2] = __ddiv_rn(sz, c);
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Layout {
  size_t mins, overflow, keys_a, keys_b, ord_a, ord_b, head, row, cub, total;
};

Layout layout(int n) {
  Layout L;
  size_t off = 0;
  const size_t nn = (size_t)(n > 0 ? n : 1);
  auto take = [&](size_t bytes) { const size_t o = off; off += align_up(bytes, 256); return o; };
  L.mins = take(3 * sizeof(long long));
  L.overflow = take(sizeof(int));
  L.keys_a = take(nn * 8); L.keys_b = take(nn * 8);
  L.ord_a = take(nn * 4); L.ord_b = take(nn * 4);
  L.head = take(nn * 4); L.row = take(nn * 4);
  size_t sort_bytes = 0, scan_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const unsigned long long *)nullptr, (unsigned long long *)nullptr,
                                  (const int *)nullptr, (int *)nullptr, (int)nn, 0, 63);
  cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, (const int *)nullptr, (int *)nullptr, (int)nn);
  L.cub = take(sort_bytes > scan_bytes ? sort_bytes : scan_bytes);
  L.total = off;
  return L;
}

__global__ void voxel_init_kernel(long long *mins, int *overflow, int *nvox) {
  mins[0] = mins[1] = mins[2] = 0x7fffffffffffffffLL;
  *overflow = 0;
  *nvox = 0;
}

__global__ void __launch_bounds__(VT)
voxel_row_fix_kernel(int *__restrict__ row, int n) {    // inclusive sum of heads -> 0-based row of each element
  const int i = blockIdx.x * VT + threadIdx.x;
  if (i < n) row[i] -= 1;
}


// ---------------------------------------------------------------------------------------------------------
// The single-CTA path.  Everything the sort touches lives in shared memory when the cloud's voxel coordinates fit
// 32 bits together (a 0.6 m scene at 1 cm: 6 + 6 + 4) and n <= SMEM_N: a compact 32-bit key per point and two
// 16-bit index arrays that the passes permute (8 n bytes); otherwise the same passes run on the 64-bit keys in the
// workspace (slower, same result).
constexpr int ST = 1024;                 // threads
constexpr int SW = ST / 32;              // warps
constexpr int SMALL_N = 32768;           // points at most
constexpr int SMEM_N = 22528;            // ... with the sort in shared memory (8 B per point + 35 KB of tables)

__device__ __forceinline__ long long warp_min_ll(long long v) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int warp_max_i(int v) { return __reduce_max_sync(0xffffffffu, v); }
__device__ __forceinline__ int bits_of(unsigned v) { return 32 - __clz(v); }      // 0 for v == 0

// One stable counting-sort pass on the `width` (<= 8) key bits from `shift`; element i of the input order has the
// key key_at(i) and is moved by put(pos, i).  Chunks of ST consecutive elements keep their order: an element's slot
// is its bin's running base + the number of equal digits in the warps before it (table column scan) + the number in
// the lanes before it (match mask).
template <class KeyAt, class Put>
__device__ __forceinline__ void small_sort_pass(int n, int shift, int width, int (*s_tab)[256], int *s_base,
                                                KeyAt key_at, Put put) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned dmask = (1u << width) - 1u;
  if (tid < 256) s_base[tid] = 0;
  for (int e = tid; e < SW * 256; e += ST) (&s_tab[0][0])[e] = 0;
  __syncthreads();
  for (int c0 = 0; c0 < n; c0 += ST) {                     // bin counts, one shared-memory atomic per warp and digit
    const int i = c0 + tid;
    const unsigned d = i < n ? (unsigned)(key_at(i) >> shift) & dmask : 0xffffu;
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    if (i < n && (peers & ((1u << lane) - 1u)) == 0) atomicAdd(&s_base[d], __popc(peers));
  }
  __syncthreads();
  if (warp == 0) {                                         // exclusive scan of the 256 bin counts: 8 bins per lane
    int c[8], tot = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e) { c[e] = s_base[8 * lane + e]; tot += c[e]; }
    int incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    int run = incl - tot;
#pragma unroll
    for (int e = 0; e < 8; ++e) { s_base[8 * lane + e] = run; run += c[e]; }
  }
  __syncthreads();
  for (int c0 = 0; c0 < n; c0 += ST) {
    const int i = c0 + tid;
    const bool valid = i < n;
    const unsigned d = valid ? (unsigned)(key_at(i) >> shift) & dmask : 0xffffu;   // lanes past the end match each other only
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    if (valid && rank == 0) s_tab[warp][d] = __popc(peers);
    __syncthreads();
    if (tid < 256) {                                       // column tid: counts -> first slot of each warp's group
      int c[SW];
#pragma unroll
      for (int w = 0; w < SW; ++w) c[w] = s_tab[w][tid];
      int run = s_base[tid];
#pragma unroll
      for (int w = 0; w < SW; ++w) {                       // (rows without this digit stay 0)
        if (c[w]) s_tab[w][tid] = run;
        run += c[w];
      }
      s_base[tid] = run;
    }
    __syncthreads();
    if (valid) put(s_tab[warp][d] + rank, i);
    __syncwarp();
    if (valid && rank == 0) s_tab[warp][d] = 0;            // this warp's row is clean again for the next chunk
  }
  __syncthreads();
}

// 4. runs of equal keys -> output rows; 5. the thread that meets a run's head walks the run (= point order):
// key_at(i) / point_at(i) = key and point number of the i-th element of the sorted order
template <class KeyAt, class PointAt, class IsBad>
__device__ __forceinline__ void small_finish(const double *__restrict__ pts, int n, int nbad_all, int *s_cnt,
                                             double *__restrict__ out, int *__restrict__ nvox, int *__restrict__ nbad,
                                             KeyAt key_at, PointAt point_at, IsBad is_bad) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int seg = (n + ST - 1) / ST;
  const int i0 = min(n, tid * seg), i1 = min(n, i0 + seg);
  int heads = 0;
  auto prev = i0 > 0 ? key_at(i0 - 1) : key_at(0);
  for (int i = i0; i < i1; ++i) {
    const auto k = key_at(i);
    heads += (i == 0 || k != prev);
    prev = k;
  }
  int incl = heads;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_cnt[warp] = incl;
  __syncthreads();
  const int c = s_cnt[lane];
  int wi = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, wi, o);
    if (lane >= o) wi += t;
  }
  const int wbase = __shfl_sync(0xffffffffu, wi - c, warp);
  const int total = __shfl_sync(0xffffffffu, wi, 31);
  int row = wbase + incl - heads;
  prev = i0 > 0 ? key_at(i0 - 1) : key_at(0);
  for (int i = i0; i < i1; ++i) {
    const auto k = key_at(i);
    const bool head = (i == 0 || k != prev);
    prev = k;
    if (!head) continue;
    const int r = row++;
    if (nbad_all && is_bad(k)) continue;
    double sx = 0.0, sy = 0.0, sz = 0.0;
    int cnt = 0;
    for (int j = i; j < n && key_at(j) == k; ++j) {
      const double *p = pts + 3 * (size_t)point_at(j);
      sx = __dadd_rn(sx, __ldg(p)); sy = __dadd_rn(sy, __ldg(p + 1)); sz = __dadd_rn(sz, __ldg(p + 2));
      ++cnt;
    }
    const double cc = (double)cnt;
    double *o = out + 3 * (size_t)r;
    o[0] = __ddiv_rn(sx, cc); o[1] = __ddiv_rn(sy, cc); o[2] = __ddiv_rn(sz, cc);
  }
  if (tid == 0) {
    *nvox = total - (nbad_all ? 1 : 0);
    if (nbad) *nbad = nbad_all;
  }
}

__global__ void __launch_bounds__(ST, 1)
voxel_small_kernel(const double *__restrict__ pts, int n, double voxel, unsigned flags, int smem_sort,
                   unsigned long long *keys_a, unsigned long long *keys_b, int *ord_a, int *ord_b,
                   double *__restrict__ out, int *__restrict__ nvox, int *__restrict__ nbad) {
  extern __shared__ __align__(16) unsigned char s_dyn[];   // smem_sort: key32[n] | ord16[2][n]
  __shared__ int s_tab[SW][256];
  __shared__ int s_base[256];
  __shared__ long long s_min[3][SW];
  __shared__ int s_max[3][SW];
  __shared__ int s_cnt[SW];
  __shared__ int s_bad;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // 1. per-axis minimum of the finite points
  long long m[3] = {0x7fffffffffffffffLL, 0x7fffffffffffffffLL, 0x7fffffffffffffffLL};
#pragma unroll 4
  for (int i = tid; i < n; i += ST) {
    const double x = __ldg(pts + 3 * (size_t)i), y = __ldg(pts + 3 * (size_t)i + 1), z = __ldg(pts + 3 * (size_t)i + 2);
    if (!(isfinite(x) && isfinite(y) && isfinite(z))) continue;
    m[0] = min(m[0], order_key(x)); m[1] = min(m[1], order_key(y)); m[2] = min(m[2], order_key(z));
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    m[a] = warp_min_ll(m[a]);
    if (lane == 0) s_min[a][warp] = m[a];
  }
  if (tid == 0) s_bad = 0;
  __syncthreads();
  const double half = (flags & B200_VOXEL_ORIGIN_AT_MIN) ? 0.0 : __dmul_rn(voxel, 0.5);
  double origin[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    long long v = s_min[a][lane];
    v = warp_min_ll(v);
    origin[a] = __dsub_rn(order_val(v), half);
  }

  // 2. keys (voxel_key_kernel's arithmetic) and the largest voxel coordinate per axis
  int qmax[3] = {0, 0, 0}, bad_here = 0;
#pragma unroll 2
  for (int i = tid; i < n; i += ST) {
    unsigned long long key = 0;
    bool bad = false;
    int qi[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const double q = floor(__ddiv_rn(__dsub_rn(__ldg(pts + 3 * (size_t)i + a), origin[a]), voxel));
      bad |= !(q >= 0.0) || q >= 2097152.0;
      qi[a] = bad ? 0 : (int)q;
      key = (key << 21) | (unsigned long long)qi[a];
    }
    if (bad) {
      key = ~0ull;
      ++bad_here;
    } else {
#pragma unroll
      for (int a = 0; a < 3; ++a) qmax[a] = max(qmax[a], qi[a]);
    }
    keys_a[i] = key;
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    qmax[a] = warp_max_i(qmax[a]);
    if (lane == 0) s_max[a][warp] = qmax[a];
  }
  bad_here = __reduce_add_sync(0xffffffffu, bad_here);
  if (lane == 0 && bad_here) atomicAdd(&s_bad, bad_here);
  __syncthreads();
#pragma unroll
  for (int a = 0; a < 3; ++a) qmax[a] = warp_max_i(s_max[a][lane]);
  const int nbad_all = s_bad;
  // the excluded points sort behind every voxel: x coordinate qmax[0] + 1 (22 bits at most) for the sort
  const int bz = bits_of((unsigned)qmax[2]), by = bits_of((unsigned)qmax[1]);
  const int bx = bits_of((unsigned)qmax[0] + (nbad_all ? 1u : 0u));

  if (smem_sort && bx + by + bz <= 32) {
    // ---- 3a. compact keys z | y << bz | x << (bz + by) in shared memory; the passes permute 16-bit point numbers
    unsigned *key32 = reinterpret_cast<unsigned *>(s_dyn);
    unsigned short *oin = reinterpret_cast<unsigned short *>(s_dyn + 4 * (size_t)n), *oout = oin + n;
    const unsigned bad32 = (unsigned)(qmax[0] + 1) << (bz + by);
#pragma unroll 4
    for (int i = tid; i < n; i += ST) {
      const unsigned long long k = keys_a[i];
      key32[i] = k == ~0ull ? bad32
                            : (unsigned)(k & 0x1fffffu) | ((unsigned)((k >> 21) & 0x1fffffu) << bz) |
                                  ((unsigned)(k >> 42) << (bz + by));
      oin[i] = (unsigned short)i;
    }
    __syncthreads();
    const int nbits = bx + by + bz;
#pragma unroll 1
    for (int b = 0; b < nbits; b += 8) {
      const unsigned short *ci = oin;
      unsigned short *co = oout;
      small_sort_pass(n, b, nbits - b < 8 ? nbits - b : 8, s_tab, s_base,
                      [&](int i) { return key32[ci[i]]; }, [&](int pos, int i) { co[pos] = ci[i]; });
      unsigned short *t = oin; oin = oout; oout = t;
    }
    const unsigned short *of = oin;
    small_finish(pts, n, nbad_all, s_cnt, out, nvox, nbad, [&](int i) { return key32[of[i]]; },
                 [&](int i) { return (int)of[i]; }, [&](unsigned k) { return k == bad32; });
    return;
  }

  // ---- 3b. the same on the 64-bit keys in the workspace: z (from bit 0), y (21), x (42), 8 bits at most per pass
  const unsigned long long bad_key = (unsigned long long)(qmax[0] + 1) << 42;
  for (int i = tid; i < n; i += ST) {
    if (nbad_all && keys_a[i] == ~0ull) keys_a[i] = bad_key;
    ord_a[i] = i;
  }
  __syncthreads();
  unsigned long long *kin = keys_a, *kout = keys_b;
  int *oin = ord_a, *oout = ord_b;
  const int nb[3] = {bz, by, bx};
#pragma unroll 1
  for (int a = 0; a < 3; ++a) {
#pragma unroll 1
    for (int b = 0; b < nb[a]; b += 8) {
      const unsigned long long *ki = kin;
      unsigned long long *ko = kout;
      const int *oi = oin;
      int *oo = oout;
      small_sort_pass(n, 21 * a + b, nb[a] - b < 8 ? nb[a] - b : 8, s_tab, s_base, [&](int i) { return ki[i]; },
                      [&](int pos, int i) { ko[pos] = ki[i]; oo[pos] = oi[i]; });
      unsigned long long *tk = kin; kin = kout; kout = tk;
      int *to = oin; oin = oout; oout = to;
    }
  }
  const unsigned long long *kf = kin;
  const int *of = oin;
  small_finish(pts, n, nbad_all, s_cnt, out, nvox, nbad, [&](int i) { return kf[i]; }, [&](int i) { return of[i]; },
               [&](unsigned long long k) { return k == bad_key; });
}

}  // namespace

static unsigned g_voxel_flags = 0u;
extern "C" void b200_voxel_set_options(unsigned flags) { g_voxel_flags = flags; }

extern "C" size_t b200_voxel_down_sample_workspace(int n) { return layout(n).total; }

extern "C" int b200_voxel_down_sample_checked(const double *points, int n, double voxel_size, double *out,
                                              int32_t *nvox, int32_t *nbad, void *workspace, size_t workspace_bytes,
                                              void *stream) {
  if (n < 0 || !nvox || !(voxel_size > 0.0)) return B200_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    B200_CUDA(cudaMemsetAsync(nvox, 0, sizeof(int32_t), st));
    if (nbad) B200_CUDA(cudaMemsetAsync(nbad, 0, sizeof(int32_t), st));
    return 0;
  }
  if (!points || !out || !workspace) return B200_EINVAL;
  const Layout L = layout(n);
  if (workspace_bytes < L.total) return B200_EWORKSPACE;
  char *ws = (char *)workspace;
  long long *mins = (long long *)(ws + L.mins);
  int *overflow = (int *)(ws + L.overflow);
  unsigned long long *keys_a = (unsigned long long *)(ws + L.keys_a), *keys_b = (unsigned long long *)(ws + L.keys_b);
  int *ord_a = (int *)(ws + L.ord_a), *ord_b = (int *)(ws + L.ord_b);
  int *head = (int *)(ws + L.head), *row = (int *)(ws + L.row);
  if (n <= SMALL_N && !(g_voxel_flags & B200_VOXEL_MULTI_CTA)) {
    const int smem_sort = n <= SMEM_N ? 1 : 0;
    static B200DeviceOnce attr_set;
    const int dev = b200_device();
    if (attr_set.need(dev)) {
      B200_CUDA(cudaFuncSetAttribute(voxel_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * SMEM_N));
      attr_set.mark(dev);
    }
    voxel_small_kernel<<<1, ST, smem_sort ? 8 * (size_t)n : 0, st>>>(points, n, voxel_size, g_voxel_flags, smem_sort, keys_a,
                                                                      keys_b, ord_a, ord_b, out, nvox, nbad);
    B200_LAUNCH_CHECK();
    return 0;
  }
  const int grid = ceil_div(n, VT);
  voxel_init_kernel<<<1, 1, 0, st>>>(mins, overflow, nvox);
  B200_LAUNCH_CHECK();
  voxel_bounds_kernel<<<min(grid, 4 * b200_sm_count()), VT, 0, st>>>(points, n, mins);
  B200_LAUNCH_CHECK();
  voxel_key_kernel<<<grid, VT, 0, st>>>(points, n, voxel_size, mins, keys_a, ord_a, overflow, g_voxel_flags);
  B200_LAUNCH_CHECK();
  size_t cub_bytes = L.total - L.cub;
  B200_CUDA(cub::DeviceRadixSort::SortPairs(ws + L.cub, cub_bytes, keys_a, keys_b, ord_a, ord_b, n, 0, 63, st));
  voxel_head_kernel<<<grid, VT, 0, st>>>(keys_b, n, head);
  B200_LAUNCH_CHECK();
  cub_bytes = L.total - L.cub;
  B200_CUDA(cub::DeviceScan::InclusiveSum(ws + L.cub, cub_bytes, head, row, n, st));
  voxel_row_fix_kernel<<<grid, VT, 0, st>>>(row, n);
  B200_LAUNCH_CHECK();
  voxel_mean_kernel<<<grid, VT, 0, st>>>(points, keys_b, ord_b, head, row, n, out, nvox, overflow, nbad);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_voxel_down_sample(const double *points, int n, double voxel_size, double *out,
                                      int32_t *nvox, void *workspace, size_t workspace_bytes, void *stream) {
  return b200_voxel_down_sample_checked(points, n, voxel_size, out, nvox, nullptr, workspace, workspace_bytes, stream);
}
