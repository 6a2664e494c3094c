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
