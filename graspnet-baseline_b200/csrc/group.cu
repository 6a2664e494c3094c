// group.cu -- gather / grouping and their backward scatters for sm_100a.
//
// Replaces gather_points_kernel / gather_points_grad_kernel
// (reference pointnet2/_ext_src/src/sampling_gpu.cu:13-52) and group_points_kernel /
// group_points_grad_kernel (group_points_gpu.cu:13-69).
//
// out[b,c,j,k] = points[b,c,idx[b,j,k]] is the same operation as the gather
// out[b,c,t] = points[b,c,idx[b,t]] over the flattened t = j*ns + k, so one
// kernel serves both.  The legacy group kernel runs one block per batch element
// and has each thread write ns consecutive floats (a warp's stores are ns-strided)
// while re-reading idx for every channel.  Here: grid over (t-tiles, channel-tiles,
// batch); a thread owns 4 consecutive t, loads their indices ONCE as an int4 and
// reuses them for CT channels; every store is a coalesced 128-bit store; the
// random reads hit a C-row of `points` that stays L1/L2 resident.
// Backward: same mapping, 128-bit loads of grad_out, red.global.add.f32 scatter
// into a callee-zeroed grad_points (order differs from the legacy atomics =>
// 1e-5-relative, not bit-exact, like the reference itself run twice).
#include "common.cuh"

namespace {

constexpr int GT = 256;   // threads per CTA
constexpr int CT = 8;     // channels per CTA

template <bool VEC>
__global__ void __launch_bounds__(GT)
gather_kernel(const float *__restrict__ points, const int32_t *__restrict__ idx, int c, int n,
              long long mt, float *__restrict__ out) {
  const int bi = blockIdx.z;
  const int c0 = blockIdx.y * CT;
  const int cn = min(CT, c - c0);
  const float *src = points + ((size_t)bi * c + c0) * n;
  float *dst = out + ((size_t)bi * c + c0) * mt;
  const int32_t *id = idx + (size_t)bi * mt;
  if (VEC) {
    const long long t = ((long long)blockIdx.x * GT + threadIdx.x) * 4;
    if (t >= mt) return;
    const int4 i4 = __ldg(reinterpret_cast<const int4 *>(id + t));
#pragma unroll 4
    for (int l = 0; l < cn; ++l) {
      const float *row = src + (size_t)l * n;
      float4 v;
      v.x = __ldg(row + i4.x); v.y = __ldg(row + i4.y); v.z = __ldg(row + i4.z); v.w = __ldg(row + i4.w);
      __stcs(reinterpret_cast<float4 *>(dst + (size_t)l * mt + t), v);
    }
  } else {
    const long long t = (long long)blockIdx.x * GT + threadIdx.x;
    if (t >= mt) return;
    const int i = __ldg(id + t);
    for (int l = 0; l < cn; ++l) dst[(size_t)l * mt + t] = __ldg(src + (size_t)l * n + i);
  }
}

template <bool VEC>
__global__ void __launch_bounds__(GT)
scatter_add_kernel(const float *__restrict__ grad_out, const int32_t *__restrict__ idx, int c, int n,
                   long long mt, float *__restrict__ grad_points) {
  const int bi = blockIdx.z;
  const int c0 = blockIdx.y * CT;
  const int cn = min(CT, c - c0);
  const float *src = grad_out + ((size_t)bi * c + c0) * mt;
  float *dst = grad_points + ((size_t)bi * c + c0) * n;
  const int32_t *id = idx + (size_t)bi * mt;
  if (VEC) {
    const long long t = ((long long)blockIdx.x * GT + threadIdx.x) * 4;
    if (t >= mt) return;
    const int4 i4 = __ldg(reinterpret_cast<const int4 *>(id + t));
#pragma unroll 4
    for (int l = 0; l < cn; ++l) {
      const float4 g = __ldcs(reinterpret_cast<const float4 *>(src + (size_t)l * mt + t));
      float *row = dst + (size_t)l * n;
      atomicAdd(row + i4.x, g.x); atomicAdd(row + i4.y, g.y);
      atomicAdd(row + i4.z, g.z); atomicAdd(row + i4.w, g.w);
    }
  } else {
    const long long t = (long long)blockIdx.x * GT + threadIdx.x;
    if (t >= mt) return;
    const int i = __ldg(id + t);
    for (int l = 0; l < cn; ++l) atomicAdd(dst + (size_t)l * n + i, src[(size_t)l * mt + t]);
  }
}

bool aligned16(const void *p) { return ((uintptr_t)p & 15u) == 0; }

int gather_launch(const float *points, const int32_t *idx, int b, int c, int n, long long mt,
                  float *out, cudaStream_t st) {
  if (b == 0 || c == 0 || mt == 0) return 0;
  if (!points || !idx || !out) return B200_EINVAL;
  if (b > 65535 || ceil_div(c, CT) > 65535) return B200_ECAPACITY;
  const bool vec = (mt % 4 == 0) && aligned16(idx) && aligned16(out);
  const long long per = vec ? (long long)GT * 4 : GT;
  const long long gx = ceil_div_ll(mt, per);
  if (gx > 0x7fffffffLL) return B200_ECAPACITY;
  dim3 grid((unsigned)gx, (unsigned)ceil_div(c, CT), (unsigned)b);
  if (vec) gather_kernel<true><<<grid, GT, 0, st>>>(points, idx, c, n, mt, out);
  else gather_kernel<false><<<grid, GT, 0, st>>>(points, idx, c, n, mt, out);
  B200_LAUNCH_CHECK();
  return 0;
}

int scatter_launch(const float *grad_out, const int32_t *idx, int b, int c, int n, long long mt,
                   float *grad_points, cudaStream_t st) {
  if (b == 0 || c == 0 || n == 0) return 0;
  if (!grad_points) return B200_EINVAL;
  B200_CUDA(cudaMemsetAsync(grad_points, 0, sizeof(float) * (size_t)b * c * n, st));
  if (mt == 0) return 0;
  if (!grad_out || !idx) return B200_EINVAL;
  if (b > 65535 || ceil_div(c, CT) > 65535) return B200_ECAPACITY;
  const bool vec = (mt % 4 == 0) && aligned16(idx) && aligned16(grad_out);
  const long long per = vec ? (long long)GT * 4 : GT;
  const long long gx = ceil_div_ll(mt, per);
  if (gx > 0x7fffffffLL) return B200_ECAPACITY;
  dim3 grid((unsigned)gx, (unsigned)ceil_div(c, CT), (unsigned)b);
  if (vec) scatter_add_kernel<true><<<grid, GT, 0, st>>>(grad_out, idx, c, n, mt, grad_points);
  else scatter_add_kernel<false><<<grid, GT, 0, st>>>(grad_out, idx, c, n, mt, grad_points);
  B200_LAUNCH_CHECK();
  return 0;
}

// new_xyz = gather_operation(xyz^T, inds)^T with both layouts written by one launch: the reference's SA module
// runs transpose+contiguous, gather, transpose+contiguous (pointnet2_modules.py:230-240) -- three launches on the
// serial FPS chain per level.  thread = one sampled point.
__global__ void __launch_bounds__(GT)
gather_xyz_kernel(const float *__restrict__ xyz, const int32_t *__restrict__ idx, int n, int m,
                  float *__restrict__ out, float *__restrict__ out_t) {
  const int bi = blockIdx.y;
  const int j = blockIdx.x * GT + threadIdx.x;
  if (j >= m) return;
  const float *p = xyz + ((size_t)bi * n + __ldg(idx + (size_t)bi * m + j)) * 3;
  const float x = __ldg(p), y = __ldg(p + 1), z = __ldg(p + 2);
  float *o = out + ((size_t)bi * m + j) * 3;
  o[0] = x; o[1] = y; o[2] = z;
  if (out_t) {
    float *t = out_t + (size_t)bi * 3 * m + j;
    t[0] = x; t[m] = y; t[2 * (size_t)m] = z;
  }
}

}  // namespace

extern "C" int b200_gather_xyz(const float *xyz, const int32_t *idx, int b, int n, int m, float *out, float *out_t,
                               void *stream) {
  if (b < 0 || n < 0 || m < 0) return B200_EINVAL;
  if (b == 0 || m == 0) return 0;
  if (!xyz || !idx || !out || n == 0) return B200_EINVAL;
  if (b > 65535) return B200_ECAPACITY;
  gather_xyz_kernel<<<dim3((unsigned)ceil_div(m, GT), (unsigned)b), GT, 0, (cudaStream_t)stream>>>(xyz, idx, n, m, out,
                                                                                                  out_t);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_gather_points(const float *points, const int32_t *idx, int b, int c, int n, int m,
                                  float *out, void *stream) {
  if (b < 0 || c < 0 || n < 0 || m < 0) return B200_EINVAL;
  return gather_launch(points, idx, b, c, n, m, out, (cudaStream_t)stream);
}

extern "C" int b200_gather_points_grad(const float *grad_out, const int32_t *idx, int b, int c, int n,
                                       int m, float *grad_points, void *stream) {
  if (b < 0 || c < 0 || n < 0 || m < 0) return B200_EINVAL;
  return scatter_launch(grad_out, idx, b, c, n, m, grad_points, (cudaStream_t)stream);
}

extern "C" int b200_group_points(const float *points, const int32_t *idx, int b, int c, int n, int m,
                                 int ns, float *out, void *stream) {
  if (b < 0 || c < 0 || n < 0 || m < 0 || ns < 0) return B200_EINVAL;
  return gather_launch(points, idx, b, c, n, (long long)m * ns, out, (cudaStream_t)stream);
}

extern "C" int b200_group_points_grad(const float *grad_out, const int32_t *idx, int b, int c, int n,
                                      int m, int ns, float *grad_points, void *stream) {
  if (b < 0 || c < 0 || n < 0 || m < 0 || ns < 0) return B200_EINVAL;
  return scatter_launch(grad_out, idx, b, c, n, (long long)m * ns, grad_points, (cudaStream_t)stream);
}
