// interp.cu -- three_nn / three_interpolate (+ backward) for sm_100a.
//
// Replaces three_nn_kernel, three_interpolate_kernel and three_interpolate_grad_kernel
// (reference pointnet2/_ext_src/src/interpolate_gpu.cu:14-64, 77-106, 121-148), all of
// which run one block per batch element.
//
// three_nn: grid over (unknown-tiles, batch); the known cloud is staged through
// shared memory in tiles; L lanes share one unknown point, each keeps the three
// best (d, index) pairs of the known points k = lane (mod L) and a butterfly of
// list merges combines them.  The legacy kernel is a strict-'<' cascade in index
// order, i.e. the three smallest under the order (d, index): that order is what
// the per-lane cascade and the merge use, so the result does not depend on L.
// (FP1/FP2 of the backbone have 512/1024 unknowns against 256/512 knowns: with a
// thread per unknown the op is one long serial scan on a handful of SMs.)
// The legacy running bests are doubles initialised to 1e40 and compared with the
// promoted fp32 distance: since every candidate is an fp32 value this is the same
// order as fp32 compares against +inf (1e40 stores back as +inf in fp32), so the
// cascade is done in fp32; a candidate that is +inf or NaN never enters a list.
// three_interpolate: thread per output column j, CT channels per CTA; idx and
// weight rows are loaded once and reused across channels; coalesced stores.
#include <math_constants.h>

#include "common.cuh"

namespace {

constexpr int NT3 = 128;
constexpr int KT = 1024;   // known points per tile
constexpr int CT = 8;

struct Best3 {
  float b1, b2, b3;
  int i1, i2, i3;
};

// (d, k) enters the list iff it precedes an entry under (distance, index); d must be < +inf
__device__ __forceinline__ void best3_insert(Best3 &t, float d, int k) {
  if (d < t.b1 || (d == t.b1 && k < t.i1)) {
    t.b3 = t.b2; t.i3 = t.i2; t.b2 = t.b1; t.i2 = t.i1; t.b1 = d; t.i1 = k;
  } else if (d < t.b2 || (d == t.b2 && k < t.i2)) {
    t.b3 = t.b2; t.i3 = t.i2; t.b2 = d; t.i2 = k;
  } else if (d < t.b3 || (d == t.b3 && k < t.i3)) {
    t.b3 = d; t.i3 = k;
  }
}

template <int L>
__global__ void __launch_bounds__(NT3)
three_nn_kernel(const float *__restrict__ unknown, const float *__restrict__ known, int n, int m,
                float *__restrict__ dist2, int32_t *__restrict__ idx) {
  __shared__ float tile[KT * 3];
  const int bi = blockIdx.y;
  const int j = blockIdx.x * (NT3 / L) + threadIdx.x / L;
  const int sub = threadIdx.x % L;
  const bool active = j < n;
  const float *kn = known + (size_t)bi * m * 3;
  float ux = 0.f, uy = 0.f, uz = 0.f;
  if (active) {
    const float *u = unknown + ((size_t)bi * n + j) * 3;
    ux = __ldg(u); uy = __ldg(u + 1); uz = __ldg(u + 2);
  }
  Best3 t{CUDART_INF_F, CUDART_INF_F, CUDART_INF_F, 0, 0, 0};
  for (int t0 = 0; t0 < m; t0 += KT) {
    const int tn = min(KT, m - t0);
    if (t0) __syncthreads();
    for (int i = threadIdx.x; i < tn * 3; i += NT3) tile[i] = __ldg(kn + (size_t)t0 * 3 + i);
    __syncthreads();
    if (active) {
#pragma unroll 4
      for (int k = sub; k < tn; k += L) {
        const float d = sq3_p2(__fsub_rn(ux, tile[3 * k]), __fsub_rn(uy, tile[3 * k + 1]),
                               __fsub_rn(uz, tile[3 * k + 2]));
        const int kk = t0 + k;
        // within a lane the indices ascend, so the strict compare alone is the (d, index) order
        if (d < t.b1) {
          t.b3 = t.b2; t.i3 = t.i2; t.b2 = t.b1; t.i2 = t.i1; t.b1 = d; t.i1 = kk;
        } else if (d < t.b2) {
          t.b3 = t.b2; t.i3 = t.i2; t.b2 = d; t.i2 = kk;
        } else if (d < t.b3) {
          t.b3 = d; t.i3 = kk;
        }
      }
    }
  }
  if (L > 1) {
    // every lane of the warp takes part (inactive ones carry empty lists)
#pragma unroll
    for (int o = 1; o < L; o <<= 1) {
      const float c1 = __shfl_xor_sync(0xffffffffu, t.b1, o), c2 = __shfl_xor_sync(0xffffffffu, t.b2, o),
                  c3 = __shfl_xor_sync(0xffffffffu, t.b3, o);
      const int k1 = __shfl_xor_sync(0xffffffffu, t.i1, o), k2 = __shfl_xor_sync(0xffffffffu, t.i2, o),
                k3 = __shfl_xor_sync(0xffffffffu, t.i3, o);
      if (c1 < CUDART_INF_F) best3_insert(t, c1, k1);
      if (c2 < CUDART_INF_F) best3_insert(t, c2, k2);
      if (c3 < CUDART_INF_F) best3_insert(t, c3, k3);
    }
  }
  if (active && sub == 0) {
    float *od = dist2 + ((size_t)bi * n + j) * 3;
    int32_t *oi = idx + ((size_t)bi * n + j) * 3;
    od[0] = t.b1; od[1] = t.b2; od[2] = t.b3;
    oi[0] = t.i1; oi[1] = t.i2; oi[2] = t.i3;
  }
}

__global__ void __launch_bounds__(NT3)
three_interpolate_kernel(const float *__restrict__ points, const int32_t *__restrict__ idx,
                         const float *__restrict__ weight, int c, int m, int n,
                         float *__restrict__ out) {
  const int bi = blockIdx.z;
  const int c0 = blockIdx.y * CT;
  const int cn = min(CT, c - c0);
  const int j = blockIdx.x * NT3 + threadIdx.x;
  if (j >= n) return;
  const int32_t *id = idx + ((size_t)bi * n + j) * 3;
  const float *w = weight + ((size_t)bi * n + j) * 3;
  const int a0 = __ldg(id), a1 = __ldg(id + 1), a2 = __ldg(id + 2);
  const float w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
  const float *src = points + ((size_t)bi * c + c0) * m;
  float *dst = out + ((size_t)bi * c + c0) * n + j;
#pragma unroll 4
  for (int l = 0; l < cn; ++l) {
    const float *row = src + (size_t)l * m;
    dst[(size_t)l * n] =
        __fmaf_rn(__ldg(row + a2), w2, __fmaf_rn(__ldg(row + a1), w1, __fmul_rn(__ldg(row + a0), w0)));
  }
}

__global__ void __launch_bounds__(NT3)
three_interpolate_grad_kernel(const float *__restrict__ grad_out, const int32_t *__restrict__ idx,
                              const float *__restrict__ weight, int c, int n, int m,
                              float *__restrict__ grad_points) {
  const int bi = blockIdx.z;
  const int c0 = blockIdx.y * CT;
  const int cn = min(CT, c - c0);
  const int j = blockIdx.x * NT3 + threadIdx.x;
  if (j >= n) return;
  const int32_t *id = idx + ((size_t)bi * n + j) * 3;
  const float *w = weight + ((size_t)bi * n + j) * 3;
  const int a0 = __ldg(id), a1 = __ldg(id + 1), a2 = __ldg(id + 2);
  const float w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
  const float *src = grad_out + ((size_t)bi * c + c0) * n + j;
  float *dst = grad_points + ((size_t)bi * c + c0) * m;
#pragma unroll 4
  for (int l = 0; l < cn; ++l) {
    const float g = __ldg(src + (size_t)l * n);
    float *row = dst + (size_t)l * m;
    atomicAdd(row + a0, __fmul_rn(g, w0));
    atomicAdd(row + a1, __fmul_rn(g, w1));
    atomicAdd(row + a2, __fmul_rn(g, w2));
  }
}

}  // namespace

extern "C" int b200_three_nn(const float *unknown, const float *known, int b, int n, int m,
                             float *dist2, int32_t *idx, void *stream) {
  if (b < 0 || n < 0 || m < 0) return B200_EINVAL;
  if (b == 0 || n == 0) return 0;
  if (!unknown || !dist2 || !idx || (m > 0 && !known)) return B200_EINVAL;
  if (b > 65535) return B200_ECAPACITY;
  // lanes per unknown point: as many as keep the grid within a few waves
  const long long work = (long long)b * n;
  if (work <= 16384 && m >= 64) {
    dim3 grid((unsigned)ceil_div(n, NT3 / 32), (unsigned)b);
    three_nn_kernel<32><<<grid, NT3, 0, (cudaStream_t)stream>>>(unknown, known, n, m, dist2, idx);
  } else if (work <= 131072 && m >= 16) {
    dim3 grid((unsigned)ceil_div(n, NT3 / 8), (unsigned)b);
    three_nn_kernel<8><<<grid, NT3, 0, (cudaStream_t)stream>>>(unknown, known, n, m, dist2, idx);
  } else {
    dim3 grid((unsigned)ceil_div(n, NT3), (unsigned)b);
    three_nn_kernel<1><<<grid, NT3, 0, (cudaStream_t)stream>>>(unknown, known, n, m, dist2, idx);
  }
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_three_interpolate(const float *points, const int32_t *idx, const float *weight,
                                      int b, int c, int m, int n, float *out, void *stream) {
  if (b < 0 || c < 0 || n < 0 || m < 0) return B200_EINVAL;
  if (b == 0 || c == 0 || n == 0) return 0;
  if (!points || !idx || !weight || !out) return B200_EINVAL;
  if (b > 65535 || ceil_div(c, CT) > 65535) return B200_ECAPACITY;
  dim3 grid((unsigned)ceil_div(n, NT3), (unsigned)ceil_div(c, CT), (unsigned)b);
  three_interpolate_kernel<<<grid, NT3, 0, (cudaStream_t)stream>>>(points, idx, weight, c, m, n, out);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_three_interpolate_grad(const float *grad_out, const int32_t *idx,
                                           const float *weight, int b, int c, int n, int m,
                                           float *grad_points, void *stream) {
  if (b < 0 || c < 0 || n < 0 || m < 0) return B200_EINVAL;
  if (b == 0 || c == 0 || m == 0) return 0;
  if (!grad_points) return B200_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaMemsetAsync(grad_points, 0, sizeof(float) * (size_t)b * c * m, st));
  if (n == 0) return 0;
  if (!grad_out || !idx || !weight) return B200_EINVAL;
  if (b > 65535 || ceil_div(c, CT) > 65535) return B200_ECAPACITY;
  dim3 grid((unsigned)ceil_div(n, NT3), (unsigned)ceil_div(c, CT), (unsigned)b);
  three_interpolate_grad_kernel<<<grid, NT3, 0, st>>>(grad_out, idx, weight, c, n, m, grad_points);
  B200_LAUNCH_CHECK();
  return 0;
}
