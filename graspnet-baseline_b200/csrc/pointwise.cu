// pointwise.cu -- 1x1 convolution over point columns with the bias (= folded eval-mode BatchNorm) and ReLU in the
// epilogue (sm_100a):   y[b, co, j] = act( bias[co] + sum_ci W[co, ci] * x[b, ci, j] )
//
// The layers between the geometry ops of GraspNet -- PointnetFPModule's SharedMLP (pointnet2_modules.py:409-414),
// ApproachNet (models/modules.py:52-57), OperationNet / ToleranceNet (:170-177, :215-222) -- are 13 such products per
// scene with 12..302 output channels over 512..4096 columns (1.9 GFLOP in all): far too small to fill a library
// SGEMM, and as torch calls each costs three launches (bias broadcast, GEMM, ReLU).  Here a layer is ONE launch of
// 64 x 32 output tiles (so that even the 256 x 1024 layers give 128 CTAs), exact fp32 FMAs in channel order within
// a K-slab, and x may be the channel concatenation of two tensors (FP modules: [interpolated | skip features],
// pointnet2_modules.py:403-404) so that torch.cat disappears as well.
//
// A CTA is 256 threads = two K-groups of 128: group g takes the 16-channel slabs g, g+2, ... of the reduction
// (its own shared-memory stages, its own named barrier), each thread a 4 x 4 register tile.  The layers are
// latency-bound (a CTA walks 8..16 slabs of a weight matrix that the L2 flush / another scene evicted), so the
// slabs move global -> shared with 16-byte cp.async through a 3-stage ring per group: two slabs are in flight
// while one is multiplied, one barrier per slab.  The two partial sums meet in shared memory in the epilogue.
// Summation order: slabs of a group in order, then group 0 + group 1 (fp32; the layers' parity bound is 1e-5
// relative, tests/test_gpu_graspnet.py).  Inputs that cannot be moved in 16-byte pieces (n or the weight row stride
// not a multiple of 4, unaligned pointers) take the register-prefetch kernel below.
#include "common.cuh"

namespace {

constexpr int BM = 64, BN = 32, BK = 16, GT = 128;       // tile: BM output channels x BN columns, K-slab, threads/group
constexpr int WS = BM + 4;                               // padded row of the transposed weight tile (16-byte aligned)

__device__ __forceinline__ void group_barrier(int g) {
  asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(GT) : "memory");
}

struct PwParams {
  const float *w, *bias, *x1, *x2;
  float *y;
  int c1, c2, n, cout, relu, ldw;
};

constexpr int STAGES = 3;
constexpr int WK = BK + 4;                               // padded row (k contiguous) of the cp.async weight tile
struct __align__(16) PwStage {
  float w[BM][WK];
  float x[BK][BN];
};

__device__ __forceinline__ void cp_async16(void *dst, const void *src, bool valid) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  const int bytes = valid ? 16 : 0;                      // 0 source bytes = zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(bytes) : "memory");
}

// fast path: n % 4 == 0, ldw % 4 == 0, 16-byte aligned pointers
__global__ void __launch_bounds__(2 * GT) pointwise_async_kernel(PwParams p) {
  __shared__ PwStage stage[2][STAGES];
  const int g = threadIdx.x / GT, t = threadIdx.x % GT;
  const int tx = t & 7, ty = t >> 3;                     // columns 4 tx.., output channels ty + 16 i
  const int co0 = blockIdx.y * BM, j0 = blockIdx.x * BN;
  const int cin = p.c1 + p.c2, n = p.n;
  const float *x1 = p.x1 + (size_t)blockIdx.z * p.c1 * n;
  const float *x2 = p.x2 ? p.x2 + (size_t)blockIdx.z * p.c2 * n : nullptr;
  const int nslab = (cin + BK - 1) / BK;
  const int ni = nslab > g ? (nslab - g + 1) / 2 : 0;    // slabs of this group: g, g+2, ...

  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;

  auto issue = [&](int i) {
    PwStage &st = stage[g][i % STAGES];
    const int k0 = (g + 2 * i) * BK;
#pragma unroll
    for (int h = 0; h < 2; ++h) {                        // weight tile: 64 rows x 4 pieces of 4 k
      const int e = t + GT * h, r = e >> 2, q = e & 3;
      const int co = co0 + r, k = k0 + 4 * q;
      const bool ok = co < p.cout && k < p.ldw;          // columns cin..ldw-1 of w are zero (host-padded)
      cp_async16(&st.w[r][4 * q], ok ? p.w + (size_t)co * p.ldw + k : p.w, ok);
    }
    const int kk = t >> 3, q = t & 7, ci = k0 + kk, j = j0 + 4 * q;
    const bool ok = ci < cin && j < n;
    const float *row = !ok ? x1 : (ci < p.c1 ? x1 + (size_t)ci * n : x2 + (size_t)(ci - p.c1) * n);
    cp_async16(&st.x[kk][4 * q], ok ? row + j : x1, ok);
  };

#pragma unroll
  for (int i = 0; i < STAGES - 1; ++i) {
    if (i < ni) issue(i);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (int i = 0; i < ni; ++i) {
    asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 2) : "memory");
    group_barrier(g);                                    // slab i has landed; everyone is done with slab i-1's stage
    if (i + STAGES - 1 < ni) issue(i + STAGES - 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    const PwStage &st = stage[g][i % STAGES];
#pragma unroll
    for (int kq = 0; kq < BK / 4; ++kq) {
      float4 a[4], b[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) a[r] = *reinterpret_cast<const float4 *>(&st.w[ty + 16 * r][4 * kq]);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) b[kk] = *reinterpret_cast<const float4 *>(&st.x[4 * kq + kk][4 * tx]);
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float av[4] = {a[r].x, a[r].y, a[r].z, a[r].w};
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          acc[r][0] = __fmaf_rn(av[kk], b[kk].x, acc[r][0]);
          acc[r][1] = __fmaf_rn(av[kk], b[kk].y, acc[r][1]);
          acc[r][2] = __fmaf_rn(av[kk], b[kk].z, acc[r][2]);
          acc[r][3] = __fmaf_rn(av[kk], b[kk].w, acc[r][3]);
        }
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();                                       // the stages are free: group 1 parks its partial sums there
  float(*part)[BN] = reinterpret_cast<float(*)[BN]>(&stage[0][0]);
  if (g == 1) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
      *reinterpret_cast<float4 *>(&part[ty + 16 * r][4 * tx]) = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
  }
  __syncthreads();
  if (g == 0) {
    float *y = p.y + (size_t)blockIdx.z * p.cout * n;
    const int j = j0 + 4 * tx;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int co = co0 + ty + 16 * r;
      if (co >= p.cout || j >= n) continue;
      const float4 o = *reinterpret_cast<const float4 *>(&part[ty + 16 * r][4 * tx]);
      const float bias = __ldg(p.bias + co);
      float v[4] = {acc[r][0] + o.x + bias, acc[r][1] + o.y + bias, acc[r][2] + o.z + bias, acc[r][3] + o.w + bias};
      if (p.relu) {
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = fmaxf(v[e], 0.f);
      }
      *reinterpret_cast<float4 *>(y + (size_t)co * n + j) = make_float4(v[0], v[1], v[2], v[3]);
    }
  }
}

template <bool VW, bool VX>
__global__ void __launch_bounds__(2 * GT) pointwise_kernel(PwParams p) {
  __shared__ __align__(16) float w_s[2][BK][WS];
  __shared__ __align__(16) float x_s[2][BK][BN];
  __shared__ __align__(16) float part[BM][BN];
  const int g = threadIdx.x / GT, t = threadIdx.x % GT;
  const int tx = t & 7, ty = t >> 3;                     // columns 4 tx.., output channels 4 ty..
  const int co0 = blockIdx.y * BM, j0 = blockIdx.x * BN;
  const int cin = p.c1 + p.c2, n = p.n;
  const float *x1 = p.x1 + (size_t)blockIdx.z * p.c1 * n;
  const float *x2 = p.x2 ? p.x2 + (size_t)blockIdx.z * p.c2 * n : nullptr;
  const int nslab = (cin + BK - 1) / BK;

  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;

  // this thread's share of a slab: two float4 of W (rows r, 4 consecutive k) and one float4 of x (row k, 4 columns)
  const int wr[2] = {t >> 2, (t + GT) >> 2}, wq = t & 3;
  const int xk = t >> 3, xq = t & 7;
  float4 wreg[2], xreg;

  auto fetch = [&](int s) {
    const int k0 = s * BK;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int co = co0 + wr[i], k = k0 + 4 * wq;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (co < p.cout) {
        const float *src = p.w + (size_t)co * p.ldw + k;
        if (VW) {
          if (k < cin) v = __ldg(reinterpret_cast<const float4 *>(src));
        } else {
          if (k + 0 < cin) v.x = __ldg(src + 0);
          if (k + 1 < cin) v.y = __ldg(src + 1);
          if (k + 2 < cin) v.z = __ldg(src + 2);
          if (k + 3 < cin) v.w = __ldg(src + 3);
        }
      }
      wreg[i] = v;
    }
    const int ci = k0 + xk, j = j0 + 4 * xq;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ci < cin) {
      const float *row = ci < p.c1 ? x1 + (size_t)ci * n : x2 + (size_t)(ci - p.c1) * n;
      if (VX) {
        if (j < n) v = __ldg(reinterpret_cast<const float4 *>(row + j));
      } else {
        if (j + 0 < n) v.x = __ldg(row + j + 0);
        if (j + 1 < n) v.y = __ldg(row + j + 1);
        if (j + 2 < n) v.z = __ldg(row + j + 2);
        if (j + 3 < n) v.w = __ldg(row + j + 3);
      }
    }
    xreg = v;
  };

  int s = g;
  if (s < nslab) fetch(s);
  for (; s < nslab; s += 2) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      w_s[g][4 * wq + 0][wr[i]] = wreg[i].x;
      w_s[g][4 * wq + 1][wr[i]] = wreg[i].y;
      w_s[g][4 * wq + 2][wr[i]] = wreg[i].z;
      w_s[g][4 * wq + 3][wr[i]] = wreg[i].w;
    }
    *reinterpret_cast<float4 *>(&x_s[g][xk][4 * xq]) = xreg;
    group_barrier(g);
    if (s + 2 < nslab) fetch(s + 2);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a = *reinterpret_cast<const float4 *>(&w_s[g][k][4 * ty]);
      const float4 b = *reinterpret_cast<const float4 *>(&x_s[g][k][4 * tx]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = __fmaf_rn(av[i], bv[j], acc[i][j]);
    }
    group_barrier(g);
  }

  if (g == 1) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      *reinterpret_cast<float4 *>(&part[4 * ty + i][4 * tx]) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  }
  __syncthreads();
  if (g == 0) {
    float *y = p.y + (size_t)blockIdx.z * p.cout * n;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int co = co0 + 4 * ty + i;
      if (co >= p.cout) continue;
      const float4 o = *reinterpret_cast<const float4 *>(&part[4 * ty + i][4 * tx]);
      const float bias = __ldg(p.bias + co);
      float r[4] = {acc[i][0] + o.x + bias, acc[i][1] + o.y + bias, acc[i][2] + o.z + bias, acc[i][3] + o.w + bias};
      if (p.relu) {
#pragma unroll
        for (int j = 0; j < 4; ++j) r[j] = fmaxf(r[j], 0.f);
      }
      const int j = j0 + 4 * tx;
      float *dst = y + (size_t)co * n + j;
      if (VX) {
        if (j < n) *reinterpret_cast<float4 *>(dst) = make_float4(r[0], r[1], r[2], r[3]);
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (j + e < n) dst[e] = r[e];
      }
    }
  }
}

}  // namespace

extern "C" int b200_pointwise_conv(const float *w, int ldw, const float *bias, const float *x1, int c1, const float *x2,
                                   int c2, int b, int n, int cout, int relu, float *y, void *stream) {
  if (b < 0 || n < 0 || cout < 0 || c1 < 1 || c2 < 0 || ldw < c1 + c2) return B200_EINVAL;
  if (b == 0 || n == 0 || cout == 0) return 0;
  if (!w || !bias || !x1 || !y || (c2 > 0 && !x2)) return B200_EINVAL;
  if (b > 65535 || (cout + BM - 1) / BM > 65535) return B200_ECAPACITY;
  PwParams p{w, bias, x1, c2 > 0 ? x2 : nullptr, y, c1, c2, n, cout, relu, ldw};
  const dim3 grid((n + BN - 1) / BN, (cout + BM - 1) / BM, b);
  auto aligned = [](const void *q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
  const bool vx = n % 4 == 0 && aligned(x1) && aligned(y) && (c2 == 0 || aligned(x2));
  const cudaStream_t st = (cudaStream_t)stream;
  // the cp.async path reads whole 16-byte pieces of a weight row up to ldw: the caller zero-pads rows to ldw % 4 == 0
  if (vx && ldw % 4 == 0 && aligned(w)) pointwise_async_kernel<<<grid, 2 * GT, 0, st>>>(p);
  else if (vx) pointwise_kernel<false, true><<<grid, 2 * GT, 0, st>>>(p);
  else pointwise_kernel<false, false><<<grid, 2 * GT, 0, st>>>(p);
  B200_LAUNCH_CHECK();
  return 0;
}
