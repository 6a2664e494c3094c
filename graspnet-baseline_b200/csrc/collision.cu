// collision.cu -- model-free gripper/scene collision counting for sm_100a.
//
// Integer core of ModelFreeCollisionDetector.detect (reference
// utils/collision_detector.py:69-100,119), which on the CPU materialises an
// (M,N,3) fp64 tensor, a batched matmul and ~20 (M,N) boolean temporaries.
//
// Mapping: grid over (grasp-tiles x point-tiles).  A thread owns ONE grasp: its
// translation, rotation and ten half-space bounds stay in fp64 registers; the
// CTA's slice of the scene is staged through shared memory and read back as
// warp-wide broadcasts; the eight masks of the reference are evaluated in fp64
// and accumulated into six per-thread integer counters, added to counts[M,6]
// with one integer atomic per counter per CTA slice.  Integer counts make the
// result independent of summation order, so masks/IoUs derived from them on the
// host are bit-identical to the reference's.
//
// Rounding: numpy evaluates (p - T) . R for an (N,3)x(3,3) fp64 product with the
// BLAS micro-kernel's FMA chain  t_c = fma(z, R[2][c], fma(y, R[1][c], x * R[0][c]))
// (pinned against the reference itself in tests/golden/collision_*.npz).
#include <cmath>

#include "post_common.cuh"

namespace {

constexpr int CTH = 128;    // grasps per CTA (one per thread)
constexpr int CPT = 256;    // scene points per CTA slice (6 KB of fp64 xyz): enough CTAs to fill every SM several times
constexpr int CU = 4;       // points per inner step: CU independent fp64 chains per thread

__global__ void __launch_bounds__(CTH)
collision_kernel(const double *__restrict__ scene, int n, const int32_t *__restrict__ n_dev,
                 const double *__restrict__ T, const double *__restrict__ R,
                 const double *__restrict__ thr, int st, int sr, int sh, int m,
                 const int32_t *__restrict__ m_dev, int32_t *__restrict__ counts,
                 const float *__restrict__ rows = nullptr, float fl = 0.f, float fw = 0.f, float ad = 0.f) {
  __shared__ double pts[CPT * 3];
  if (n_dev) n = min(n, __ldg(n_dev));               // scene length produced on the device
  if (m_dev) m = min(m, __ldg(m_dev));               // grasp count produced on the device (decode / compaction)
  const int g = blockIdx.x * CTH + threadIdx.x;
  // the scene is split EVENLY over the point tiles of the grid (sized for n_max / the SM count by the host): when the
  // device-side n is a fraction of n_max every CTA still gets a slice, instead of a few CTAs of CPT points with one
  // warp per scheduler waiting on fp64 latency while the rest of the grid exits
  const int per = min(CPT, ((n + (int)gridDim.y - 1) / (int)gridDim.y + CU - 1) / CU * CU);
  const int p0 = blockIdx.y * per;
  if (p0 >= n) return;
  const int pn = min(per, n - p0);
  // the tail of a short slice is NaN: every comparison below is false for it
  for (int i = threadIdx.x; i < per * 3; i += CTH)
    pts[i] = i < pn * 3 ? __ldg(scene + (size_t)p0 * 3 + i) : __longlong_as_double(0x7ff8000000000000LL);
  __syncthreads();
  if (g >= m) return;
  double tq[3], rq[9], hq[10];
  const double *t = tq, *r = rq, *h = hq;
  if (rows) {                                        // fp32 grasp rows: the bounds are derived here (no prep launch)
    collision_row_params(rows + (size_t)g * 17, fl, fw, ad, tq, rq, hq);
  } else {
    t = T + (size_t)g * st; r = R + (size_t)g * sr; h = thr + (size_t)g * sh;              // row strides
  }
  const double t0 = t[0], t1 = t[1], t2 = t[2];
  const double r0 = r[0], r1 = r[1], r2 = r[2], r3 = r[3], r4 = r[4], r5 = r[5], r6 = r[6],
               r7 = r[7], r8 = r[8];
  const double h0 = h[0], h1 = h[1], h2 = h[2], h3 = h[3], h4 = h[4], h5 = h[5], h6 = h[6],
               h7 = h[7], h8 = h[8], h9 = h[9];
  int cg_ = 0, cl = 0, cr = 0, cb = 0, cs = 0, ci = 0;
  const int steps = (pn + CU - 1) / CU;
  for (int k4 = 0; k4 < steps; ++k4) {
    double x[CU], y[CU], z[CU];
    bool m1[CU], any = false;
#pragma unroll
    for (int u = 0; u < CU; ++u) {
      const int k = k4 * CU + u;
      x[u] = __dsub_rn(pts[3 * k], t0); y[u] = __dsub_rn(pts[3 * k + 1], t1); z[u] = __dsub_rn(pts[3 * k + 2], t2);
      const double tz = __fma_rn(z[u], r8, __fma_rn(y[u], r5, __dmul_rn(x[u], r2)));
      m1[u] = (tz > h0) & (tz < h1);
      any = any | m1[u];
    }
    if (!__any_sync(__activemask(), any)) continue;    // every mask below needs the height slab
#pragma unroll
    for (int u = 0; u < CU; ++u) {
      if (!m1[u]) continue;
      const double tx = __fma_rn(z[u], r6, __fma_rn(y[u], r3, __dmul_rn(x[u], r0)));
      const double ty = __fma_rn(z[u], r7, __fma_rn(y[u], r4, __dmul_rn(x[u], r1)));
      const bool m2 = (tx > h2) & (tx < h3);
      const bool m3 = ty > h4;
      const bool m4 = ty < h5;
      const bool m5 = ty < h6;
      const bool m6 = ty > h7;
      const bool m7 = (tx <= h2) & (tx > h8);
      const bool m8 = (tx <= h8) & (tx > h9);
      const bool left = m2 & m3 & m4;
      const bool right = m2 & m5 & m6;
      const bool bottom = m3 & m5 & m7;
      const bool shifting = m3 & m5 & m8;
      cg_ += (left | right | bottom | shifting);
      cl += left; cr += right; cb += bottom; cs += shifting;
      ci += (m2 & (!m4) & (!m6));
    }
  }
  int32_t *o = counts + (size_t)g * 6;
  if (cg_) atomicAdd(o + 0, cg_);
  if (cl) atomicAdd(o + 1, cl);
  if (cr) atomicAdd(o + 2, cr);
  if (cb) atomicAdd(o + 3, cb);
  if (cs) atomicAdd(o + 4, cs);
  if (ci) atomicAdd(o + 5, ci);
}

}  // namespace

static int collision_launch(const double *scene, int n, const int32_t *n_dev, const double *T,
                            const double *R, const double *thr, int st, int sr, int sh, int m,
                            const int32_t *m_dev, int32_t *counts, void *stream);

// point tiles of the grid: slices of CPT points, or shorter ones (>= 32 points) while the grid is under four CTAs
// per SM -- the kernel shares the scene evenly among whatever number of tiles it is given
static int collision_point_tiles(int m, int n) {
  const int gx = ceil_div(m, CTH);
  int gy = ceil_div(n, CPT);
  const int want = ceil_div(4 * b200_sm_count(), gx);
  if (gy < want) gy = want < ceil_div(n, 32) ? want : ceil_div(n, 32);
  return gy < 1 ? 1 : gy;
}

extern "C" int b200_collision_counts(const double *scene, int n, const double *T, const double *R,
                                     const double *thr, int m, int32_t *counts, void *stream) {
  return collision_launch(scene, n, nullptr, T, R, thr, 3, 9, 10, m, nullptr, counts, stream);
}

extern "C" int b200_collision_counts_dn(const double *scene, const int32_t *n_dev, int n_max,
                                        const double *T, const double *R, const double *thr, int m,
                                        int32_t *counts, void *stream) {
  if (!n_dev) return B200_EINVAL;
  return collision_launch(scene, n_max, n_dev, T, R, thr, 3, 9, 10, m, nullptr, counts, stream);
}

extern "C" int b200_collision_counts_packed(const double *scene, const int32_t *n_dev, int n_max,
                                            const double *packed, int m, int32_t *counts, void *stream) {
  if (!packed && m > 0) return B200_EINVAL;
  return collision_launch(scene, n_max, n_dev, packed, packed + 3, packed + 12, 22, 22, 22, m, nullptr, counts, stream);
}

static int collision_launch(const double *scene, int n, const int32_t *n_dev, const double *T,
                            const double *R, const double *thr, int st, int sr, int sh, int m,
                            const int32_t *m_dev, int32_t *counts, void *stream) {
  if (n < 0 || m < 0) return B200_EINVAL;
  if (m == 0) return 0;
  if (!counts) return B200_EINVAL;
  cudaStream_t cs = (cudaStream_t)stream;
  B200_CUDA(cudaMemsetAsync(counts, 0, sizeof(int32_t) * 6 * (size_t)m, cs));
  if (n == 0) return 0;
  if (!scene || !T || !R || !thr) return B200_EINVAL;
  if (ceil_div(n, CPT) > 65535) return B200_ECAPACITY;
  dim3 grid((unsigned)ceil_div(m, CTH), (unsigned)collision_point_tiles(m, n));
  collision_kernel<<<grid, CTH, 0, cs>>>(scene, n, n_dev, T, R, thr, st, sr, sh, m, m_dev, counts);
  B200_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Device-resident detect(): grasp rows that are already on the device (pred_decode's output) -> collision mask,
// with the reference's own arithmetic and dtype rules for everything around the point-in-gripper counts
// (utils/collision_detector.py:69-93 thresholds, :103-112 volumes / IoU / mask).  The rows are fp32 (what
// `.cpu().numpy()` of the network output gives the reference, test.py:88-89), so numpy evaluates the per-grasp
// expressions in fp32 (Python floats are weak scalars) and only the IoU division in fp64 (int64 / float32).
namespace {

constexpr double FINGER_WIDTH = 0.01;    // collision_detector.py:29
constexpr double FINGER_LENGTH = 0.06;   // collision_detector.py:30

// one thread per grasp: T | R | thr of detect() as a packed fp64 row (the layout collision_kernel reads)
__global__ void collision_prep_kernel(const float *__restrict__ grasps, const int32_t *__restrict__ m_dev, int m,
                                      float fl, float fw, float ad, double *__restrict__ packed) {
  if (m_dev) m = min(m, __ldg(m_dev));
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= m) return;
  double t[3], r[9], h[10];
  collision_row_params(grasps + (size_t)g * 17, fl, fw, ad, t, r, h);
  double *o = packed + (size_t)g * 22;
#pragma unroll
  for (int e = 0; e < 3; ++e) o[e] = t[e];
#pragma unroll
  for (int e = 0; e < 9; ++e) o[3 + e] = r[e];
#pragma unroll
  for (int e = 0; e < 10; ++e) o[12 + e] = h[e];
}

// counts -> mask: global_iou = count / (volume + 1e-6) > collision_thresh
__global__ void collision_finalize_kernel(const float *__restrict__ grasps, const int32_t *__restrict__ counts,
                                          const int32_t *__restrict__ m_dev, int m, float fl, float fw, float ad,
                                          float vox3, double thresh, uint8_t *__restrict__ mask) {
  if (m_dev) m = min(m, __ldg(m_dev));
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= m) return;
  mask[g] = collision_row_collides(grasps + (size_t)g * 17, counts[(size_t)g * 6], fl, fw, ad, vox3, thresh) ? 1 : 0;
}

}  // namespace

extern "C" size_t b200_collision_mask_workspace(int m_max) {
  const size_t m = (size_t)(m_max > 0 ? m_max : 1);
  return (m * 22 * sizeof(double) + 255) / 256 * 256 + (m * 6 * sizeof(int32_t) + 255) / 256 * 256;
}

extern "C" int b200_collision_mask_device(const double *scene, const int32_t *n_dev, int n_max, const float *grasps,
                                          const int32_t *m_dev, int m_max, double voxel_size, double approach_dist,
                                          double collision_thresh, uint8_t *mask, void *workspace,
                                          size_t workspace_bytes, void *stream) {
  if (m_max < 0 || n_max < 0 || !(voxel_size > 0.0)) return B200_EINVAL;
  if (m_max == 0) return 0;
  if (!grasps || !mask || !workspace) return B200_EINVAL;
  if (workspace_bytes < b200_collision_mask_workspace(m_max)) return B200_EWORKSPACE;
  cudaStream_t cs = (cudaStream_t)stream;
  double *packed = (double *)workspace;
  int32_t *counts = (int32_t *)((char *)workspace + ((size_t)m_max * 22 * sizeof(double) + 255) / 256 * 256);
  if (approach_dist < FINGER_WIDTH) approach_dist = FINGER_WIDTH;            // collision_detector.py:62
  // Python floats enter the reference's fp32 expressions as fp32 scalars
  const float fl = (float)FINGER_LENGTH, fw = (float)FINGER_WIDTH, ad = (float)approach_dist;
  const float vox3 = (float)std::pow(voxel_size, 3.0);                      // voxel_size ** 3 (a Python float)
  const int blocks = ceil_div(m_max, 128);
  collision_prep_kernel<<<blocks, 128, 0, cs>>>(grasps, m_dev, m_max, fl, fw, ad, packed);
  B200_LAUNCH_CHECK();
  const int rc = collision_launch(scene, n_max, n_dev, packed, packed + 3, packed + 12, 22, 22, 22, m_max, m_dev,
                                  counts, stream);
  if (rc) return rc;
  collision_finalize_kernel<<<blocks, 128, 0, cs>>>(grasps, counts, m_dev, m_max, fl, fw, ad, vox3, collision_thresh,
                                                    mask);
  B200_LAUNCH_CHECK();
  return 0;
}

// the point-in-gripper counts of fp32 grasp rows, their bounds derived inside the kernel (post_chain.cu)
int b200i_collision_counts_rows(const double *scene, const int32_t *n_dev, int n_max, const float *rows,
                                const int32_t *m_dev, int m_max, float fl, float fw, float ad, int32_t *counts,
                                cudaStream_t st) {
  if (n_max < 0 || m_max < 0) return B200_EINVAL;
  if (m_max == 0) return 0;
  if (!counts || !rows) return B200_EINVAL;
  B200_CUDA(cudaMemsetAsync(counts, 0, sizeof(int32_t) * 6 * (size_t)m_max, st));
  if (n_max == 0) return 0;
  if (!scene) return B200_EINVAL;
  if (ceil_div(n_max, CPT) > 65535) return B200_ECAPACITY;
  dim3 grid((unsigned)ceil_div(m_max, CTH), (unsigned)collision_point_tiles(m_max, n_max));
  collision_kernel<<<grid, CTH, 0, st>>>(scene, n_max, n_dev, nullptr, nullptr, nullptr, 0, 0, 0, m_max, m_dev, counts, rows,
                                         fl, fw, ad);
  B200_LAUNCH_CHECK();
  return 0;
}
