// query_pred.cuh -- the per-(centre, point) tests shared by query.cu (linear scan) and query_grid.cu (grid walk).
// The arithmetic (operation order, fused multiply-adds) is the legacy kernels' as compiled by nvcc:
// ball_query_gpu.cu:33-37, cylinder_query_gpu.cu:36-46.
#pragma once
#include "common.cuh"

namespace {

constexpr int MAXH = 8;

struct HList { float v[MAXH]; };

struct BallPred {
  float cx, cy, cz, r2;
  __device__ __forceinline__ void init(const float *c, const float *, float radius, float) {
    cx = c[0]; cy = c[1]; cz = c[2];
    r2 = __fmul_rn(radius, radius);
  }
  // returns whether the point passes every test that does not depend on the list;
  // `h` receives the value compared against each list's upper bound.
  __device__ __forceinline__ bool test(float x, float y, float z, float &h) const {
    h = 0.f;
    return sq3_p2(__fsub_rn(cx, x), __fsub_rn(cy, y), __fsub_rn(cz, z)) < r2;
  }
};

struct CylPred {
  float cx, cy, cz, r2, hmin;
  float r0, r1, r2_, r3, r4, r5, r6, r7, r8;
  __device__ __forceinline__ void init(const float *c, const float *rot, float radius, float hmin_) {
    cx = c[0]; cy = c[1]; cz = c[2];
    r2 = __fmul_rn(radius, radius);
    hmin = hmin_;
    r0 = rot[0]; r1 = rot[1]; r2_ = rot[2]; r3 = rot[3]; r4 = rot[4];
    r5 = rot[5]; r6 = rot[6]; r7 = rot[7]; r8 = rot[8];
  }
  __device__ __forceinline__ bool test(float px, float py, float pz, float &h) const {
    const float x = __fsub_rn(px, cx), y = __fsub_rn(py, cy), z = __fsub_rn(pz, cz);
    const float xr = __fmaf_rn(r6, z, __fmaf_rn(r0, x, __fmul_rn(r3, y)));
    const float yr = __fmaf_rn(r7, z, __fmaf_rn(r1, x, __fmul_rn(r4, y)));
    const float zr = __fmaf_rn(r8, z, __fmaf_rn(r2_, x, __fmul_rn(r5, y)));
    const float d2 = __fmaf_rn(yr, yr, __fmul_rn(zr, zr));
    h = xr;
    return d2 < r2 && xr > hmin;
  }
};

}  // namespace
