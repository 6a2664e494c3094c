// post_common.cuh -- pieces of the post-processing chain shared by its kernels (collision.cu, nms.cu, post_chain.cu),
// so that the fused chain and the call-by-call chain evaluate the SAME expressions (they are compared bit for bit).
#pragma once
#include <math.h>

#include "common.cuh"

// detect()'s per-grasp gripper bounds from an fp32 grasp row, with numpy's dtype rules (the rows are fp32, Python
// floats are weak scalars: everything in fp32, utils/collision_detector.py:69-93): T (3), R (9), thr (10) as fp64
__device__ __forceinline__ void collision_row_params(const float *__restrict__ r, float fl, float fw, float ad,
                                                     double (&t)[3], double (&rot)[9], double (&h)[10]) {
  t[0] = (double)r[13]; t[1] = (double)r[14]; t[2] = (double)r[15];
#pragma unroll
  for (int e = 0; e < 9; ++e) rot[e] = (double)r[4 + e];
  const float w = r[1], hh = r[2], d = r[3];
  const float h2 = __fdiv_rn(hh, 2.f), w2 = __fdiv_rn(w, 2.f);
  const float dl = __fsub_rn(d, fl), dlw = __fsub_rn(dl, fw);
  h[0] = (double)(-h2);                       // -h/2
  h[1] = (double)h2;                          //  h/2
  h[2] = (double)dl;                          //  d - finger_length
  h[3] = (double)d;
  h[4] = (double)(-__fadd_rn(w2, fw));        // -(w/2 + finger_width)
  h[5] = (double)(-w2);
  h[6] = (double)__fadd_rn(w2, fw);
  h[7] = (double)w2;
  h[8] = (double)dlw;                         //  d - finger_length - finger_width
  h[9] = (double)__fsub_rn(dlw, ad);          //  ... - approach_dist
}

// counts -> collision flag: global_iou = count / (volume + 1e-6) > collision_thresh (collision_detector.py:103-112)
__device__ __forceinline__ bool collision_row_collides(const float *__restrict__ r, int count, float fl, float fw, float ad,
                                                       float vox3, double thresh) {
  const float w = r[1], h = r[2];
  const float fw2 = __fmul_rn(2.f, fw);                                                   // 2*finger_width (weak scalar)
  const float lrv = __fdiv_rn(__fmul_rn(__fmul_rn(h, fl), fw), vox3);
  const float hw = __fmul_rn(h, __fadd_rn(w, fw2));
  const float bot = __fdiv_rn(__fmul_rn(hw, fw), vox3);
  const float shf = __fdiv_rn(__fmul_rn(hw, ad), vox3);
  const float vol = __fadd_rn(__fadd_rn(__fmul_rn(lrv, 2.f), bot), shf);
  const double iou = __ddiv_rn((double)count, (double)__fadd_rn(vol, 1e-6f));
  return iou > thresh;
}

// NMS score order: descending under a TOTAL order (NaN counts as -inf, ties by row number)
__device__ __forceinline__ float nms_score_key(float s) { return isnan(s) ? -INFINITY : s; }

// internal launchers shared across translation units (not part of the C ABI)
int b200i_collision_counts_rows(const double *scene, const int32_t *n_dev, int n_max, const float *rows,
                                const int32_t *m_dev, int m_max, float fl, float fw, float ad, int32_t *counts,
                                cudaStream_t st);
// mask + sweep of grasp NMS for rows whose score order is already known (order[rank] = row); the sweep also gathers
// the kept rows into out_rows (if not NULL).  mask_ws: m_max x ceil(m_max / 64) words.
int b200i_nms_ranked(const float *grasps, const int32_t *m_dev, int m_max, const int32_t *order, double translation_thresh,
                     double cos_rotation_thresh, int32_t *keep, int32_t *nkeep, float *out_rows,
                     unsigned long long *mask_ws, cudaStream_t st);
