// labels.cu -- training-path label generation fused with the nearest-grasp-point search.
//
// Reference: utils/label_generation.py.  process_grasp_labels (:18-122) finds, for every seed point, the nearest
// annotated grasp point of the scene (knn, k = 1, :84-86) and then index_selects, for that point, the whole
// (V, A, D) = (300, 12, 4) block of labels, offsets (x3) and tolerances into (B, Ns, V, A, D[, 3]) batch tensors
// (:89-94: 59 + 177 + 59 MB per scene at Ns = 1024, after the per-object tensors were already copied once by the
// index_select over views :64-66 and the torch.cat :76-81), rewrites the labels as log(u_max / label) under a mask
// (:113-117) and takes the maximum over (A, D) as the view label (:118).  match_grasp_view_and_label (:125-149) then
// keeps ONE view of those tensors per seed.  The loss (models/loss.py) only ever reads the view labels (B, Ns, V) and
// the matched slices -- so here the big tensors are never built:
//   b200_label_view_stats   one pass over the gathered label / offset rows, straight from the per-object arrays
//                           through a pointer table: per (seed, view) the smallest label that passes the mask, and
//                           the global maximum u_max of all gathered labels;
//   b200_label_view_scores  view label = log(u_max / min) (x -> log(u_max / x) is monotone, so the maximum over
//                           (A, D) of the transformed labels is the transform of the smallest masked label);
//   b200_label_top_gather   the matched slices: transformed labels, offsets, tolerances (A, D[, 3]) and the rotated
//                           template frame (3, 3) of each seed's top view.
// Traffic per scene: the 236 MB of label + offset rows are read once (HBM-bound: 36 us at 6.5 TB/s) and ~1.5 MB are
// written, against ~1.2 GB of reads + writes for the index_select chain.
#include <math.h>

#include "common.cuh"

namespace {

struct ObjTable {                       // per object of a scene (device array, built by the host wrapper)
  const float *labels;                  // (Np, V, A, D)
  const float *offsets;                 // (Np, V, A, D, 3)
  const float *tolerance;               // (Np, V, A, D)
  const int32_t *view_inds;             // (V): template view v of the scene frame <- view view_inds[v] of the object frame
  const float *rot;                     // the object's rotation (pose[:3, :3]), 9 floats row-major, device memory
};

__device__ __forceinline__ unsigned ordered_key(float f) {            // monotone float -> unsigned
  const unsigned b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_to_float(unsigned k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

constexpr int LV_THREADS = 256;

// grid (Ns, B); one CTA per seed, one warp per view in turn; lanes over the A * D entries of the row
__global__ void __launch_bounds__(LV_THREADS)
label_view_stats_kernel(const ObjTable *__restrict__ objs, const int32_t *__restrict__ obj_first,
                        const int32_t *__restrict__ point_obj, const int32_t *__restrict__ point_local,
                        const int32_t *__restrict__ point_first, const long long *__restrict__ nn, int ns, int V,
                        int AD, float max_width, float *__restrict__ lmin, unsigned *__restrict__ umax_key) {
  const int s = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pt = point_first[b] + (int)nn[(size_t)b * ns + s];        // merged-point number within the batch tables
  const ObjTable &o = objs[obj_first[b] + point_obj[pt]];
  const size_t prow = (size_t)point_local[pt] * V;
  unsigned best = 0u;                                                 // ordered key of -NaN..: below every real label
  for (int v = warp; v < V; v += LV_THREADS / 32) {
    const size_t row = (prow + (size_t)__ldg(o.view_inds + v)) * AD;
    const float *lab = o.labels + row;
    const float *off = o.offsets + row * 3;
    float mn = INFINITY;
    for (int e = lane; e < AD; e += 32) {
      const float l = __ldg(lab + e), w = __ldg(off + 3 * e + 2);
      best = max(best, ordered_key(l));
      if (l > 0.f && w <= max_width) mn = fminf(mn, l);
    }
    mn = __uint_as_float(__reduce_min_sync(0xffffffffu, __float_as_uint(mn)));    // mn > 0 or +inf: bit order = value order
    if (lane == 0) lmin[((size_t)b * ns + s) * V + v] = mn;
  }
  best = __reduce_max_sync(0xffffffffu, best);
  __shared__ unsigned sbest[LV_THREADS / 32];
  if (lane == 0) sbest[warp] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned m = sbest[0];
    for (int w = 1; w < LV_THREADS / 32; ++w) m = max(m, sbest[w]);
    atomicMax(umax_key, m);
  }
}

__global__ void label_view_scores_kernel(const float *__restrict__ lmin, const unsigned *__restrict__ umax_key,
                                         long long n, float *__restrict__ scores, float *__restrict__ umax_out) {
  const float umax = key_to_float(*umax_key);
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0 && umax_out) *umax_out = umax;
  if (i >= n) return;
  const float m = lmin[i];
  scores[i] = isinf(m) ? 0.f : logf(__fdiv_rn(umax, m));
}

// grid (ceil(Ns / 4), B), 128 threads: one warp per seed; the seed's top view only
__global__ void __launch_bounds__(128)
label_top_gather_kernel(const ObjTable *__restrict__ objs, const int32_t *__restrict__ obj_first,
                        const int32_t *__restrict__ point_obj, const int32_t *__restrict__ point_local,
                        const int32_t *__restrict__ point_first, const long long *__restrict__ nn,
                        const long long *__restrict__ top_view, const float *__restrict__ view_rot_tmpl,
                        const unsigned *__restrict__ umax_key, int ns, int V, int AD, float max_width,
                        float *__restrict__ out_label, float *__restrict__ out_offset, float *__restrict__ out_tol,
                        float *__restrict__ out_rot) {
  const int s = blockIdx.x * 4 + (threadIdx.x >> 5), b = blockIdx.y, lane = threadIdx.x & 31;
  if (s >= ns) return;
  const size_t seed = (size_t)b * ns + s;
  const int pt = point_first[b] + (int)nn[seed];
  const ObjTable &o = objs[obj_first[b] + point_obj[pt]];
  const int src = __ldg(o.view_inds + (int)top_view[seed]);
  const size_t row = ((size_t)point_local[pt] * V + src) * AD;
  const float umax = key_to_float(*umax_key);
  for (int e = lane; e < AD; e += 32) {
    const float l = __ldg(o.labels + row + e);
    const float o0 = __ldg(o.offsets + (row + e) * 3), o1 = __ldg(o.offsets + (row + e) * 3 + 1),
                o2 = __ldg(o.offsets + (row + e) * 3 + 2);
    out_label[seed * AD + e] = (l > 0.f && o2 <= max_width) ? logf(__fdiv_rn(umax, l)) : 0.f;
    out_offset[(seed * AD + e) * 3] = o0;
    out_offset[(seed * AD + e) * 3 + 1] = o1;
    out_offset[(seed * AD + e) * 3 + 2] = o2;
    out_tol[seed * AD + e] = __ldg(o.tolerance + row + e);
  }
  if (lane < 9) {                                                     // pose rotation times the template frame of view src
    const int r = lane / 3, c = lane % 3;
    const float *t = view_rot_tmpl + (size_t)src * 9;
    out_rot[seed * 9 + lane] = __fmaf_rn(__ldg(o.rot + 3 * r + 2), __ldg(t + 6 + c),
                                         __fmaf_rn(__ldg(o.rot + 3 * r + 1), __ldg(t + 3 + c),
                                                   __fmul_rn(__ldg(o.rot + 3 * r), __ldg(t + c))));
  }
}

}  // namespace

extern "C" size_t b200_label_object_table_bytes(int n_objects) { return sizeof(ObjTable) * (size_t)(n_objects > 0 ? n_objects : 0); }

// HOST helper: fills entry i of a host-side object table (the caller copies the table to the device)
extern "C" int b200_label_object_table_set(void *host_table, int i, const float *labels, const float *offsets,
                                           const float *tolerance, const int32_t *view_inds, const float *rot9) {
  if (!host_table || i < 0 || !labels || !offsets || !tolerance || !view_inds || !rot9) return B200_EINVAL;
  ObjTable &o = static_cast<ObjTable *>(host_table)[i];
  o.labels = labels; o.offsets = offsets; o.tolerance = tolerance; o.view_inds = view_inds; o.rot = rot9;
  return 0;
}

extern "C" int b200_label_view_stats(const void *objs, const int32_t *obj_first, const int32_t *point_obj,
                                     const int32_t *point_local, const int32_t *point_first, const long long *nn, int b,
                                     int ns, int v, int ad, float max_width, float *lmin, uint32_t *umax_key,
                                     void *stream) {
  if (b < 0 || ns < 0 || v < 1 || ad < 1) return B200_EINVAL;
  if (!umax_key) return B200_EINVAL;
  B200_CUDA(cudaMemsetAsync(umax_key, 0, sizeof(uint32_t), (cudaStream_t)stream));
  if (b == 0 || ns == 0) return 0;
  if (!objs || !obj_first || !point_obj || !point_local || !point_first || !nn || !lmin) return B200_EINVAL;
  if (b > 65535) return B200_ECAPACITY;
  label_view_stats_kernel<<<dim3((unsigned)ns, (unsigned)b), LV_THREADS, 0, (cudaStream_t)stream>>>(
      static_cast<const ObjTable *>(objs), obj_first, point_obj, point_local, point_first, nn, ns, v, ad, max_width, lmin,
      umax_key);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_label_view_scores(const float *lmin, const uint32_t *umax_key, long long n, float *scores,
                                      float *umax_out, void *stream) {
  if (n < 0 || !umax_key) return B200_EINVAL;
  if (n > 0 && (!lmin || !scores)) return B200_EINVAL;
  const long long blocks = n > 0 ? ceil_div_ll(n, 256) : 1;
  if (blocks > 0x7fffffffLL) return B200_ECAPACITY;
  label_view_scores_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(lmin, umax_key, n, scores, umax_out);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_label_top_gather(const void *objs, const int32_t *obj_first, const int32_t *point_obj,
                                     const int32_t *point_local, const int32_t *point_first, const long long *nn,
                                     const long long *top_view, const float *view_rot_tmpl, const uint32_t *umax_key,
                                     int b, int ns, int v, int ad, float max_width, float *out_label,
                                     float *out_offset, float *out_tol, float *out_rot, void *stream) {
  if (b < 0 || ns < 0 || v < 1 || ad < 1) return B200_EINVAL;
  if (b == 0 || ns == 0) return 0;
  if (!objs || !obj_first || !point_obj || !point_local || !point_first || !nn || !top_view || !view_rot_tmpl ||
      !umax_key || !out_label || !out_offset || !out_tol || !out_rot)
    return B200_EINVAL;
  if (b > 65535) return B200_ECAPACITY;
  label_top_gather_kernel<<<dim3((unsigned)ceil_div(ns, 4), (unsigned)b), 128, 0, (cudaStream_t)stream>>>(
      static_cast<const ObjTable *>(objs), obj_first, point_obj, point_local, point_first, nn, top_view, view_rot_tmpl,
      umax_key, ns, v, ad, max_width, out_label, out_offset, out_tol, out_rot);
  B200_LAUNCH_CHECK();
  return 0;
}
