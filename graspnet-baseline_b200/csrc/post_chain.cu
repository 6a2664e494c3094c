// post_chain.cu -- the reference's post-processing sequence after pred_decode as ONE entry point and five graph nodes.
//
//   gg = gg[~mfcdetector.detect(gg, approach_dist, collision_thresh)]     (test.py:93-96, demo_pointcloud.py:228-232)
//   gg = gg.nms()                                                          (demo_pointcloud.py:249)
//
// Call by call (b200_collision_mask_device, b200_compact_rows, b200_grasp_nms_device, b200_gather_rows) the sequence is
// nine launches of 2-6 us each on one stream: at batch 1 their launch gaps, not their work, are what the scene's
// latency pays.  Here the same arithmetic (shared through post_common.cuh, results bit-identical -- tests/
// test_gpu_postprocess.py) runs as
//   1. memset of the six counters per grasp
//   2. collision_kernel        gripper bounds derived from the fp32 rows inside the kernel (no prep launch)
//   3. post_rank_kernel        counts -> collision flag -> the score order of the collision-free rows, written as row
//                              numbers of the ORIGINAL array: the compaction gg[~mask] never materialises
//   4. nms_mask_kernel         pair tests through order[]
//   5. nms_sweep_kernel        greedy sweep + gather of the kept rows
// and the caller's row count is passed along (nrows_out) so that no copy node is needed for it either.
#include <cmath>

#include "post_common.cuh"

namespace {

constexpr int G = 17;
constexpr int PT = 128;                  // rows ranked per CTA
constexpr int PMAX = B200_POST_CHAIN_MAX_M;

// grid ceil(m_max / PT).  Every CTA rebuilds the (small) list of collision-free scores in shared memory -- 2048 rows
// at most, a few flops per row -- and ranks its own PT entries of that list against all of it: no inter-CTA step.
__global__ void __launch_bounds__(PT)
post_rank_kernel(const float *__restrict__ rows, const int32_t *__restrict__ counts, const int32_t *__restrict__ m_dev,
                 int m, float fl, float fw, float ad, float vox3, double thresh, uint8_t *__restrict__ mask,
                 int32_t *__restrict__ order, int32_t *__restrict__ nfree, int32_t *__restrict__ nrows_out) {
  __shared__ __align__(16) float s_key[PMAX];          // scores of the free rows, in row order
  __shared__ int16_t s_row[PMAX];                      // their row numbers
  __shared__ int s_wsum[PT / 32];
  __shared__ int s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  m = min(m, __ldg(m_dev));
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int i0 = 0; i0 < m; i0 += PT) {
    const int i = i0 + tid;
    bool free_row = false;
    float key = 0.f;
    if (i < m) {
      const float *r = rows + (size_t)i * G;
      const bool hit = collision_row_collides(r, __ldg(counts + (size_t)i * 6), fl, fw, ad, vox3, thresh);
      if (blockIdx.x == 0) mask[i] = hit ? 1 : 0;
      free_row = !hit;
      key = nms_score_key(__ldg(r));
    }
    const unsigned bal = __ballot_sync(0xffffffffu, free_row);
    if (lane == 0) s_wsum[warp] = __popc(bal);
    __syncthreads();
    int pos = s_base;
    for (int w = 0; w < warp; ++w) pos += s_wsum[w];
    pos += __popc(bal & lanemask_lt());
    if (free_row) { s_key[pos] = key; s_row[pos] = (int16_t)i; }
    __syncthreads();
    if (tid == PT - 1) s_base = pos + (free_row ? 1 : 0);
    __syncthreads();
  }
  const int nf = s_base;
  if (blockIdx.x == 0 && tid == 0) {
    *nfree = nf;
    if (nrows_out) *nrows_out = m;
  }
  const int p = blockIdx.x * PT + tid;                 // this thread's entry of the free list
  if (blockIdx.x * PT >= nf) return;
  const float kp = p < nf ? s_key[p] : 0.f;
  // rank = #{q : key_q > key_p} + #{q < p : key_q == key_p}  (descending, ties by row number; keys are never NaN)
  int rank = 0;
  const int nf4 = nf & ~3;
  for (int q = 0; q < nf4; q += 4) {
    const float4 k = *reinterpret_cast<const float4 *>(s_key + q);
    rank += (k.x > kp) | ((k.x == kp) & (q < p));
    rank += (k.y > kp) | ((k.y == kp) & (q + 1 < p));
    rank += (k.z > kp) | ((k.z == kp) & (q + 2 < p));
    rank += (k.w > kp) | ((k.w == kp) & (q + 3 < p));
  }
  for (int q = nf4; q < nf; ++q) {
    const float k = s_key[q];
    rank += (k > kp) | ((k == kp) & (q < p));
  }
  if (p < nf) order[rank] = (int32_t)s_row[p];
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

extern "C" size_t b200_grasp_post_chain_workspace(int m_max) {
  if (m_max <= 0) return 256;
  const size_t m = (size_t)m_max, W = (m + 63) / 64;
  return align_up(sizeof(int32_t) * 6 * m, 256)        // collision counters
         + align_up(sizeof(int32_t) * m, 256)          // order
         + align_up(sizeof(int32_t) * m, 256)          // keep
         + 256                                         // nfree
         + align_up(8 * W * m, 256);                   // pair mask
}

extern "C" int b200_grasp_post_chain(const double *scene, const int32_t *n_dev, int n_max, const float *grasps,
                                     const int32_t *m_dev, int m_max, double voxel_size, double approach_dist,
                                     double collision_thresh, double translation_thresh, double cos_rotation_thresh,
                                     uint8_t *collision_mask, float *kept_rows, int32_t *nkept, int32_t *nrows_out,
                                     void *workspace, size_t workspace_bytes, void *stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (n_max < 0 || m_max < 0 || !nkept || !m_dev || !(voxel_size > 0.0)) return B200_EINVAL;
  if (m_max > PMAX) return B200_ECAPACITY;
  if (m_max == 0) {
    B200_CUDA(cudaMemsetAsync(nkept, 0, sizeof(int32_t), st));
    if (nrows_out) B200_CUDA(cudaMemsetAsync(nrows_out, 0, sizeof(int32_t), st));
    return 0;
  }
  if (!grasps || !collision_mask || !kept_rows || !workspace) return B200_EINVAL;
  if (workspace_bytes < b200_grasp_post_chain_workspace(m_max)) return B200_EWORKSPACE;
  const size_t m = (size_t)m_max;
  char *ws = static_cast<char *>(workspace);
  int32_t *counts = reinterpret_cast<int32_t *>(ws); ws += align_up(sizeof(int32_t) * 6 * m, 256);
  int32_t *order = reinterpret_cast<int32_t *>(ws);  ws += align_up(sizeof(int32_t) * m, 256);
  int32_t *keep = reinterpret_cast<int32_t *>(ws);   ws += align_up(sizeof(int32_t) * m, 256);
  int32_t *nfree = reinterpret_cast<int32_t *>(ws);  ws += 256;
  unsigned long long *pair_mask = reinterpret_cast<unsigned long long *>(ws);
  // the gripper constants of detect() (utils/collision_detector.py:29-30, :62) with the dtype rules
  // b200_collision_mask_device documents: Python floats enter the reference's fp32 expressions as fp32 scalars
  if (approach_dist < 0.01) approach_dist = 0.01;
  const float fl = (float)0.06, fw = (float)0.01, ad = (float)approach_dist;
  const float vox3 = (float)std::pow(voxel_size, 3.0);
  int rc = b200i_collision_counts_rows(scene, n_dev, n_max, grasps, m_dev, m_max, fl, fw, ad, counts, st);
  if (rc) return rc;
  post_rank_kernel<<<ceil_div(m_max, PT), PT, 0, st>>>(grasps, counts, m_dev, m_max, fl, fw, ad, vox3, collision_thresh,
                                                        collision_mask, order, nfree, nrows_out);
  B200_LAUNCH_CHECK();
  return b200i_nms_ranked(grasps, nfree, m_max, order, translation_thresh, cos_rotation_thresh, keep, nkept, kept_rows,
                          pair_mask, st);
}
