// rows.cu -- order-preserving row compaction and row gather with device-side counts (sm_100a).
//
// The two pieces of host glue between the reference's post-processing calls, kept on the device:
//   gg = gg[~collision_mask]         (test.py:96, demo_pointcloud.py:232)      -> b200_compact_rows
//   GraspGroup(array[kept_indices])  (what grasp_nms returns, in score order)  -> b200_gather_rows
// Row counts come from and go to device memory, so decode -> collision -> NMS never synchronises with the host.
#include "common.cuh"

namespace {

constexpr int CT = 1024;

__global__ void __launch_bounds__(CT)
compact_rows_kernel(const float *__restrict__ rows, int cols, const uint8_t *__restrict__ drop,
                    const int32_t *__restrict__ m_dev, int m, float *__restrict__ out, int32_t *__restrict__ nout) {
  __shared__ int warp_sum[CT / 32];
  __shared__ int s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (m_dev) m = min(m, __ldg(m_dev));
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int i0 = 0; i0 < m; i0 += CT) {
    const int i = i0 + tid;
    const bool keep = i < m && !drop[i];
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_sum[warp] = __popc(bal);
    __syncthreads();
    int before = s_base;
    for (int w = 0; w < warp; ++w) before += warp_sum[w];
    const int pos = before + __popc(bal & lanemask_lt());
    if (keep)
      for (int c = 0; c < cols; ++c) out[(size_t)pos * cols + c] = rows[(size_t)i * cols + c];
    __syncthreads();
    if (tid == CT - 1) s_base = pos + (keep ? 1 : 0);
    __syncthreads();
  }
  if (tid == 0) *nout = s_base;
}

__global__ void gather_rows_kernel(const float *__restrict__ rows, int cols, const int32_t *__restrict__ index,
                                   const int32_t *__restrict__ k_dev, int k, float *__restrict__ out) {
  if (k_dev) k = min(k, __ldg(k_dev));
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= k * cols) return;
  const int i = t / cols, c = t - i * cols;
  out[t] = rows[(size_t)__ldg(index + i) * cols + c];
}

}  // namespace

extern "C" int b200_compact_rows(const float *rows, int cols, const uint8_t *drop, const int32_t *m_dev, int m_max,
                                 float *out, int32_t *nout, void *stream) {
  if (m_max < 0 || cols < 1 || !nout) return B200_EINVAL;
  if (m_max > 0 && (!rows || !drop || !out)) return B200_EINVAL;
  compact_rows_kernel<<<1, CT, 0, (cudaStream_t)stream>>>(rows, cols, drop, m_dev, m_max, out, nout);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_gather_rows(const float *rows, int cols, const int32_t *index, const int32_t *k_dev, int k_max,
                                float *out, void *stream) {
  if (k_max < 0 || cols < 1) return B200_EINVAL;
  if (k_max == 0) return 0;
  if (!rows || !index || !out) return B200_EINVAL;
  const long long total = (long long)k_max * cols;
  if (total > 0x7fffffffLL) return B200_ECAPACITY;
  gather_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(rows, cols, index, k_dev, k_max, out);
  B200_LAUNCH_CHECK();
  return 0;
}
