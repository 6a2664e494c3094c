// decode.cu -- head outputs -> grasp rows (sm_100a).
//
// Replaces pred_decode (reference models/graspnet.py:77-134) and the rotation construction it calls
// (utils/loss_utils.py:68-96, batch_viewpoint_params_to_matrix): per seed point the best in-plane angle class per
// depth (argmax over A), the best depth (argmax over D of the gathered scores), width = clamp(1.2 w, 0, 0.1),
// score = score * tolerance / 0.05, rotation from the (negated) approach view and the angle, and the rows of the
// seeds whose objectness argmax is 1 compacted IN SEED ORDER into the (M,17) layout the collision detector and
// NMS read: [score, width, 0.02, depth, R(9 row-major), centre(3), -1].  The reference does this with ~40 small
// torch launches and a boolean-mask compaction per scene; here it is one launch for the whole batch, 128 seeds per
// CTA, and the row count stays on the device.
//
// Exactness: classes, depth, width, score and centre are selections of inputs or single-rounded fp32 expressions
// in the reference's order (bit-exact against the reference, tests/golden/decode_*.npz); the rotation goes through
// norms, sin/cos and a 3x3 product and is compared to 1e-6.  torch.argmax returns the FIRST maximum; the strict `>`
// scans below do the same.  The reference calls torch.cross without `dim`: with exactly three surviving seeds the
// cross product runs down the rows (first dimension of size 3) -- reproduced (the `total == 3` branch).
#include "common.cuh"

namespace {

constexpr int DT = 128;       // seeds per CTA
constexpr int MAXD = 8;

struct Frame { float ax[3], ay[3], s, c; };
struct Decoded { float score, width, depth, angle; };

// one seed: best in-plane angle class per depth (first maximum over A), best depth (first maximum over the gathered
// scores), then width / score / depth / angle of that (angle, depth) cell   [models/graspnet.py:92-118]
__device__ __forceinline__ Decoded decode_seed(const float *__restrict__ sc, const float *__restrict__ cl,
                                               const float *__restrict__ wd, const float *__restrict__ tl, int i, int ns,
                                               int na, int nd, int vec4) {
  int a_of[MAXD];
  if (vec4) {
    // D = 4: the four depths of (angle class a, seed i) are one aligned float4; the A loads are independent
    const float4 *c4 = reinterpret_cast<const float4 *>(cl) + i;
    float4 cb = __ldg(c4);
    a_of[0] = a_of[1] = a_of[2] = a_of[3] = 0;
#pragma unroll 4
    for (int a = 1; a < na; ++a) {
      const float4 c = __ldg(c4 + (size_t)a * ns);
      if (c.x > cb.x) { cb.x = c.x; a_of[0] = a; }
      if (c.y > cb.y) { cb.y = c.y; a_of[1] = a; }
      if (c.z > cb.z) { cb.z = c.z; a_of[2] = a; }
      if (c.w > cb.w) { cb.w = c.w; a_of[3] = a; }
    }
  } else {
    for (int d = 0; d < nd; ++d) {
      int a_best = 0;
      float c_best = __ldg(cl + ((size_t)0 * ns + i) * nd + d);
      for (int a = 1; a < na; ++a) {
        const float c = __ldg(cl + ((size_t)a * ns + i) * nd + d);
        if (c > c_best) { c_best = c; a_best = a; }
      }
      a_of[d] = a_best;
    }
  }
  float s_of[MAXD];
#pragma unroll
  for (int d = 0; d < MAXD; ++d)
    if (d < nd) s_of[d] = __ldg(sc + ((size_t)a_of[d] * ns + i) * nd + d);
  int d_best = 0;
#pragma unroll
  for (int d = 1; d < MAXD; ++d)
    if (d < nd && s_of[d] > s_of[d_best]) d_best = d;
  // (static indexing keeps a_of / s_of in registers)
  int a_best = a_of[0];
  float best = s_of[0];
#pragma unroll
  for (int d = 1; d < MAXD; ++d)
    if (d == d_best) { a_best = a_of[d]; best = s_of[d]; }
  const size_t at = ((size_t)a_best * ns + i) * nd + d_best;
  Decoded o;
  o.width = fminf(fmaxf(__fmul_rn(1.2f, __ldg(wd + at)), 0.f), 0.1f);
  o.score = __fdiv_rn(__fmul_rn(best, __ldg(tl + at)), 0.05f);
  o.depth = __fmul_rn((float)(d_best + 1), 0.01f);
  o.angle = __fmul_rn(__fdiv_rn((float)a_best, 12.f), 3.14159265358979323846f);
  return o;
}

// approach axis, in-plane y axis and sin / cos of the angle   [utils/loss_utils.py:68-96]
__device__ __forceinline__ Frame seed_frame(const float *__restrict__ vw, int i, float angle) {
  const float tx = -__ldg(vw + 3 * (size_t)i), ty = -__ldg(vw + 3 * (size_t)i + 1), tz = -__ldg(vw + 3 * (size_t)i + 2);
  float yx = -ty, yy = tx;                                      // axis_y = (-x1, x0, 0)
  if (sqrtf(yx * yx + yy * yy) == 0.f) yy = 1.f;
  const float nx = sqrtf(tx * tx + ty * ty + tz * tz), ny = sqrtf(yx * yx + yy * yy);
  Frame f;
  f.ax[0] = tx / nx; f.ax[1] = ty / nx; f.ax[2] = tz / nx;
  f.ay[0] = yx / ny; f.ay[1] = yy / ny; f.ay[2] = 0.f / ny;
  f.s = sinf(angle); f.c = cosf(angle);
  return f;
}

// grid (ceil(ns / DT), scenes).  The rows are compacted in seed order across the CTAs of a scene without any
// exchange: every CTA counts the kept seeds in front of its own range itself (2 floats per seed, L2-resident).
__global__ void __launch_bounds__(DT)
pred_decode_kernel(const float *__restrict__ objectness, const float *__restrict__ score,
                   const float *__restrict__ angle_cls, const float *__restrict__ width,
                   const float *__restrict__ tol, const float *__restrict__ xyz, const float *__restrict__ view,
                   int ns, int na, int nd, int vec4, float *__restrict__ rows, int32_t *__restrict__ nrows) {
  __shared__ int warp_sum[DT / 32], warp_before[DT / 32], warp_total[DT / 32];
  __shared__ int s_three[3], s_nthree;
  __shared__ Frame first3[3];
  const int scene = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int i_first = blockIdx.x * DT;
  const size_t head = (size_t)scene * na * ns * nd;
  const float *obj = objectness + (size_t)scene * 2 * ns;
  const float *sc = score + head, *cl = angle_cls + head, *wd = width + head, *tl = tol + head;
  const float *ctr = xyz + (size_t)scene * ns * 3, *vw = view + (size_t)scene * ns * 3;
  float *out = rows + (size_t)scene * ns * 17;
  auto kept = [&](int i) { return __ldg(obj + ns + i) > __ldg(obj + i); };   // argmax over 2 == 1 (first maximum wins a tie)
  // kept seeds in front of this CTA's range, and in the whole scene
  int n_before = 0, n_total = 0;
  for (int i = tid; i < ns; i += DT) {
    const int k = kept(i) ? 1 : 0;
    n_total += k;
    n_before += i < i_first ? k : 0;
  }
  n_before = __reduce_add_sync(0xffffffffu, n_before);
  n_total = __reduce_add_sync(0xffffffffu, n_total);
  if (tid == 0) s_nthree = 0;
  if (lane == 0) { warp_before[warp] = n_before; warp_total[warp] = n_total; }
  __syncthreads();
  int before = 0, total = 0;
#pragma unroll
  for (int w = 0; w < DT / 32; ++w) { before += warp_before[w]; total += warp_total[w]; }
  if (blockIdx.x == 0 && tid == 0) nrows[scene] = total;
  if (total == 3) {
    // torch.cross(axis_x, axis_y) without `dim` on (3,3) inputs is the cross product of the COLUMNS: each of the
    // three rows needs the frames of the other two.  Every CTA decodes the three seeds again (rare, and cheap).
    for (int i = tid; i < ns; i += DT)
      if (kept(i)) s_three[atomicAdd(&s_nthree, 1)] = i;
    __syncthreads();
    if (tid == 0) {                                                     // seed order
      if (s_three[0] > s_three[1]) { const int t = s_three[0]; s_three[0] = s_three[1]; s_three[1] = t; }
      if (s_three[1] > s_three[2]) { const int t = s_three[1]; s_three[1] = s_three[2]; s_three[2] = t; }
      if (s_three[0] > s_three[1]) { const int t = s_three[0]; s_three[0] = s_three[1]; s_three[1] = t; }
    }
    __syncthreads();
    if (tid < 3) first3[tid] = seed_frame(vw, s_three[tid], decode_seed(sc, cl, wd, tl, s_three[tid], ns, na, nd, vec4).angle);
    __syncthreads();
  }
  const int i = i_first + tid;
  const bool keep = i < ns && kept(i);
  // order-preserving compaction inside the CTA
  const unsigned bal = __ballot_sync(0xffffffffu, keep);
  if (lane == 0) warp_sum[warp] = __popc(bal);
  __syncthreads();
  int pos = before + __popc(bal & lanemask_lt());
  for (int w = 0; w < warp; ++w) pos += warp_sum[w];
  if (!keep) return;
  const Decoded g = decode_seed(sc, cl, wd, tl, i, ns, na, nd, vec4);
  const Frame f = seed_frame(vw, i, g.angle);
  const float ax0 = f.ax[0], ax1 = f.ax[1], ax2 = f.ax[2], ay0 = f.ay[0], ay1 = f.ay[1], ay2 = f.ay[2], s = f.s, c = f.c;
  float az0 = ax1 * ay2 - ax2 * ay1, az1 = ax2 * ay0 - ax0 * ay2, az2 = ax0 * ay1 - ax1 * ay0;
  if (total == 3) {                                                     // z[pos] of the column-wise cross products
    float z[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const float a0 = first3[0].ax[j], a1 = first3[1].ax[j], a2 = first3[2].ax[j];
      const float b0 = first3[0].ay[j], b1 = first3[1].ay[j], b2 = first3[2].ay[j];
      const float zz[3] = {a1 * b2 - a2 * b1, a2 * b0 - a0 * b2, a0 * b1 - a1 * b0};
      z[j] = pos == 0 ? zz[0] : (pos == 1 ? zz[1] : zz[2]);
    }
    az0 = z[0]; az1 = z[1]; az2 = z[2];
  }
  float *r = out + (size_t)pos * 17;
  r[0] = g.score; r[1] = g.width; r[2] = 0.02f; r[3] = g.depth;
  r[4] = ax0; r[5] = ay0 * c + az0 * s; r[6] = az0 * c - ay0 * s;
  r[7] = ax1; r[8] = ay1 * c + az1 * s; r[9] = az1 * c - ay1 * s;
  r[10] = ax2; r[11] = ay2 * c + az2 * s; r[12] = az2 * c - ay2 * s;
  r[13] = __ldg(ctr + 3 * (size_t)i); r[14] = __ldg(ctr + 3 * (size_t)i + 1); r[15] = __ldg(ctr + 3 * (size_t)i + 2);
  r[16] = -1.f;
}

// ApproachNet's tail (reference models/modules.py:60-73): per seed the best of the V template views (torch.max over
// the view scores = first maximum), its unit direction and the approach frame built from it with in-plane angle 0.
// The V directions (generate_grasp_views, utils/loss_utils.py:41-66) and their frames
// (batch_viewpoint_params_to_matrix(-view, 0), :68-96) depend on V only: the caller computes both tables once and
// this kernel looks them up -- the reference rebuilds the lattice in a Python loop and expands it to (B,S,V,3)
// on every forward.  features is conv3's output [b, 2+V, s]; lanes = consecutive seeds, so every view row is read
// coalesced.
// A CTA is 8 seeds x 32 view slices (128 CTAs per 1024 seeds, ~10 independent loads per thread): slice g scans the
// views g, g+32, ... (strict `>`, so the first maximum of its own ascending sub-sequence), then the 32 candidates of a
// seed meet in shared memory (larger value, then lower index).
// NaN is treated as the largest value and the first NaN wins (torch.max propagates it).
__device__ __forceinline__ bool view_better(float v, int k, float bv, int bk) {
  const bool vn = v != v, bn = bv != bv;
  if (vn != bn) return vn;
  if (!vn && v != bv) return v > bv;
  return k < bk;
}

constexpr int VSEEDS = 8;      // seeds per CTA (a 32-byte sector of every view row)
constexpr int VSL = 32;        // view slices per CTA

__global__ void __launch_bounds__(VSEEDS * VSL)
view_select_kernel(const float *__restrict__ features, const float *__restrict__ views,
                   const float *__restrict__ frames, int s, int v, long long *__restrict__ top_inds,
                   float *__restrict__ top_score, float *__restrict__ vp_xyz, float *__restrict__ vp_rot) {
  __shared__ float s_val[VSL][VSEEDS];
  __shared__ int s_idx[VSL][VSEEDS];
  const int scene = blockIdx.y, sx = threadIdx.x % VSEEDS, g = threadIdx.x / VSEEDS;
  const int i = blockIdx.x * VSEEDS + sx;
  const float *f = features + ((size_t)scene * (2 + v) + 2) * s + min(i, s - 1);
  int best = g < v ? g : 0;
  float bv = __ldg(f + (size_t)best * s);
#pragma unroll 10
  for (int k = g + VSL; k < v; k += VSL) {
    const float x = __ldg(f + (size_t)k * s);
    if (view_better(x, k, bv, best)) { bv = x; best = k; }
  }
  s_val[g][sx] = bv;
  s_idx[g][sx] = best;
  __syncthreads();
  if (g == 0 && i < s) {
#pragma unroll
    for (int q = 1; q < VSL; ++q)
      if (view_better(s_val[q][sx], s_idx[q][sx], bv, best)) { bv = s_val[q][sx]; best = s_idx[q][sx]; }
    s_idx[0][sx] = best;
    const size_t o = (size_t)scene * s + i;
    top_inds[o] = best;
    top_score[o] = bv;
  }
  __syncthreads();
  // the 12 table floats of each seed
  for (int e = threadIdx.x; e < VSEEDS * 12; e += VSEEDS * VSL) {
    const int q = e / 12, c = e % 12, ii = blockIdx.x * VSEEDS + q;
    if (ii >= s) break;
    const int k = s_idx[0][q];
    const size_t o = (size_t)scene * s + ii;
    if (c < 3) vp_xyz[o * 3 + c] = __ldg(views + k * 3 + c);
    else vp_rot[o * 9 + (c - 3)] = __ldg(frames + k * 9 + (c - 3));
  }
}

// PointnetFPModule's interpolation weights (reference pointnet2/pointnet2_modules.py:394-397, after the sqrt of
// pointnet2_utils.py:140): dist = sqrt(dist2); r = 1 / (dist + 1e-8); weight = r / (r0 + r1 + r2).
__global__ void __launch_bounds__(256)
three_nn_weights_kernel(const float *__restrict__ dist2, int n, float *__restrict__ dist, float *__restrict__ weight) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float d[3], r[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    d[k] = __fsqrt_rn(__ldg(dist2 + 3 * (size_t)i + k));
    r[k] = __fdiv_rn(1.0f, __fadd_rn(d[k], 1e-8f));
  }
  const float norm = __fadd_rn(__fadd_rn(r[0], r[1]), r[2]);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if (dist) dist[3 * (size_t)i + k] = d[k];
    weight[3 * (size_t)i + k] = __fdiv_rn(r[k], norm);
  }
}

}  // namespace

extern "C" int b200_pred_decode(const float *objectness, const float *score, const float *angle_cls,
                                const float *width, const float *tolerance, const float *xyz, const float *view_xyz,
                                int b, int ns, int num_angle, int num_depth, float *rows, int32_t *nrows,
                                void *stream) {
  if (b < 0 || ns < 0 || num_angle < 1 || num_depth < 1) return B200_EINVAL;
  if (num_depth > MAXD) return B200_ECAPACITY;
  if (b == 0) return 0;
  if (!nrows) return B200_EINVAL;
  if (ns > 0 && (!objectness || !score || !angle_cls || !width || !tolerance || !xyz || !view_xyz || !rows))
    return B200_EINVAL;
  const int vec4 = num_depth == 4 && (reinterpret_cast<uintptr_t>(angle_cls) & 15u) == 0;
  if (b > 65535) return B200_ECAPACITY;
  if (ns == 0) {
    B200_CUDA(cudaMemsetAsync(nrows, 0, sizeof(int32_t) * (size_t)b, (cudaStream_t)stream));
    return 0;
  }
  pred_decode_kernel<<<dim3((unsigned)ceil_div(ns, DT), (unsigned)b), DT, 0, (cudaStream_t)stream>>>(objectness, score, angle_cls, width, tolerance, xyz, view_xyz,
                                                        ns, num_angle, num_depth, vec4, rows, nrows);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_view_select(const float *features, const float *views, const float *frames, int b, int s, int v,
                                long long *top_inds, float *top_score, float *vp_xyz, float *vp_rot, void *stream) {
  if (b < 0 || s < 0 || v < 1) return B200_EINVAL;
  if (b == 0 || s == 0) return 0;
  if (!features || !views || !frames || !top_inds || !top_score || !vp_xyz || !vp_rot) return B200_EINVAL;
  if (b > 65535) return B200_ECAPACITY;
  view_select_kernel<<<dim3((s + VSEEDS - 1) / VSEEDS, b), VSEEDS * VSL, 0, (cudaStream_t)stream>>>(features, views, frames, s, v,
                                                                                 top_inds, top_score, vp_xyz, vp_rot);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_three_nn_weights(const float *dist2, int n, float *dist, float *weight, void *stream) {
  if (n < 0) return B200_EINVAL;
  if (n == 0) return 0;
  if (!dist2 || !weight) return B200_EINVAL;
  three_nn_weights_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(dist2, n, dist, weight);
  B200_LAUNCH_CHECK();
  return 0;
}
