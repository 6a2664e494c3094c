// query_grid.cu -- ball / cylinder query through a uniform grid (sm_100a).
//
// Same results as query.cu (and as the reference's query_ball_point_kernel / query_cylinder_point_kernel,
// pointnet2/_ext_src/src/ball_query_gpu.cu:14-49, cylinder_query_gpu.cu:9-57): for every centre the FIRST
// nsample points IN INDEX ORDER that pass the test, padded with the first hit, zeros if none.  The linear
// kernels test every (centre, point) pair: 20.5 M tests per 20k-point scene for CloudCrop, of which ~2 % hit.
// Here the cloud is binned once into a uniform grid over its bounding box (cell edge = half the query's
// bounding radius, at most 64 cells per axis and 32768 in all):
//   1. bounds kernel   : per-cloud bounding box (ordered-integer atomics)
//   2. count kernel    : grid origin / cell size from the box, cell of every point, per-cell histogram
//   3. scan kernel     : exclusive prefix sum of the histogram (one CTA per cloud)
//   4. scatter kernel  : points copied cell by cell as {x, y, z, index}
//   5. query kernel    : one CTA per centre walks only the cell rows that intersect the query's bounding
//                        sphere and applies the SAME predicate arithmetic; every hit sets its bit in a
//                        shared-memory BITMAP over point indices (one per output list), and a popcount
//                        prefix over that bitmap is the hit's rank in index order -- "first nsample in
//                        index order" without a comparison sort or a hit buffer.
// Conservative pruning keeps it exact: a row of cells is skipped only if its distance to the centre exceeds
// R' (a hair above the true bounding radius), measured in the same monotone cell coordinate the points were
// binned with.  The cylinder bound assumes `rot` is orthonormal; each centre checks that, and a centre whose
// matrix is not falls back to the plain linear scan of the whole cloud inside the same kernel.
#include "common.cuh"
#include "query_pred.cuh"

namespace {

constexpr int GMAX = 64;                 // at most GMAX cells per axis
constexpr int GCELLS = 32768;            // and GCELLS per cloud
constexpr int GT = 256;

struct GridParams {                      // one per cloud, in the workspace
  float ox, oy, oz;                      // origin
  float inv, cell;                       // 1 / cell edge, cell edge
  int nx, ny, nz;
};

// monotone float -> unsigned (for atomicMax); 0 is below every float
__device__ __forceinline__ unsigned ordered_uint(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_float(unsigned k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
// cell coordinate as a float: monotone non-decreasing in x
__device__ __forceinline__ float cell_pos(float x, float o, float inv) { return __fmul_rn(__fsub_rn(x, o), inv); }
__device__ __forceinline__ int cell_clamp(float t, int n) {
  const int c = (int)floorf(t);                                 // saturating; NaN -> 0
  return max(0, min(n - 1, c));
}

// ---- 1. bounds: box[6*cloud + a] = max over points of key(a): a < 3: ~ordered(v) (i.e. the minimum), else ordered(v).
//         Zero-initialised = "no finite point".
__global__ void __launch_bounds__(GT)
grid_bounds_kernel(const float *__restrict__ xyz, int n, unsigned *__restrict__ box) {
  __shared__ unsigned red[GT / 32][6];
  const int cloud = blockIdx.y;
  const float *p = xyz + (size_t)cloud * n * 3;
  unsigned m[6];
#pragma unroll
  for (int a = 0; a < 6; ++a) m[a] = 0u;
  for (int i = blockIdx.x * GT + threadIdx.x; i < n; i += gridDim.x * GT) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float v = __ldg(p + 3 * (size_t)i + a);
      if (fabsf(v) < 1e30f) {                                   // NaN / inf / absurd values do not shape the grid
        const unsigned k = ordered_uint(v);
        m[a] = max(m[a], ~k);
        m[3 + a] = max(m[3 + a], k);
      }
    }
  }
#pragma unroll
  for (int a = 0; a < 6; ++a) {
    m[a] = __reduce_max_sync(0xffffffffu, m[a]);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][a] = m[a];
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    unsigned v = red[0][threadIdx.x];
#pragma unroll
    for (int w = 1; w < GT / 32; ++w) v = max(v, red[w][threadIdx.x]);
    atomicMax(box + cloud * 6 + threadIdx.x, v);
  }
}

__device__ GridParams make_grid(const unsigned *__restrict__ box, float cell) {
  float lo[3], ext[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const unsigned kmin = box[a], kmax = box[3 + a];
    float l = 0.f, h = 0.f;
    if (kmax != 0u) { l = ordered_float(~kmin); h = ordered_float(kmax); }
    lo[a] = l;
    ext[a] = fmaxf(h - l, 0.f);
  }
  float c = cell;
  int d[3];
  for (int it = 0; it < 400; ++it) {                            // grow the cell until the grid fits
    bool fits = true;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float cells = floorf(ext[a] / c) + 1.f;
      fits = fits && cells <= (float)GMAX;
      d[a] = (int)fminf(cells, (float)GMAX);
    }
    if (fits && d[0] * d[1] * d[2] <= GCELLS) break;
    c *= 1.25f;
  }
  if (d[0] * d[1] * d[2] > GCELLS) d[0] = d[1] = d[2] = 1;      // (not reached with finite extents)
  GridParams g;
  g.ox = lo[0]; g.oy = lo[1]; g.oz = lo[2];
  g.cell = c;
  g.inv = 1.0f / c;
  g.nx = d[0]; g.ny = d[1]; g.nz = d[2];
  return g;
}

// ---- 2. count (the first CTA of each cloud also publishes the grid parameters)
__global__ void __launch_bounds__(GT)
grid_count_kernel(const float *__restrict__ xyz, int n, const unsigned *__restrict__ box, float cell,
                  GridParams *__restrict__ gp, int *__restrict__ cell_of, int *__restrict__ counts) {
  __shared__ GridParams gs;
  const int cloud = blockIdx.y;
  if (threadIdx.x == 0) {
    gs = make_grid(box + cloud * 6, cell);
    if (blockIdx.x == 0) gp[cloud] = gs;
  }
  __syncthreads();
  const int i = blockIdx.x * GT + threadIdx.x;
  if (i >= n) return;
  const GridParams g = gs;
  const float *p = xyz + ((size_t)cloud * n + i) * 3;
  const int cx = cell_clamp(cell_pos(__ldg(p), g.ox, g.inv), g.nx);
  const int cy = cell_clamp(cell_pos(__ldg(p + 1), g.oy, g.inv), g.ny);
  const int cz = cell_clamp(cell_pos(__ldg(p + 2), g.oz, g.inv), g.nz);
  const int c = (cz * g.ny + cy) * g.nx + cx;
  cell_of[(size_t)cloud * n + i] = c;
  atomicAdd(counts + (size_t)cloud * (GCELLS + 1) + c, 1);
}

// ---- 3. exclusive scan of the used counters of each cloud, in place (one CTA of 1024 threads per cloud);
//         starts[ncell] = n closes the last cell.
__global__ void __launch_bounds__(1024)
grid_scan_kernel(int *__restrict__ counts, const GridParams *__restrict__ gp) {
  __shared__ int warp_sum[32];
  const GridParams g = gp[blockIdx.x];
  const int ncell = g.nx * g.ny * g.nz;
  const int per = (ncell + 1023) / 1024;
  int *c = counts + (size_t)blockIdx.x * (GCELLS + 1);
  const int t = threadIdx.x;
  const int lo = min(ncell, t * per), hi = min(ncell, lo + per);
  int s = 0;
  for (int k = lo; k < hi; ++k) s += c[k];
  int inc = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, inc, o);
    if ((t & 31) >= o) inc += u;
  }
  if ((t & 31) == 31) warp_sum[t >> 5] = inc;
  __syncthreads();
  if (t < 32) {
    int w = warp_sum[t];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, w, o);
      if (t >= o) w += u;
    }
    warp_sum[t] = w;
  }
  __syncthreads();
  int run = inc - s + ((t >> 5) ? warp_sum[(t >> 5) - 1] : 0);
  for (int k = lo; k < hi; ++k) { const int v = c[k]; c[k] = run; run += v; }
  if (t == 1023) c[ncell] = run;
}

// ---- 4. scatter (order inside a cell is arbitrary: the query restores index order itself)
__global__ void __launch_bounds__(GT)
grid_scatter_kernel(const float *__restrict__ xyz, int n, const int *__restrict__ cell_of,
                    const int *__restrict__ starts, int *__restrict__ fill, float4 *__restrict__ sorted) {
  const int cloud = blockIdx.y;
  const int i = blockIdx.x * GT + threadIdx.x;
  if (i >= n) return;
  const int c = cell_of[(size_t)cloud * n + i];
  const int pos = starts[(size_t)cloud * (GCELLS + 1) + c] + atomicAdd(fill + (size_t)cloud * GCELLS + c, 1);
  const float *p = xyz + ((size_t)cloud * n + i) * 3;
  sorted[(size_t)cloud * n + pos] = make_float4(__ldg(p), __ldg(p + 1), __ldg(p + 2), __int_as_float(i));
}

// ---- 5. query.  A hit word is  point index | (list mask << 24).
template <class Pred, int NH, bool BALL>
__device__ __forceinline__ void linear_scan(const Pred &pred, const float *__restrict__ pts, int n, const HList &hmax,
                                            int nsample, size_t list_stride, int32_t *__restrict__ out, int lane) {
  int cnt[NH], first[NH];
#pragma unroll
  for (int h = 0; h < NH; ++h) { cnt[h] = 0; first[h] = 0; }
  for (int i0 = 0; i0 < n; i0 += 32) {
    const int k = i0 + lane;
    bool ok = false;
    float hv = 0.f;
    if (k < n) ok = pred.test(__ldg(pts + 3 * (size_t)k), __ldg(pts + 3 * (size_t)k + 1), __ldg(pts + 3 * (size_t)k + 2), hv);
    if (!__any_sync(0xffffffffu, ok)) continue;
    bool all_full = true;
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      if (cnt[h] < nsample) {
        const bool hit = BALL ? ok : (ok && hv < hmax.v[h]);
        const unsigned bal = __ballot_sync(0xffffffffu, hit);
        if (bal) {
          if (cnt[h] == 0) first[h] = i0 + __ffs(bal) - 1;
          const int pos = cnt[h] + __popc(bal & lanemask_lt());
          if (hit && pos < nsample) out[h * list_stride + pos] = k;
          cnt[h] += __popc(bal);
        }
        if (cnt[h] < nsample) all_full = false;
      }
    }
    if (all_full) break;
  }
#pragma unroll
  for (int h = 0; h < NH; ++h) {
    const int fill = cnt[h] > 0 ? first[h] : 0;
    for (int s = min(cnt[h], nsample) + lane; s < nsample; s += 32) out[h * list_stride + s] = fill;
  }
}

__device__ __forceinline__ bool orthonormal(const float *__restrict__ r) {
  bool ok = true;
#pragma unroll
  for (int a = 0; a < 3; ++a)
#pragma unroll
    for (int c = a; c < 3; ++c) {
      // columns a and c of the row-major 3x3
      const float d = r[a] * r[c] + r[3 + a] * r[3 + c] + r[6 + a] * r[6 + c];
      ok = ok && (fabsf(d - (a == c ? 1.f : 0.f)) < 1e-3f);   // false for NaN
    }
  return ok;
}

constexpr int ROWS = 64;                 // (y, z) cell rows a centre's bounding sphere can touch (<= 6 x 6)
constexpr int CHUNKS = 256;              // window of the chunk table (a chunk = up to 32 consecutive binned points)

// distance (in cells) from coordinate t to the row of cells k of an axis with n cells; the outermost cells also
// hold everything that was clamped into them
__device__ __forceinline__ float row_dist(float t, int k, int n) {
  float d = 0.f;
  if (k > 0 && t < (float)k) d = (float)k - t;
  if (k < n - 1 && t > (float)(k + 1)) d = t - (float)(k + 1);
  return d;
}

// One CTA of NT threads per centre.  Dynamic shared memory: one bitmap over point indices per output list.
//   rows    : warp 0 turns the bounding sphere into cell rows (x-contiguous, hence contiguous in the binned
//             array) and cuts them into 32-point chunks; the other warps clear the bitmaps meanwhile
//   walk    : warps take chunks round-robin; a point that passes the test sets its bit in every list it belongs to
//   rank    : a block scan of the word popcounts gives every bitmap word the number of hits before it, i.e. the
//             rank IN INDEX ORDER of its first hit -- a sort by a bitmap, no comparisons, no hit buffer to overflow
//   emit    : every thread writes the hits of its own words whose rank is below nsample; the rest of the list is
//             padded with the lowest set bit
template <class Pred, int NH, bool BALL, int NT>
__global__ void __launch_bounds__(NT)
grid_query_kernel(const float *__restrict__ new_xyz, const float *__restrict__ xyz, const float *__restrict__ rot,
                  const GridParams *__restrict__ gp, const int *__restrict__ starts,
                  const float4 *__restrict__ sorted_pts, int b, int n, int m, float radius, float hmin, HList hmax,
                  float rb, int nsample, int32_t *__restrict__ idx) {
  constexpr int NW = NT / 32;
  extern __shared__ unsigned bm[];                              // [NH][Wp]
  __shared__ unsigned chunk_tab[CHUNKS];
  __shared__ int row_s[ROWS], row_e[ROWS], row_k[ROWS];
  __shared__ int wsum[NH][NW], wmin[NH][NW];
  __shared__ int s_rows, s_more, s_chunks, s_linear;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int W = (n + 31) >> 5, Wp = (W + 3) & ~3;
  const int bi = blockIdx.x / m, j = blockIdx.x - bi * m;

  const float *ctr = new_xyz + ((size_t)bi * m + j) * 3;
  const float *rj = BALL ? nullptr : rot + ((size_t)bi * m + j) * 9;
  const size_t list_stride = (size_t)b * m * nsample;
  int32_t *out = idx + ((size_t)bi * m + j) * nsample;

  Pred pred;
  pred.init(ctr, rj, radius, hmin);
  const float4 *sp = sorted_pts + (size_t)bi * n;
  for (int rbase = 0;; rbase += ROWS) {                         // ROWS cell rows at a time (one round unless cells << R')
    if (warp == 0) {
      bool linear = false;
      if (!BALL) linear = !orthonormal(rj);
      const GridParams g = gp[bi];
      const int *st = starts + (size_t)bi * (GCELLS + 1);
      const float tx = cell_pos(ctr[0], g.ox, g.inv), ty = cell_pos(ctr[1], g.oy, g.inv), tz = cell_pos(ctr[2], g.oz, g.inv);
      const float rc = rb * g.inv * 1.001f + 1e-3f;             // bounding radius in cells, with slack for the binning
      const int y0 = cell_clamp(ty - rc, g.ny), y1 = cell_clamp(ty + rc, g.ny);
      const int z0 = cell_clamp(tz - rc, g.nz), z1 = cell_clamp(tz + rc, g.nz);
      const int ry = y1 - y0 + 1, nrows = min(ROWS, ry * (z1 - z0 + 1) - rbase);
      int carry = 0;
      if (!linear) {
        for (int r0 = 0; r0 < nrows; r0 += 32) {
          const int r = r0 + lane;
          int s = 0, e = 0;
          if (r < nrows) {
            const int qz = (rbase + r) / ry;
            const int ky = y0 + (rbase + r - qz * ry), kz = z0 + qz;
            const float dy = row_dist(ty, ky, g.ny), dz = row_dist(tz, kz, g.nz);
            const float rem = rc * rc - dy * dy - dz * dz;
            if (rem >= 0.f) {
              const float wx = sqrtf(rem);
              const int x0 = cell_clamp(tx - wx, g.nx), x1 = cell_clamp(tx + wx, g.nx);
              const int row = (kz * g.ny + ky) * g.nx;
              s = __ldg(st + row + x0);
              e = __ldg(st + row + x1 + 1);
            }
          }
          const int nc = (e - s + 31) >> 5;
          int inc = nc;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
          }
          if (r < nrows) { row_s[r] = s; row_e[r] = e; row_k[r] = carry + inc - nc; }
          carry += __shfl_sync(0xffffffffu, inc, 31);
        }
      }
      if (lane == 0) {
        s_rows = nrows;
        s_more = ry * (z1 - z0 + 1) - rbase > ROWS;
        s_chunks = carry;
        s_linear = linear ? 1 : 0;
      }
    } else if (rbase == 0) {
      uint4 *z = (uint4 *)bm;
      for (int w = tid - 32; w < NH * Wp / 4; w += NT - 32) z[w] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    if (s_linear) {                                             // rare: one warp scans the whole cloud
      if (warp == 0)
        linear_scan<Pred, NH, BALL>(pred, xyz + (size_t)bi * n * 3, n, hmax, nsample, list_stride, out, lane);
      return;
    }
    const int nrows = s_rows, nchunks = s_chunks;
    const bool more = s_more != 0;
    for (int base = 0; base < nchunks; base += CHUNKS) {
      if (base) __syncthreads();                                // the previous window has been consumed
      if (tid < nrows) {
        const int s = row_s[tid], e = row_e[tid], k0 = row_k[tid] - base;
        for (int c = max(0, -k0); s + 32 * c < e && k0 + c < CHUNKS; ++c)
          chunk_tab[k0 + c] = (unsigned)(s + 32 * c) | ((unsigned)(min(32, e - s - 32 * c) - 1) << 24);
      }
      __syncthreads();
      const int nwin = min(CHUNKS, nchunks - base);
      for (int k = warp; k < nwin; k += NW) {
        const unsigned ent = chunk_tab[k];
        if (lane <= (int)(ent >> 24)) {
          const float4 p = __ldg(sp + (ent & 0xffffffu) + lane);
          float hv;
          if (pred.test(p.x, p.y, p.z, hv)) {
            const unsigned id = (unsigned)__float_as_int(p.w);
            unsigned *wp = bm + (id >> 5);
            const unsigned bit = 1u << (id & 31u);
#pragma unroll
            for (int h = 0; h < NH; ++h)
              if (BALL || hv < hmax.v[h]) atomicOr(wp + h * Wp, bit);
          }
        }
      }
    }
    if (!more) break;
    __syncthreads();                                            // row and chunk tables are free again
  }
  __syncthreads();
  const int cw = (W + NT - 1) / NT;
  const int w0 = min(W, tid * cw), w1 = min(W, w0 + cw);
  int cnt[NH], incl[NH];
#pragma unroll
  for (int h = 0; h < NH; ++h) {
    const unsigned *bh = bm + h * Wp;
    int c = 0, lowest = 0x7fffffff;
    for (int w = w1 - 1; w >= w0; --w) {
      const unsigned v = bh[w];
      c += __popc(v);
      if (v) lowest = w * 32 + __ffs(v) - 1;
    }
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += u;
    }
    lowest = __reduce_min_sync(0xffffffffu, lowest);
    if (lane == 31) { wsum[h][warp] = inc; wmin[h][warp] = lowest; }
    cnt[h] = c;
    incl[h] = inc;
  }
  __syncthreads();
#pragma unroll
  for (int h = 0; h < NH; ++h) {
    int run = incl[h] - cnt[h], total = 0, first = 0x7fffffff;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const int ws = wsum[h][w];
      run += (w < warp) ? ws : 0;
      total += ws;
      first = min(first, wmin[h][w]);
    }
    int32_t *o = out + h * list_stride;
    if (cnt[h] > 0 && run < nsample) {
      const unsigned *bh = bm + h * Wp;
      for (int w = w0; w < w1 && run < nsample; ++w) {
        unsigned v = bh[w];
        while (v && run < nsample) {
          o[run++] = w * 32 + __ffs(v) - 1;
          v &= v - 1;
        }
      }
    }
    const int fill = total > 0 ? first : 0;
    for (int s = total + tid; s < nsample; s += NT) o[s] = fill;
  }
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Layout { size_t gp, box, counts, fill, cell_of, sorted, total; };

Layout layout(int b, int n) {
  Layout L;
  size_t off = 0;
  const size_t bb = (size_t)(b > 0 ? b : 1), nn = (size_t)(n > 0 ? n : 1);
  auto take = [&](size_t bytes) { const size_t o = off; off += align_up(bytes, 256); return o; };
  L.gp = take(bb * sizeof(GridParams));
  L.box = take(bb * 6 * sizeof(unsigned));                     // box, counts and fill are cleared by one memset
  L.counts = take(bb * (GCELLS + 1) * sizeof(int));
  L.fill = take(bb * GCELLS * sizeof(int));
  L.cell_of = take(bb * nn * sizeof(int));
  L.sorted = take(bb * nn * sizeof(float4));
  L.total = off;
  return L;
}

struct Grid { const GridParams *gp; const int *starts; const float4 *sorted; };

// bins `xyz` (b clouds of n points) into the workspace; `cell` is the requested cell edge
int build_grid(const float *xyz, int b, int n, float cell, char *ws, const Layout &L, cudaStream_t st, Grid &out) {
  GridParams *gp = (GridParams *)(ws + L.gp);
  unsigned *box = (unsigned *)(ws + L.box);
  int *counts = (int *)(ws + L.counts), *fill = (int *)(ws + L.fill);
  int *cell_of = (int *)(ws + L.cell_of);
  float4 *sorted = (float4 *)(ws + L.sorted);
  B200_CUDA(cudaMemsetAsync(box, 0, L.cell_of - L.box, st));
  const int per_cloud = ceil_div(n, GT);
  grid_bounds_kernel<<<dim3(max(1, min(per_cloud, 16)), b), GT, 0, st>>>(xyz, n, box);
  B200_LAUNCH_CHECK();
  grid_count_kernel<<<dim3(per_cloud, b), GT, 0, st>>>(xyz, n, box, cell, gp, cell_of, counts);
  B200_LAUNCH_CHECK();
  grid_scan_kernel<<<b, 1024, 0, st>>>(counts, gp);
  B200_LAUNCH_CHECK();
  grid_scatter_kernel<<<dim3(per_cloud, b), GT, 0, st>>>(xyz, n, cell_of, counts, fill, sorted);
  B200_LAUNCH_CHECK();
  out.gp = gp; out.starts = counts; out.sorted = sorted;
  return 0;
}

size_t query_smem(int n, int nh) {
  const size_t Wp = ((size_t)((n + 31) / 32) + 3) / 4 * 4;
  return (size_t)nh * Wp * 4;
}

template <class Pred, int NH, bool BALL, int NT>
int launch_kernel(const float *new_xyz, const float *xyz, const float *rot, int b, int n, int m, float radius,
                  float hmin, const HList &hmax, float rb, int nsample, int32_t *idx, const Grid &g, size_t smem,
                  cudaStream_t st) {
  auto kern = grid_query_kernel<Pred, NH, BALL, NT>;
  if (smem > 48 * 1024) B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)(b * m), NT, smem, st>>>(new_xyz, xyz, rot, g.gp, g.starts, g.sorted, b, n, m, radius, hmin, hmax,
                                           rb, nsample, idx);
  B200_LAUNCH_CHECK();
  return 0;
}

int g_query_threads = 0;                 // 0 = automatic; b200_query_grid_set_threads() pins it (tuning hook)

// `cell` > 0: bin the cloud first (cell edge `cell`); `cell` == 0: the workspace already holds the grid of `xyz`
template <class Pred, int NH, bool BALL>
int launch_grid_query(const float *new_xyz, const float *xyz, const float *rot, int b, int n, int m, float radius,
                      float hmin, const HList &hmax, float rb, float cell, int nsample, int32_t *idx, void *ws,
                      size_t ws_bytes, cudaStream_t st) {
  const Layout L = layout(b, n);
  if (!ws || ws_bytes < L.total) return B200_EWORKSPACE;
  const size_t smem = query_smem(n, NH);
  const long long grid = (long long)b * m;
  if (smem > 200 * 1024 || n >= (1 << 24) || b > 65535 || grid > 0x7fffffffLL) return B200_ECAPACITY;
  Grid g;
  if (cell > 0.f) {
    const int rc = build_grid(xyz, b, n, cell, (char *)ws, L, st, g);
    if (rc) return rc;
  } else {
    g.gp = (const GridParams *)((char *)ws + L.gp);
    g.starts = (const int *)((char *)ws + L.counts);
    g.sorted = (const float4 *)((char *)ws + L.sorted);
  }
  // few centres: 8 warps each so that the SMs are busy and a centre finishes sooner; many: 4 warps each
  const int nt = g_query_threads ? g_query_threads : (grid * 8 <= 64LL * b200_sm_count() ? 256 : 128);
  if (nt == 256)
    return launch_kernel<Pred, NH, BALL, 256>(new_xyz, xyz, rot, b, n, m, radius, hmin, hmax, rb, nsample, idx, g, smem, st);
  if (nt == 64)
    return launch_kernel<Pred, NH, BALL, 64>(new_xyz, xyz, rot, b, n, m, radius, hmin, hmax, rb, nsample, idx, g, smem, st);
  return launch_kernel<Pred, NH, BALL, 128>(new_xyz, xyz, rot, b, n, m, radius, hmin, hmax, rb, nsample, idx, g, smem, st);
}

// bounding radius with head-room for the fp32 rounding of the tests (and, for the cylinder, for a rotation
// matrix that is orthonormal only to 1e-3); false when the grid cannot prune (non-finite or non-positive)
bool usable_radius(float r, float *out) {
  const float rb = r * 1.01f;
  if (!(rb > 0.f) || !(rb < 1e30f)) return false;
  *out = rb;
  return true;
}

int ball_any(const float *new_xyz, const float *xyz, int b, int n, int m, float radius, int nsample, int32_t *idx,
             void *ws, size_t ws_bytes, bool prebuilt, cudaStream_t st) {
  if (b < 0 || n < 0 || m < 0 || nsample < 0) return B200_EINVAL;
  if (b == 0 || m == 0 || nsample == 0) return 0;
  if (!new_xyz || !idx || (n > 0 && !xyz)) return B200_EINVAL;
  float rb;
  if (n == 0 || !usable_radius(fabsf(radius), &rb))
    return b200_ball_query(new_xyz, xyz, b, n, m, radius, nsample, idx, st);
  HList h = {};
  const int rc = launch_grid_query<BallPred, 1, true>(new_xyz, xyz, nullptr, b, n, m, radius, 0.f, h, rb,
                                                      prebuilt ? 0.f : 0.5f * rb, nsample, idx, ws, ws_bytes, st);
  if (rc == B200_ECAPACITY) return b200_ball_query(new_xyz, xyz, b, n, m, radius, nsample, idx, st);
  return rc;
}

int cylinder_any(const float *new_xyz, const float *xyz, const float *rot, int b, int n, int m, float radius,
                 float hmin, const float *hmax, int nh, int nsample, int32_t *idx, void *ws, size_t ws_bytes,
                 bool prebuilt, cudaStream_t st) {
  if (b < 0 || n < 0 || m < 0 || nsample < 0 || nh < 0) return B200_EINVAL;
  if (nh > MAXH) return B200_ECAPACITY;
  if (b == 0 || m == 0 || nsample == 0 || nh == 0) return 0;
  if (!new_xyz || !rot || !idx || !hmax || (n > 0 && !xyz)) return B200_EINVAL;
  HList h = {};
  float top = 0.f;                                             // a hit has hmin < x_rot < max(hmax)
  for (int i = 0; i < nh; ++i) { h.v[i] = hmax[i]; top = fmaxf(top, fabsf(hmax[i])); }
  const float ax = fmaxf(top, fabsf(hmin));
  float rb;
  if (n == 0 || !usable_radius(sqrtf(radius * radius + ax * ax), &rb))
    return b200_cylinder_query_multi(new_xyz, xyz, rot, b, n, m, radius, hmin, hmax, nh, nsample, idx, st);
  const float cell = prebuilt ? 0.f : 0.5f * rb;
  int rc = B200_EINVAL;
#define CYL_CASE(N)                                                                                               \
  case N:                                                                                                         \
    rc = launch_grid_query<CylPred, N, false>(new_xyz, xyz, rot, b, n, m, radius, hmin, h, rb, cell, nsample, idx, \
                                              ws, ws_bytes, st);                                                  \
    break;
  switch (nh) { CYL_CASE(1) CYL_CASE(2) CYL_CASE(3) CYL_CASE(4) CYL_CASE(5) CYL_CASE(6) CYL_CASE(7) CYL_CASE(8) }
#undef CYL_CASE
  if (rc == B200_ECAPACITY)
    return b200_cylinder_query_multi(new_xyz, xyz, rot, b, n, m, radius, hmin, hmax, nh, nsample, idx, st);
  return rc;
}

}  // namespace

extern "C" size_t b200_query_grid_workspace(int b, int n) { return layout(b, n).total; }

extern "C" void b200_query_grid_set_threads(int threads) {
  g_query_threads = (threads == 64 || threads == 128 || threads == 256) ? threads : 0;
}

extern "C" int b200_query_grid_build(const float *xyz, int b, int n, float cell, void *workspace,
                                     size_t workspace_bytes, void *stream) {
  if (b <= 0 || n <= 0 || !xyz || !(cell > 0.f) || !(cell < 1e30f) || b > 65535) return B200_EINVAL;
  const Layout L = layout(b, n);
  if (!workspace || workspace_bytes < L.total) return B200_EWORKSPACE;
  Grid g;
  return build_grid(xyz, b, n, cell, (char *)workspace, L, (cudaStream_t)stream, g);
}

extern "C" int b200_ball_query_grid(const float *new_xyz, const float *xyz, int b, int n, int m, float radius,
                                    int nsample, int32_t *idx, void *workspace, size_t workspace_bytes,
                                    void *stream) {
  return ball_any(new_xyz, xyz, b, n, m, radius, nsample, idx, workspace, workspace_bytes, false, (cudaStream_t)stream);
}

extern "C" int b200_ball_query_prebuilt(const float *new_xyz, const float *xyz, int b, int n, int m, float radius,
                                        int nsample, int32_t *idx, const void *grid, size_t grid_bytes,
                                        void *stream) {
  return ball_any(new_xyz, xyz, b, n, m, radius, nsample, idx, (void *)grid, grid_bytes, true, (cudaStream_t)stream);
}

extern "C" int b200_cylinder_query_multi_grid(const float *new_xyz, const float *xyz, const float *rot, int b,
                                              int n, int m, float radius, float hmin, const float *hmax, int nh,
                                              int nsample, int32_t *idx, void *workspace,
                                              size_t workspace_bytes, void *stream) {
  return cylinder_any(new_xyz, xyz, rot, b, n, m, radius, hmin, hmax, nh, nsample, idx, workspace, workspace_bytes,
                      false, (cudaStream_t)stream);
}

extern "C" int b200_cylinder_query_multi_prebuilt(const float *new_xyz, const float *xyz, const float *rot, int b,
                                                  int n, int m, float radius, float hmin, const float *hmax,
                                                  int nh, int nsample, int32_t *idx, const void *grid,
                                                  size_t grid_bytes, void *stream) {
  return cylinder_any(new_xyz, xyz, rot, b, n, m, radius, hmin, hmax, nh, nsample, idx, (void *)grid, grid_bytes, true,
                      (cudaStream_t)stream);
}
