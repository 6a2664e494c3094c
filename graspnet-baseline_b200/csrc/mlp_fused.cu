// mlp_fused.cu -- group -> centre / normalise / rotate -> shared MLP (3 x [1x1 conv + BN + ReLU]) -> max-pool over the
// neighbourhood, as ONE sm_100a kernel with the two wide layers on the 5th-generation tensor cores.
//
// Replaces, for the modules whose grouped input is xyz only (SA1 of the backbone and CloudCrop), the chain
//   QueryAndGroup / CylinderQueryAndGroup.forward   (reference pointnet2/pointnet2_utils.py:313-372, 488-555)
//   -> pt_utils.SharedMLP                            (pointnet2/pytorch_utils.py:11-36: Conv2d 1x1, BatchNorm2d, ReLU)
//   -> F.max_pool2d(kernel=[1, nsample])             (pointnet2/pointnet2_modules.py:243-257, models/modules.py:127-141)
// in eval mode (BatchNorm folded into the conv weights by the host wrapper).  The reference materialises the
// (B, C, npoint, nsample) activations of every layer in HBM (CloudCrop: 262,144 rows x (64 + 128 + 256) channels x 4 B,
// written and re-read by conv, BN and ReLU kernels); here a row's activations never leave the SM.
//
// Shape of the work: a "row" is one neighbour of one neighbourhood list; a tile is 128 rows (= 128 / nsample lists).
//   layer 1 (3 -> C1)  : K = 3 is no tensor-core shape: each row thread gathers its point, centres / scales /
//                        rotates it and evaluates the C1 outputs with FFMA.
//   layer 2 (C1 -> C2) : D2[128, C2] = A2[128, C1] . W2^T     tcgen05.mma kind::tf32, M = 128, N = C2, K = 8 per MMA
//   layer 3 (C2 -> C3) : D3[128, C3] = A3[128, C2] . W3^T     accumulators in TMEM (columns [0, C2) and [C2, C2 + C3))
//   max-pool           : over the nsample rows of a list, on the layer-3 epilogue registers.
// fp32-grade results from tf32 tensor cores ("3xTF32"): x = hi + lo with hi = tf32(x), lo = tf32(x - hi), and
// a.b ~= a_lo.b_hi + a_hi.b_lo + a_hi.b_hi, accumulated in fp32; the dropped a_lo.b_lo term is 2^-22 relative.
// `passes` = 1 runs plain TF32 (what cuDNN does for the reference's convs under torch's default allow_tf32).
// Modules with point features (SA2-SA4: grouped input = [xyz | C features], K = 131 / 259): the first layer is a
// tensor-core layer too.  Its A operand is gathered straight into K-slabs -- the features are read CHANNELS-LAST
// (B, N, C), one neighbour's 16 channels = four 16-byte loads, the K order is [features | x y z | zero pad to 16]
// (the host packs W1's columns the same way) -- and the three accumulators (C1 + C2 + C3 = 512 columns for
// 128/128/256) fill TMEM exactly.  The kernel writes its output a second time channels-last for the next module.
//
// Pipeline inside a CTA (persistent over tiles), 10 warps:
//   warps 0-7  row threads.  Warp w works on rows 32 (w % 4) .. + 31 (the TMEM lanes a warp may read) and on the
//              K-slabs s with s % 2 == w / 4, so every scheduler holds two row warps.  They gather + run layer 1, then
//              the epilogues, and PRODUCE the A operand of the next layer in K-slabs of 16 channels (hi and lo parts,
//              16 KB) into a 4-slot ring: the epilogue of layer l for channels [16 s, 16 s + 16) is the K-slab s of
//              layer l + 1, so the tensor core starts layer l + 1 while the CUDA cores are still converting layer l,
//              and no layer's activations are ever held whole.
//   warp 8     lane 0 issues the MMAs: per slab 2 K-steps x 3 passes, then tcgen05.commit releases the A slot and the
//              weight slot; after a layer's last slab a commit tells the row threads its accumulators are complete.
//   warp 9     lane 0 streams the weight slabs (host-packed: hi | lo, core-matrix order) global/L2 -> shared with
//              cp.async.bulk into a 4-slot ring (W2 + W3 in 3xTF32 form are 320 KB: they cannot stay resident).
// Max-pool: max commutes with the (monotone) bias add and ReLU, so the layer-3 epilogue reduces the RAW accumulators
// over the 32 rows of a warp with one redux.sync.max.NaN.f32 per channel and bias + ReLU are applied once per
// (list, channel) when the warps' partial maxima are combined.
// Shared-memory operand layout (both A and B are "K-major, no swizzle" UMMA operands): for each 16-byte K-chunk
// (4 tf32) the rows are contiguous, i.e. byte offset = chunk * rows * 16 + row * 16; a thread stores its row's
// chunk with one st.shared.v4 and a warp's stores are 512 contiguous bytes.
#include "tc_common.cuh"

namespace {

constexpr int TM = 128;                        // rows per tile = TMEM lanes = row threads
constexpr int KS = 16;                         // channels per K-slab
constexpr int NA = 4;                          // A ring slots
constexpr int NWS = 4;                         // weight ring slots
constexpr int A_PART = TM * KS * 4;            // 8 KB: the hi (or lo) part of an A slab
constexpr int A_SLOT = 2 * A_PART;
constexpr int MAXC = 256;                      // widest layer
constexpr int W_SLOT = MAXC * KS * 4 * 2;      // 32 KB: hi | lo of a weight slab of the widest layer
constexpr int MAXE = 8;                        // pooling entries per tile (128 / nsample, at least one per warp)
constexpr int ROW_THREADS = 256;               // two per row: even / odd K-slabs
constexpr int NTHREADS = ROW_THREADS + 64;

struct Params {
  const float *xyz, *new_xyz, *rot;
  const int32_t *idx;
  int G, B, N, m, ns;
  float radius;                                // > 0: grouped_xyz /= radius (normalize_xyz)
  const float *feat_t;                         // (B, N, C) channels-last point features, or NULL (xyz-only module)
  int C;
  const float *w1, *b1;                        // xyz-only: first layer (C0, 3), (C0) for the CUDA cores, BN folded
  int C0;
  int NG;                                      // tensor-core layers: 2 (xyz-only) or 3 (with features)
  const float *wp[3];                          // packed weights (mlp_pack_kernel) of the tensor-core layers
  const float *bias[3];
  int Kd[3], Nd[3];
  int passes;
  long long Q;                                 // lists = B * m * G
  int tiles;
  float *out;                                  // (B, C3, m, G)
  float *out_t;                                // (B, m, G, C3) channels-last copy for the next module, or NULL
};

struct Smem {
  // offsets into the dynamic shared memory block
  static constexpr int a_ring = 0;
  static constexpr int w_ring = a_ring + NA * A_SLOT;
  static constexpr int consts = w_ring + NWS * W_SLOT;               // {w1[c][0..2], b1[c]} (float4 x MAXC) | bias[3][MAXC]
  static constexpr int pool = consts + (MAXC * 4 + 3 * MAXC) * 4;   // MAXE x C3
  static constexpr int bars = pool + 2 * MAXE * MAXC * 4;            // (two pools: tiles alternate); then a_full,
                                                                     // a_empty, w_full, w_empty, acc_full[3]
  static constexpr int holder = bars + (2 * NA + 2 * NWS + 3) * 8;
  static constexpr int total = holder + 16;
};

#ifdef MLP_PROFILE
// timeline of CTA 0 (tools/mlp_prof.py): clock64 stamps [role][tile iteration][event]
constexpr int PROF_TILES = 16, PROF_EV = 8;
__device__ unsigned long long g_mlp_prof[3][PROF_TILES][PROF_EV];
#define PROF(role, it, ev)                                                                   \
  do {                                                                                       \
    if (blockIdx.x == 0 && (it) < PROF_TILES) g_mlp_prof[role][it][ev] = (unsigned long long)clock64(); \
  } while (0)
#else
#define PROF(role, it, ev) do {} while (0)
#endif

__device__ __forceinline__ float relu_keep_nan(float v) { return (v <= 0.f) ? 0.f : v; }

__device__ __forceinline__ float max_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ float warp_max_nan(float v) {
  float r;
  asm volatile("redux.sync.max.NaN.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}

// one 16-channel slab of this row into ring slot (na % NA): waits for the slot, splits hi / lo, publishes.
// hi = tf32(v) rounded to nearest; lo = v - hi exactly (the tensor core drops its low 13 mantissa bits: 2^-21 of v).
__device__ __forceinline__ void put_slab(uint32_t sm, const float (&v)[16], int row, uint32_t na, int passes) {
  const uint32_t slot = na % NA, ph = (na / NA) & 1u;
  const uint32_t a_full = sm + Smem::bars + slot * 8, a_empty = sm + Smem::bars + (NA + slot) * 8;
  tc::mbar_wait(a_empty, ph ^ 1u);
  const uint32_t base = sm + Smem::a_ring + slot * A_SLOT + row * 16;
#pragma unroll
  for (int kc = 0; kc < 4; ++kc) {
    float hi[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      hi[e] = tc::tf32_rna(v[4 * kc + e]);
      lo[e] = __fsub_rn(v[4 * kc + e], hi[e]);
    }
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(base + kc * (TM * 16)), "f"(hi[0]), "f"(hi[1]),
                 "f"(hi[2]), "f"(hi[3]) : "memory");
    if (passes == 3)
      asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(base + A_PART + kc * (TM * 16)), "f"(lo[0]),
                   "f"(lo[1]), "f"(lo[2]), "f"(lo[3]) : "memory");
  }
  tc::fence_proxy_async();
  tc::fence_before_sync();
  tc::mbar_arrive(a_full);
}

__global__ void __launch_bounds__(NTHREADS, 1)
group_mlp_maxpool_kernel(const Params p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t sm = tc::smem_u32(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float4 *w1s = reinterpret_cast<float4 *>(smem_raw + Smem::consts);
  float *biases = reinterpret_cast<float *>(w1s + MAXC);              // [layer][MAXC]
  float *pools = reinterpret_cast<float *>(smem_raw + Smem::pool);
  volatile uint32_t *holder = reinterpret_cast<volatile uint32_t *>(smem_raw + Smem::holder);
  const bool with_feat = p.feat_t != nullptr;

  if (!with_feat)
    for (int i = tid; i < p.C0; i += NTHREADS)
      w1s[i] = make_float4(__ldg(p.w1 + 3 * i), __ldg(p.w1 + 3 * i + 1), __ldg(p.w1 + 3 * i + 2), __ldg(p.b1 + i));
  for (int l = 0; l < p.NG; ++l)
    for (int i = tid; i < p.Nd[l]; i += NTHREADS) biases[l * MAXC + i] = __ldg(p.bias[l] + i);
  if (tid == 0) {
    for (int s = 0; s < NA; ++s) {
      tc::mbar_init(sm + Smem::bars + s * 8, TM);                    // a_full: the 128 row threads of one half
      tc::mbar_init(sm + Smem::bars + (NA + s) * 8, 1);              // a_empty: one tcgen05.commit
    }
    for (int s = 0; s < NWS; ++s) {
      tc::mbar_init(sm + Smem::bars + (2 * NA + s) * 8, 1);          // w_full: expect_tx arrival + bytes
      tc::mbar_init(sm + Smem::bars + (2 * NA + NWS + s) * 8, 1);    // w_empty: one tcgen05.commit
    }
    for (int l = 0; l < 3; ++l)                                      // acc_full[l]: accumulators of tensor-core
      tc::mbar_init(sm + Smem::bars + (2 * NA + 2 * NWS + l) * 8, 1);  // layer l complete (one tcgen05.commit)
    tc::fence_barrier_init();
  }
  if (warp == 8) tc::tmem_alloc_512(sm + Smem::holder);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *holder;
  const uint32_t acc_full = sm + Smem::bars + (2 * NA + 2 * NWS) * 8;     // + 8 l
  uint32_t col[3];                                                    // first TMEM column of layer l's accumulators
  col[0] = 0; col[1] = (uint32_t)p.Nd[0]; col[2] = (uint32_t)(p.Nd[0] + p.Nd[1]);
  const int C3 = p.Nd[p.NG - 1];

  if (warp < 8) {
    // ------------------------------------------------------------------ row threads
    const int half = warp >> 2;                                       // K-slabs s with (s & 1) == half
    const int row = (warp & 3) * 32 + lane;
    const int lpt = TM / p.ns;                                        // lists per tile
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const int s0 = with_feat ? p.Kd[0] / KS : p.C0 / KS;             // slabs gather (+ layer 1) produces per tile
    uint32_t na = 0;                                                  // A slabs published so far (by both halves)
    // gather + layer 1 of a tile: this row's neighbour in the frame of its list's centre, then the C1 outputs on
    // the CUDA cores -> K-slabs of layer 2
    auto gather_layer1 = [&](int tile) {
      const long long q = (long long)tile * lpt + row / p.ns;
      float g0 = 0.f, g1 = 0.f, g2 = 0.f;
      const float4 *frow = nullptr;
      if (q < p.Q) {
        const int gi = (int)(q % p.G);
        const long long t = q / p.G;
        const int seed = (int)(t % p.m), bi = (int)(t / p.m);
        const int i = __ldg(p.idx + ((((long long)gi * p.B + bi) * p.m + seed) * p.ns + row % p.ns));
        const float *pt = p.xyz + ((long long)bi * p.N + i) * 3;
        if (with_feat) frow = reinterpret_cast<const float4 *>(p.feat_t + ((long long)bi * p.N + i) * p.C);
        const float *ct = p.new_xyz + ((long long)bi * p.m + seed) * 3;
        float d0 = __fsub_rn(__ldg(pt), __ldg(ct)), d1 = __fsub_rn(__ldg(pt + 1), __ldg(ct + 1)),
              d2 = __fsub_rn(__ldg(pt + 2), __ldg(ct + 2));
        if (p.radius > 0.f) {
          d0 = __fdiv_rn(d0, p.radius); d1 = __fdiv_rn(d1, p.radius); d2 = __fdiv_rn(d2, p.radius);
        }
        if (p.rot) {                                                  // row vector times the seed's 3x3
          const float *R = p.rot + ((long long)bi * p.m + seed) * 9;
          g0 = __fmaf_rn(d2, __ldg(R + 6), __fmaf_rn(d1, __ldg(R + 3), __fmul_rn(d0, __ldg(R))));
          g1 = __fmaf_rn(d2, __ldg(R + 7), __fmaf_rn(d1, __ldg(R + 4), __fmul_rn(d0, __ldg(R + 1))));
          g2 = __fmaf_rn(d2, __ldg(R + 8), __fmaf_rn(d1, __ldg(R + 5), __fmul_rn(d0, __ldg(R + 2))));
        } else {
          g0 = d0; g1 = d1; g2 = d2;
        }
      }
      if (with_feat) {
        // K order of layer 1: [features (C) | x y z | zero pad]: whole feature slabs are four 16-byte loads
        const int sf = p.C / KS;
        for (int s = half; s < s0; s += 2) {
          float v[16];
          if (s < sf && frow) {
#pragma unroll
            for (int c4 = 0; c4 < 4; ++c4) {
              const float4 f = __ldg(frow + s * 4 + c4);
              v[4 * c4] = f.x; v[4 * c4 + 1] = f.y; v[4 * c4 + 2] = f.z; v[4 * c4 + 3] = f.w;
            }
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) v[c] = 0.f;
            if (s == sf) { v[0] = g0; v[1] = g1; v[2] = g2; }
          }
          put_slab(sm, v, row, na + s, p.passes);
        }
        na += s0;
        return;
      }
      for (int s = half; s < s0; s += 2) {
        float v[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) {
          const float4 w = w1s[s * KS + c];
          const float a = __fmaf_rn(w.z, g2, __fmaf_rn(w.y, g1, __fmul_rn(w.x, g0)));
          v[c] = relu_keep_nan(__fadd_rn(a, w.w));
        }
        put_slab(sm, v, row, na + s, p.passes);
      }
      na += s0;
    };
    // Software pipeline over tiles: the next tile's gather + layer 1 is issued BEFORE this tile's layer-3 epilogue,
    // so the CUDA cores work on tile t + 1 while the tensor core finishes layer 3 of tile t, and the tensor core
    // runs layer 2 of tile t + 1 under the layer-3 epilogue of tile t.  No extra TMEM: layer 2 of t + 1 writes the
    // columns whose epilogue (tile t) is already done, and layer 3 of t + 1 only starts on slabs produced after
    // this thread's layer-3 epilogue of tile t.
    if ((int)blockIdx.x < p.tiles) gather_layer1(blockIdx.x);
    for (int tile = blockIdx.x, it = 0; tile < p.tiles; tile += gridDim.x, ++it) {
      // the pooling scratch alternates between tiles: a warp may start tile t + 1's layer-3 epilogue while
      // another still combines tile t (they meet again at tile t + 1's barrier before the buffer is reused)
      float *pool = pools + (it & 1) * (MAXE * MAXC);
      const bool pr = lane == 0 && (warp & 3) == 0;
      if (pr) PROF(half, it, 0);
      // epilogues of all tensor-core layers but the last -> K-slabs of the next layer
      for (int l = 0; l + 1 < p.NG; ++l) {
        tc::mbar_wait(acc_full + 8 * l, it & 1);
        tc::fence_after_sync();
        if (pr && l == p.NG - 2) PROF(half, it, 1);
        const int sl = p.Nd[l] / KS;
        const float *bl = biases + l * MAXC;
        for (int s = half; s < sl; s += 2) {
          float v[16];
          tc::tmem_ld16(lane_base + col[l] + s * KS, v);
#pragma unroll
          for (int c = 0; c < 16; ++c) v[c] = relu_keep_nan(__fadd_rn(v[c], bl[s * KS + c]));
          put_slab(sm, v, row, na + s, p.passes);
        }
        na += sl;
      }
      if (pr) PROF(half, it, 2);
      if (tile + (int)gridDim.x < p.tiles) gather_layer1(tile + gridDim.x);
      if (pr) PROF(half, it, 3);
      // layer-3 epilogue: max of the raw accumulators over the rows of a pooling entry (a warp's 32 rows, or 16)
      tc::mbar_wait(acc_full + 8 * (p.NG - 1), it & 1);
      tc::fence_after_sync();
      if (pr) PROF(half, it, 4);
      for (int s = half; s < C3 / KS; s += 2) {
        float v[16];
        tc::tmem_ld16(lane_base + col[p.NG - 1] + s * KS, v);
        if (p.ns >= 32) {
#pragma unroll
          for (int c = 0; c < 16; ++c) v[c] = warp_max_nan(v[c]);
          if (lane == 0) {
            float4 *dst = reinterpret_cast<float4 *>(pool + (warp & 3) * MAXC + s * KS);
#pragma unroll
            for (int c4 = 0; c4 < 4; ++c4) dst[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
          }
        } else {                                                      // 16-row lists: two entries per warp
#pragma unroll
          for (int c = 0; c < 16; ++c) {
#pragma unroll
            for (int o = 8; o >= 1; o >>= 1) v[c] = max_nan(v[c], __shfl_xor_sync(0xffffffffu, v[c], o));
          }
          if ((lane & 15) == 0) {
            float4 *dst = reinterpret_cast<float4 *>(pool + (row / 16) * MAXC + s * KS);
#pragma unroll
            for (int c4 = 0; c4 < 4; ++c4) dst[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
          }
        }
      }
      if (pr) PROF(half, it, 5);
      tc::fence_before_sync();
      tc::named_barrier(1, ROW_THREADS);
      if (pr) PROF(half, it, 6);
      // combine the entries of each list, bias + ReLU, write (B, C3, m, G)
      const int epl = p.ns >= 32 ? p.ns / 32 : 1;                     // entries per list
      for (int c = tid; c < C3; c += ROW_THREADS) {
        const float bias = biases[(p.NG - 1) * MAXC + c];
        for (int l = 0; l < lpt; ++l) {
          const long long ql = (long long)tile * lpt + l;
          if (ql >= p.Q) break;
          float u = pool[(l * epl) * MAXC + c];
          for (int e = 1; e < epl; ++e) u = max_nan(u, pool[(l * epl + e) * MAXC + c]);
          const int gi = (int)(ql % p.G);
          const long long t = ql / p.G;
          const int seed = (int)(t % p.m), bi = (int)(t / p.m);
          const float y = relu_keep_nan(__fadd_rn(u, bias));
          p.out[(((long long)bi * C3 + c) * p.m + seed) * p.G + gi] = y;
          if (p.out_t) p.out_t[ql * C3 + c] = y;
        }
      }
      if (pr) PROF(half, it, 7);
    }
  } else if (warp == 8) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      uint32_t na = 0, nw = 0;
      for (int tile = blockIdx.x, it = 0; tile < p.tiles; tile += gridDim.x, ++it) {
        (void)it;
        for (int layer = 0; layer < p.NG; ++layer) {
          const int K = p.Kd[layer], N = p.Nd[layer];
          const uint32_t d = tmem + col[layer];
          const uint32_t idesc = tc::idesc_tf32_m128(N);
          const uint32_t b_lbo = (uint32_t)N * 16u, b_part = (uint32_t)N * KS * 4u;
          for (int s = 0; s < K / KS; ++s) {
            const uint32_t ws = nw % NWS, as = na % NA;
            tc::mbar_wait(sm + Smem::bars + (2 * NA + ws) * 8, (nw / NWS) & 1u);
            tc::mbar_wait(sm + Smem::bars + as * 8, (na / NA) & 1u);
            tc::fence_after_sync();
            if (s == 0 && layer >= p.NG - 2) PROF(2, it, 2 * (layer - (p.NG - 2)) + 0);          // first slab of the last two layers ready
            if (s == K / KS - 1 && layer >= p.NG - 2) PROF(2, it, 2 * (layer - (p.NG - 2)) + 1);  // last slab ready
            const uint32_t a_base = sm + Smem::a_ring + as * A_SLOT, w_base = sm + Smem::w_ring + ws * W_SLOT;
#pragma unroll
            for (int kk = 0; kk < KS / 8; ++kk) {
              const uint64_t ah = tc::smem_desc(a_base + kk * 2 * (TM * 16), TM * 16, 128);
              const uint64_t al = tc::smem_desc(a_base + A_PART + kk * 2 * (TM * 16), TM * 16, 128);
              const uint64_t bh = tc::smem_desc(w_base + kk * 2 * b_lbo, b_lbo, 128);
              const uint64_t bl = tc::smem_desc(w_base + b_part + kk * 2 * b_lbo, b_lbo, 128);
              const uint32_t acc = (s > 0 || kk > 0) ? 1u : 0u;
              if (p.passes == 3) {
                tc::mma_tf32(d, al, bh, idesc, acc);
                tc::mma_tf32(d, ah, bl, idesc, 1u);
                tc::mma_tf32(d, ah, bh, idesc, 1u);
              } else {
                tc::mma_tf32(d, ah, bh, idesc, acc);
              }
            }
            tc::mma_commit(sm + Smem::bars + (NA + as) * 8);              // A slot free once these MMAs have read it
            tc::mma_commit(sm + Smem::bars + (2 * NA + NWS + ws) * 8);    // weight slot too
            ++na; ++nw;
          }
          tc::mma_commit(acc_full + 8 * layer);
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ weight loader
    if (lane == 0) {
      uint32_t nw = 0;
      for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
        for (int layer = 0; layer < p.NG; ++layer) {
          const int K = p.Kd[layer], N = p.Nd[layer];
          const float *src = p.wp[layer];
          const uint32_t slab_floats = (uint32_t)N * KS * 2u;
          const uint32_t bytes = (uint32_t)N * KS * 4u * (p.passes == 3 ? 2u : 1u);
          for (int s = 0; s < K / KS; ++s) {
            const uint32_t ws = nw % NWS;
            tc::mbar_wait(sm + Smem::bars + (2 * NA + NWS + ws) * 8, ((nw / NWS) & 1u) ^ 1u);
            const uint32_t full = sm + Smem::bars + (2 * NA + ws) * 8;
            tc::mbar_arrive_expect_tx(full, bytes);
            tc::bulk_g2s(sm + Smem::w_ring + ws * W_SLOT, src + (size_t)s * slab_floats, bytes, full);
            ++nw;
          }
        }
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 8) {
    __syncwarp();
    tc::tmem_dealloc_512(tmem);
  }
}

// (cout, cin) fp32 weights -> per K-slab s of 16 input channels: [hi | lo][k-chunk kc (4)][n (cout)][e (4)],
// k = 16 s + 4 kc + e; hi = tf32(w), lo = tf32(w - hi)
__global__ void mlp_pack_kernel(const float *__restrict__ w, int cout, int cin, float *__restrict__ packed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cout * cin) return;
  const int n = i / cin, k = i % cin;
  const int s = k / KS, kc = (k % KS) / 4, e = k % 4;
  const float v = __ldg(w + i);
  const float hi = tc::tf32_rna(v), lo = tc::tf32_rna(v - hi);
  const size_t base = (size_t)s * cout * KS * 2;
  packed[base + ((size_t)kc * cout + n) * 4 + e] = hi;
  packed[base + (size_t)cout * KS + ((size_t)kc * cout + n) * 4 + e] = lo;
}

int launch(Params &p, cudaStream_t stream) {
  static B200DeviceOnce attr_set;
  const int dev = b200_device();
  if (attr_set.need(dev)) {
    B200_CUDA(cudaFuncSetAttribute(group_mlp_maxpool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem::total));
    attr_set.mark(dev);
  }
  const int grid = p.tiles < b200_sm_count() ? p.tiles : b200_sm_count();
  group_mlp_maxpool_kernel<<<grid, NTHREADS, Smem::total, stream>>>(p);
  B200_LAUNCH_CHECK();
  return 0;
}

bool width_ok(int c) { return c >= 16 && c <= MAXC && c % 16 == 0; }

}  // namespace

#ifdef MLP_PROFILE
extern "C" int b200_mlp_read_profile(unsigned long long *host) {
  return (int)cudaMemcpyFromSymbol(host, g_mlp_prof, sizeof(g_mlp_prof));
}
#endif

extern "C" int b200_mlp_pack_weights(const float *w, int cout, int cin, float *packed, void *stream) {
  if (!w || !packed || !width_ok(cout) || cin < KS || cin % KS || cin > 2 * MAXC) return B200_EINVAL;
  const int n = cout * cin;
  mlp_pack_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(w, cout, cin, packed);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_group_mlp_maxpool(const float *xyz, const float *new_xyz, const float *rot, const int32_t *idx,
                                      int g, int b, int n, int m, int ns, float radius, const float *w1,
                                      const float *b1, const float *w2_packed, const float *b2,
                                      const float *w3_packed, const float *b3, int c1, int c2, int c3, int passes,
                                      float *out, float *out_t, void *stream) {
  if (g < 1 || b < 0 || n < 1 || m < 0 || ns < 1) return B200_EINVAL;
  if (passes != 1 && passes != 3) return B200_EINVAL;
  // shapes the tensor-core path covers: whole lists per tile, K in slabs of 16, N a multiple of 16 up to 256
  if (ns != 16 && ns != 32 && ns != 64 && ns != 128) return B200_ECAPACITY;
  if (!width_ok(c1) || !width_ok(c2) || !width_ok(c3) || c2 + c3 > 512) return B200_ECAPACITY;
  const long long lists = (long long)b * m * g;
  if (lists == 0) return 0;
  if (!xyz || !new_xyz || !idx || !w1 || !b1 || !w2_packed || !b2 || !w3_packed || !b3 || !out) return B200_EINVAL;
  const long long tiles = ceil_div_ll(lists * ns, TM);
  if (tiles > 0x7fffffffLL) return B200_ECAPACITY;
  Params p = {};
  p.xyz = xyz; p.new_xyz = new_xyz; p.rot = rot; p.idx = idx;
  p.G = g; p.B = b; p.N = n; p.m = m; p.ns = ns; p.radius = radius;
  p.w1 = w1; p.b1 = b1; p.C0 = c1;
  p.NG = 2;
  p.wp[0] = w2_packed; p.bias[0] = b2; p.Kd[0] = c1; p.Nd[0] = c2;
  p.wp[1] = w3_packed; p.bias[1] = b3; p.Kd[1] = c2; p.Nd[1] = c3;
  p.passes = passes; p.Q = lists; p.tiles = (int)tiles; p.out = out; p.out_t = out_t;
  return launch(p, (cudaStream_t)stream);
}

extern "C" int b200_group_feat_mlp_maxpool(const float *xyz, const float *new_xyz, const float *feat_t,
                                           const int32_t *idx, int b, int n, int m, int ns, int c, float radius,
                                           const float *w1_packed, const float *b1, const float *w2_packed,
                                           const float *b2, const float *w3_packed, const float *b3, int c1, int c2,
                                           int c3, int passes, float *out, float *out_t, void *stream) {
  if (b < 0 || n < 1 || m < 0 || ns < 1 || c < 1) return B200_EINVAL;
  if (passes != 1 && passes != 3) return B200_EINVAL;
  if (ns != 16 && ns != 32 && ns != 64 && ns != 128) return B200_ECAPACITY;
  if (c % KS || c > 2 * MAXC - KS) return B200_ECAPACITY;
  if (!width_ok(c1) || !width_ok(c2) || !width_ok(c3) || c1 + c2 + c3 > 512) return B200_ECAPACITY;
  const long long lists = (long long)b * m;
  if (lists == 0) return 0;
  if (!xyz || !new_xyz || !feat_t || !idx || !w1_packed || !b1 || !w2_packed || !b2 || !w3_packed || !b3 || !out)
    return B200_EINVAL;
  const long long tiles = ceil_div_ll(lists * ns, TM);
  if (tiles > 0x7fffffffLL) return B200_ECAPACITY;
  Params p = {};
  p.xyz = xyz; p.new_xyz = new_xyz; p.rot = nullptr; p.idx = idx;
  p.G = 1; p.B = b; p.N = n; p.m = m; p.ns = ns; p.radius = radius;
  p.feat_t = feat_t; p.C = c;
  p.NG = 3;
  p.wp[0] = w1_packed; p.bias[0] = b1; p.Kd[0] = c + KS; p.Nd[0] = c1;
  p.wp[1] = w2_packed; p.bias[1] = b2; p.Kd[1] = c1; p.Nd[1] = c2;
  p.wp[2] = w3_packed; p.bias[2] = b3; p.Kd[2] = c2; p.Nd[2] = c3;
  p.passes = passes; p.Q = lists; p.tiles = (int)tiles; p.out = out; p.out_t = out_t;
  return launch(p, (cudaStream_t)stream);
}
