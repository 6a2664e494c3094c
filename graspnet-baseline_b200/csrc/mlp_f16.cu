// mlp_f16.cu -- the shared MLPs of GraspNet on the 5th-generation tensor cores, fp32-grade results from fp16 operands
// ("fp16x3"), one sm_100a kernel family for
//   (1) group -> centre / normalise / rotate -> 3 x (1x1 conv + BN + ReLU) -> max-pool over the neighbourhood
//       (SA1..SA4 of the backbone: reference pointnet2/pointnet2_modules.py:232-272 over pointnet2_utils.py:313-372;
//       CloudCrop: models/modules.py:118-142 over pointnet2_utils.py:488-555; SharedMLP = pytorch_utils.py:11-36),
//   (2) stacks of 1x1 convolutions over point columns (PointnetFPModule's SharedMLP pointnet2_modules.py:409-414,
//       ApproachNet models/modules.py:52-57, OperationNet :170-177, ToleranceNet :215-222), eval mode, BN folded.
// It supersedes mlp_fused.cu (3xTF32, kept as the `tf32x3` arithmetic) and pointwise.cu (fp32 SIMT) on the inference
// path.  What the measurements of round 2 said about mlp_fused.cu (profiles/r02_mlp_timeline.txt): a CloudCrop tile
// took 14.9k cycles for 7.7k cycles of MMAs because (a) the layer-3 epilogue, the pooling combine and the next
// tile's first layer-2 epilogue slab ran one after the other with the tensor core idle, and (b) the 320 KB of 3xTF32
// weights were re-streamed from L2 for every 128-row tile (42 B/clk/SM = the L2 cap).  Here:
//
// * Arithmetic: x = s (hi + lo), hi = fp16(x / s), lo = fp16(x / s - hi), s a power of two; a.b ~= lo.hi + hi.lo +
//   hi.hi on tcgen05.mma kind::f16 (K = 16 per instruction, 128 cycles at N = 256: twice the tf32 rate, measured with
//   tools/cu/mma_rate.cu), fp32 accumulation in TMEM.  hi + lo carries 22 significand bits + rounding: the
//   representation error is 2^-24 relative, the dropped lo.lo term 2^-24 -- fp32-grade, like 3xTF32, with HALF the
//   accumulation steps (the tensor core truncates at each step).  fp16's range is handled with exact power-of-two
//   scales: weights per layer (host, from max |w|), activations per WARP of 32 rows from a RIGOROUS bound
//   (|y_c| <= max_k |x_k| * max_c sum_k |w_ck| + max |b|, propagated layer to layer from the actual row maxima of the
//   kernel's input), so nothing can overflow and every value keeps an absolute error <= 2^-25 of the bound; scales
//   are undone exactly in the epilogue FMA.  fp16 subnormals are honoured by the tensor core (mma_rate.cu probe).
// * Weights are RESIDENT in shared memory when they fit (SA1, CloudCrop: 160 KB as hi | lo fp16) -- loaded once per
//   CTA; otherwise streamed through a ring as before.
// * Two row GROUPS (warps 0-3, 4-7; thread = row = TMEM lane) take alternate tiles: while one group pools tile t
//   (layer-3 epilogue + combine + output), the other has already pushed tile t+1 through layer 1/2 and staged its
//   first layer-3 slabs, so the tensor core restarts the moment the accumulator is free.
// * K-slabs are 32 channels (16 KB of A: hi | lo), half as many barrier round trips per channel.
//
// Operand layout in shared memory (A and B alike, "K-major, no swizzle" UMMA operands): for each 16-byte K-chunk
// (8 fp16) the rows are contiguous: byte offset = chunk * rows * 16 + row * 16.
#include <cuda_fp16.h>

#include "tc_common.cuh"

namespace {

constexpr int TM = 128;                        // rows per tile = TMEM lanes = threads of a row group
constexpr int KS = 32;                         // channels per K-slab
constexpr int A_PART = TM * KS * 2;            // 8 KB: the hi (or lo) part of an A slab
constexpr int A_SLOT = 2 * A_PART;
constexpr int MAXC = 256;                      // widest layer (one MMA: N <= 256)
constexpr int MAXL = 3;
constexpr int MAXE = 8;                        // pooling entries per tile
constexpr int MAX_WSLOTS = 40;                 // weight slabs resident at most (or ring slots)
constexpr int NGRP = 2;                        // row groups: slab `seq` of the global sequence is produced by group seq % 2
constexpr int ROWS = NGRP * TM;                // row threads
constexpr int MMA_WARP = 4 * NGRP;             // then the weight-loader warp
constexpr int NTHREADS = NGRP * TM + 64;
constexpr int SMEM_LIMIT = 227 * 1024;

enum { SRC_XYZ = 0, SRC_FEAT = 1, SRC_DENSE = 2 };
enum { SINK_POOL = 0, SINK_DENSE = 1 };

struct Layer {
  const __half *wp;                            // packed weights (mlp_f16_pack_kernel)
  const float *bias;                           // [N] zero padded
  int K, N;                                    // padded: K % 32 == 0, N % 16 == 0, 16 <= N <= 256
  int cout;                                    // real output channels
  int relu;
  int wexp;                                    // the weights were multiplied by 2^wexp before the split
  float wnorm, bmax;                           // max_c sum_k |w_ck|, max_c |b_c|
  uint32_t col;                                // first TMEM column of the accumulator
  float *y;                                    // dense sink / intermediate output (B, *, n) or NULL
  long long ybs;                               // its batch stride (floats)
};

struct Params {
  int src, sink;
  const float *xyz, *new_xyz, *rot;
  const int32_t *idx;
  int G, B, N, m, ns, log_ns;
  float radius;
  const float *feat_t;
  int C;
  const float *w1, *b1;                        // SRC_XYZ: the K = 3 first layer on the CUDA cores, (C0, 3), (C0)
  int C0;
  float w1norm, b1max;
  const float *x1, *x2;                        // SRC_DENSE: (B, c1, n) [| (B, c2, n)]
  int c1, c2, n, tiles_per_batch;
  int NG;
  Layer L[MAXL];                               // head 0 (the only one of the grouped modules)
  Layer L1[MAXL];                              // pointwise stacks: head 1 (same shapes, own weights / outputs)
  int heads;
  int S[MAXL];                                 // A slabs per tile feeding layer l (= L[l].K / 32)
  int P;                                       // slabs per tile
  long long Q;
  int tiles;
  float *out, *out_t;
  // shared memory geometry
  int pool_entries;                            // pooling entries per tile (one per 32 rows, or per 16-row list)
  int NA;                                      // A ring slots
  int w_resident, NWS;                         // resident: NWS = slabs of all layers, each at its own offset
  uint32_t w_slot;                             // ring slot bytes
  uint32_t woff[MAXL];                         // resident: byte offset of layer l's first slab in the weight region
  int wslab0[MAXL];                            // resident: index of layer l's first slab (its w_full barrier)
  int pipelined;                               // two tensor-core layers, layer-2 accumulators double-buffered
  int alias;                                   // layer 2's accumulator shares layer 0's columns
  uint32_t off_w, off_consts, off_pool, off_bars, off_holder;
};

// barriers (8 bytes each) from off_bars: a_full[NA], a_empty[NA], w_full[MAX_WSLOTS], w_empty[MAX_WSLOTS],
// acc_full[MAXL], acc_free
__device__ __forceinline__ uint32_t bar_a_full(const Params &p, uint32_t sm, int s) { return sm + p.off_bars + 8 * s; }
__device__ __forceinline__ uint32_t bar_a_empty(const Params &p, uint32_t sm, int s) { return sm + p.off_bars + 8 * (p.NA + s); }
__device__ __forceinline__ uint32_t bar_w_full(const Params &p, uint32_t sm, int s) { return sm + p.off_bars + 8 * (2 * p.NA + s); }
__device__ __forceinline__ uint32_t bar_w_empty(const Params &p, uint32_t sm, int s) {
  return sm + p.off_bars + 8 * (2 * p.NA + MAX_WSLOTS + s);
}
__device__ __forceinline__ uint32_t bar_acc_full(const Params &p, uint32_t sm, int l) {
  return sm + p.off_bars + 8 * (2 * p.NA + 2 * MAX_WSLOTS + l);
}
__device__ __forceinline__ uint32_t bar_acc_free(const Params &p, uint32_t sm) {
  return sm + p.off_bars + 8 * (2 * p.NA + 2 * MAX_WSLOTS + MAXL);
}
constexpr int NBARS_FIXED = 2 * MAX_WSLOTS + MAXL + 1;

#ifdef MLP_PROFILE
// timeline of CTA 0 (tools/mlp_prof.py): clock64 stamps [row group | MMA issuer][tile iteration][event]
constexpr int PROF_TILES = 16, PROF_EV = 8;
__device__ unsigned long long g_mlp_prof[3][PROF_TILES][PROF_EV];
#define PROF(role, it, ev)                                                                   \
  do {                                                                                       \
    if (blockIdx.x == 0 && (it) < PROF_TILES) g_mlp_prof[role][it][ev] = (unsigned long long)clock64(); \
  } while (0)
// inside one slab (thread 0, slabs SLAB_SEQ and SLAB_SEQ + 2): entry, TMEM read, math, slot wait, stores, proxy fence, arrive
__device__ unsigned long long g_slab_prof[2][8];
constexpr uint32_t SLAB_SEQ = 24;
#define SPROF(seq, ev)                                                                      \
  do {                                                                                      \
    if (blockIdx.x == 0 && threadIdx.x == 0 && ((seq) == SLAB_SEQ || (seq) == SLAB_SEQ + 2)) \
      g_slab_prof[((seq) - SLAB_SEQ) >> 1][ev] = (unsigned long long)clock64();             \
  } while (0)
#else
#define PROF(role, it, ev) do {} while (0)
#define SPROF(seq, ev) do {} while (0)
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
__device__ __forceinline__ float pow2f(int e) {                      // 2^e, -126 <= e <= 127
  return __int_as_float((e + 127) << 23);
}
// Exponent s with bound * 2^s < 2^14 (fp16 overflows at 65504 = 2^16 - 32), the same for all lanes of the warp:
// the largest bound of the warp decides.  bound >= 0 or NaN; NaN / inf rows end up NaN / inf (as in the reference).
__device__ __forceinline__ int warp_scale_exp(float bound) {
  unsigned bits = __float_as_uint(bound) & 0x7fffffffu;
  bits = __reduce_max_sync(0xffffffffu, bits);
  const int e = (int)(bits >> 23) - 127;                            // bound < 2^(e+1)
  int s = 13 - e;
  s = s > 60 ? 60 : s;
  return s < -60 ? -60 : s;
}

// 32 fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,"
      "%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

__host__ __device__ constexpr uint32_t idesc_f16_m128(int n) {      // D fp32, A and B fp16, both K-major, M = 128
  return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
}
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}

// One 32-channel slab of this row (already multiplied by the warp's scale) into ring slot seq % NA: waits for the
// slot, splits hi / lo, publishes.  hi = fp16(v) (round to nearest even), lo = fp16(v - hi).
__device__ __forceinline__ void put_slab(const Params &p, uint32_t sm, const float (&v)[32], int row, uint32_t seq) {
  const uint32_t slot = seq % (uint32_t)p.NA, ph = (seq / (uint32_t)p.NA) & 1u;
  SPROF(seq, 3);
  tc::mbar_wait(bar_a_empty(p, sm, slot), ph ^ 1u);
  SPROF(seq, 4);
  const uint32_t base = sm + slot * A_SLOT + row * 16;
#pragma unroll
  for (int kc = 0; kc < 4; ++kc) {
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float a = v[8 * kc + 2 * e], b = v[8 * kc + 2 * e + 1];
      const __half2 h = __floats2half2_rn(a, b);
      const float2 hf = __half22float2(h);
      const __half2 l = __floats2half2_rn(__fsub_rn(a, hf.x), __fsub_rn(b, hf.y));
      hi[e] = *reinterpret_cast<const uint32_t *>(&h);
      lo[e] = *reinterpret_cast<const uint32_t *>(&l);
    }
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(base + kc * (TM * 16)), "r"(hi[0]), "r"(hi[1]),
                 "r"(hi[2]), "r"(hi[3]) : "memory");
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(base + A_PART + kc * (TM * 16)), "r"(lo[0]),
                 "r"(lo[1]), "r"(lo[2]), "r"(lo[3]) : "memory");
  }
  SPROF(seq, 5);
  tc::fence_proxy_async();
  SPROF(seq, 6);
  tc::fence_before_sync();
  tc::mbar_arrive(bar_a_full(p, sm, slot));
  SPROF(seq, 7);
}

// One instantiation per (source, schedule): the row threads' code must stay small -- at 67 KB of SASS every warp ran
// at ~5 cycles per instruction on instruction fetch (profiles/r02_mlp_timeline.txt), so every helper below has ONE
// call site and only its own source / sink variant.
template <int SRC, bool PIPE>
__global__ void __launch_bounds__(NTHREADS, 1) mlp_f16_kernel(const Params p) {
  constexpr int SINK = SRC == SRC_DENSE ? SINK_DENSE : SINK_POOL;
  const Layer *LL = (SRC == SRC_DENSE && blockIdx.y == 1) ? p.L1 : p.L;   // this CTA's head
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t sm = tc::smem_u32(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float4 *w1s = reinterpret_cast<float4 *>(smem_raw + p.off_consts);
  float *biases = reinterpret_cast<float *>(w1s + p.C0);              // [layer][MAXC]
  float *pools = reinterpret_cast<float *>(smem_raw + p.off_pool);    // [tile parity][entries][MAXC] raw maxima, then
  float *pool_inv = pools + 2 * p.pool_entries * MAXC;                // [tile parity][entries] the entries' 1 / scale
  volatile uint32_t *holder = reinterpret_cast<volatile uint32_t *>(smem_raw + p.off_holder);

  if (SRC == SRC_XYZ)
    for (int i = tid; i < p.C0; i += NTHREADS)
      w1s[i] = make_float4(__ldg(p.w1 + 3 * i), __ldg(p.w1 + 3 * i + 1), __ldg(p.w1 + 3 * i + 2), __ldg(p.b1 + i));
  for (int l = 0; l < p.NG; ++l)
    for (int i = tid; i < LL[l].N; i += NTHREADS) biases[l * MAXC + i] = __ldg(LL[l].bias + i);
  if (tid == 0) {
    for (int s = 0; s < p.NA; ++s) {
      tc::mbar_init(bar_a_full(p, sm, s), TM);                       // the 128 threads of one row group
      tc::mbar_init(bar_a_empty(p, sm, s), 1);                       // one tcgen05.commit
    }
    for (int s = 0; s < p.NWS; ++s) {
      tc::mbar_init(bar_w_full(p, sm, s), 1);                        // expect_tx arrival + bytes
      tc::mbar_init(bar_w_empty(p, sm, s), 1);                       // one tcgen05.commit
    }
    for (int l = 0; l < MAXL; ++l) tc::mbar_init(bar_acc_full(p, sm, l), 1);
    tc::mbar_init(bar_acc_free(p, sm), ROWS);
    tc::fence_barrier_init();
  }
  if (warp == MMA_WARP) tc::tmem_alloc_512(sm + p.off_holder);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *holder;
  const int last = p.NG - 1;
  const int ntile = ((int)blockIdx.x < p.tiles) ? (p.tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;   // this CTA's tiles
  const int PS = p.S[1] < 2 ? p.S[1] : 2;                             // pipelined schedule: layer-3 slabs staged ahead

  if (warp < 4 * NGRP) {
    // ------------------------------------------------------------------ row threads
    // Both groups walk the same global slab sequence; slab `seq` is produced by group seq % 2 (so a thread's
    // previous slab is seq - 2 and every slot it waits for is at most one occupant behind: the mbarrier parity
    // waits stay meaningful).
    const int g = warp >> 2;
    const int row = (warp & 3) * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const int lpt = TM >> p.log_ns;                                   // lists per tile (pooled sinks)
    const bool pr = lane == 0 && (warp & 3) == 0;
    (void)pr;
    uint32_t seq = 0;

    // ---- the rows of a tile.  Grouped sources gather in three stages so that the two dependent global round trips
    // hide under other work: (A) the neighbour's index, (B) the point, the centre and the rotation, (C) the arithmetic.
    struct {
      bool valid;
      int bi, seed, i, j;
      float pt[3], ct[3], R[9];
      float g0, g1, g2;
      const float4 *frow;
      const float *r1, *r2;
    } gt;
    gt.valid = false; gt.bi = gt.seed = gt.i = gt.j = 0; gt.g0 = gt.g1 = gt.g2 = 0.f;
    gt.frow = nullptr; gt.r1 = gt.r2 = nullptr;
    auto gather_a = [&](int it_) {
      gt.valid = false;
      if (it_ >= ntile || SRC == SRC_DENSE) return;
      const int tile_ = blockIdx.x + it_ * gridDim.x;
      const unsigned q = (unsigned)tile_ * (unsigned)lpt + (unsigned)(row >> p.log_ns);
      gt.valid = (long long)q < p.Q;
      if (gt.valid) {
        const int gi = (int)(q % (unsigned)p.G);
        const unsigned t = q / (unsigned)p.G;
        gt.seed = (int)(t % (unsigned)p.m);
        gt.bi = (int)(t / (unsigned)p.m);
        gt.i = __ldg(p.idx + ((((long long)gi * p.B + gt.bi) * p.m + gt.seed) * p.ns + (row & (p.ns - 1))));
      }
    };
    auto gather_b = [&]() {
      if (!gt.valid || SRC == SRC_DENSE) return;
      const float *pt = p.xyz + ((long long)gt.bi * p.N + gt.i) * 3;
      const float *ct = p.new_xyz + ((long long)gt.bi * p.m + gt.seed) * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) { gt.pt[c] = __ldg(pt + c); gt.ct[c] = __ldg(ct + c); }
      if (p.rot) {
        const float *R = p.rot + ((long long)gt.bi * p.m + gt.seed) * 9;
#pragma unroll
        for (int c = 0; c < 9; ++c) gt.R[c] = __ldg(R + c);
      }
    };
    // dense rows: batch, column, validity of this thread's row of tile it_
    auto dense_row = [&](int it_, int &bi, int &j, bool &valid) {
      const int tile_ = blockIdx.x + it_ * gridDim.x;
      bi = tile_ / p.tiles_per_batch;
      j = (tile_ - bi * p.tiles_per_batch) * TM + row;
      valid = j < p.n;
    };
    // 16 channels of a dense row from ch0 on: whole blocks inside x1 or x2 are 16 unconditional loads
    auto load16 = [&](int ch0, float (&t)[16]) {
      const int cin = p.c1 + p.c2;
      if (ch0 + 16 <= p.c1) {
        const float *q = gt.r1 + (long long)ch0 * p.n;
#pragma unroll
        for (int c = 0; c < 16; ++c) t[c] = __ldg(q + (long long)c * p.n);
      } else if (ch0 >= p.c1 && ch0 + 16 <= cin) {
        const float *q = gt.r2 + (long long)(ch0 - p.c1) * p.n;
#pragma unroll
        for (int c = 0; c < 16; ++c) t[c] = __ldg(q + (long long)c * p.n);
      } else {
#pragma unroll 1
        for (int c = 0; c < 16; ++c) {
          const int ch = ch0 + c;
          const float x = ch >= cin ? 0.f : (ch < p.c1 ? __ldg(gt.r1 + (long long)ch * p.n) : __ldg(gt.r2 + (long long)(ch - p.c1) * p.n));
#pragma unroll
          for (int e = 0; e < 16; ++e)
            if (e == c) t[e] = x;
        }
      }
    };
    // stage (C) / the dense first pass: this row's inputs of tile it_ and the bound on their magnitude
    auto prepare = [&](int it_, float &bound0) {
      if (SRC == SRC_DENSE) {
        bool valid;
        dense_row(it_, gt.bi, gt.j, valid);
        gt.valid = valid;
        const int jj = valid ? gt.j : 0;                              // invalid rows read column 0 and are never written
        gt.r1 = p.x1 + (long long)gt.bi * p.c1 * p.n + jj;
        gt.r2 = p.x2 ? p.x2 + (long long)gt.bi * p.c2 * p.n + jj : gt.r1;
        const int cin = p.c1 + p.c2;
        // the row's largest |x| (first pass): the groups take alternate blocks of 32 channels (32 loads in flight
        // per thread) and meet in shared memory
        float amax = 0.f;
        for (int c0 = 32 * g; c0 < cin; c0 += 64) {
          float t[16], u[16];
          load16(c0, t);
          if (c0 + 16 < cin) {
            load16(c0 + 16, u);
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) u[c] = 0.f;
          }
#pragma unroll
          for (int c = 0; c < 16; ++c) amax = max_nan(amax, max_nan(fabsf(t[c]), fabsf(u[c])));
        }
        float *xchg = pools + (it_ & 1) * ROWS;
        xchg[g * TM + row] = amax;
        tc::named_barrier(1, ROWS);
        amax = max_nan(amax, xchg[(g ^ 1) * TM + row]);
        bound0 = valid ? amax : 0.f;
        return;
      }
      gt.g0 = gt.g1 = gt.g2 = 0.f;
      gt.frow = reinterpret_cast<const float4 *>(p.feat_t);           // invalid rows: any finite data, scaled by 0
      if (gt.valid) {
        if (SRC == SRC_FEAT) gt.frow = reinterpret_cast<const float4 *>(p.feat_t + ((long long)gt.bi * p.N + gt.i) * p.C);
        float d0 = __fsub_rn(gt.pt[0], gt.ct[0]), d1 = __fsub_rn(gt.pt[1], gt.ct[1]), d2 = __fsub_rn(gt.pt[2], gt.ct[2]);
        if (p.radius > 0.f) {
          d0 = __fdiv_rn(d0, p.radius); d1 = __fdiv_rn(d1, p.radius); d2 = __fdiv_rn(d2, p.radius);
        }
        if (p.rot) {                                                  // row vector times the seed's 3x3
          gt.g0 = __fmaf_rn(d2, gt.R[6], __fmaf_rn(d1, gt.R[3], __fmul_rn(d0, gt.R[0])));
          gt.g1 = __fmaf_rn(d2, gt.R[7], __fmaf_rn(d1, gt.R[4], __fmul_rn(d0, gt.R[1])));
          gt.g2 = __fmaf_rn(d2, gt.R[8], __fmaf_rn(d1, gt.R[5], __fmul_rn(d0, gt.R[2])));
        } else {
          gt.g0 = d0; gt.g1 = d1; gt.g2 = d2;
        }
      }
      const float gmax = max_nan(max_nan(fabsf(gt.g0), fabsf(gt.g1)), fabsf(gt.g2));
      if (SRC == SRC_FEAT) {
        // the row's largest |feature| first (a second pass over the same 16-byte loads, from L1/L2)
        const int nb = p.C >> 4;
        float amax = gmax;
        for (int b16 = 0; b16 < nb; ++b16) {
          float4 f[4];
#pragma unroll
          for (int c = 0; c < 4; ++c) f[c] = __ldg(gt.frow + 4 * b16 + c);
#pragma unroll
          for (int c = 0; c < 4; ++c)
            amax = max_nan(max_nan(amax, max_nan(fabsf(f[c].x), fabsf(f[c].y))), max_nan(fabsf(f[c].z), fabsf(f[c].w)));
        }
        bound0 = gt.valid ? amax : 0.f;
      } else {
        // the K = 3 layer on the CUDA cores: |y_c| <= max|g| * max_c (|w_c0| + |w_c1| + |w_c2|) + max |b|
        bound0 = __fmaf_rn(gmax, p.w1norm, p.b1max);
      }
    };
    // the A slabs of the first tensor-core layer (inner loops per block of 16 channels with no per-channel
    // conditions: a runtime test per element costs a convergence barrier each)
    auto emit_source = [&](int sexp0) {
      const float sc = (SRC == SRC_XYZ || gt.valid) ? pow2f(sexp0) : 0.f;
      if (SRC == SRC_DENSE) {
        // this group's slabs (every other one), the next slab's 32 loads in flight while the current one is split
        // and stored
        const int cin = p.c1 + p.c2;
        auto load32 = [&](int s, float (&t)[32]) {
          float a[16], b[16];
          load16(s * KS, a);
          if (s * KS + 16 < cin) {
            load16(s * KS + 16, b);
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) b[c] = 0.f;
          }
#pragma unroll
          for (int c = 0; c < 16; ++c) { t[c] = a[c]; t[16 + c] = b[c]; }
        };
        const int S0 = p.S[0];
        int s = (int)((seq ^ (uint32_t)g) & 1u);                      // first slab with (seq + s) % 2 == g
        float nxt[32];
        if (s < S0) load32(s, nxt);
#pragma unroll 1
        for (; s < S0; s += 2) {
          float v[32];
#pragma unroll
          for (int c = 0; c < 32; ++c) v[c] = __fmul_rn(nxt[c], sc);
          if (s + 2 < S0) load32(s + 2, nxt);
          put_slab(p, sm, v, row, seq + (uint32_t)s);
        }
        seq += (uint32_t)S0;
        return;
      }
      for (int s = 0; s < p.S[0]; ++s, ++seq) {
        if ((int)(seq & 1u) != g) continue;
        float v[32];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int ch0 = s * KS + 16 * h;
          if (SRC == SRC_FEAT) {
            // K order of layer 1: [features (C) | x y z | zero pad], C % 16 == 0
            const int b16 = ch0 >> 4, nb = p.C >> 4;
            float4 f[4];
            if (b16 < nb) {
#pragma unroll
              for (int c = 0; c < 4; ++c) f[c] = __ldg(gt.frow + 4 * b16 + c);
            } else {
#pragma unroll
              for (int c = 0; c < 4; ++c) f[c] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (b16 == nb) f[0] = make_float4(gt.g0, gt.g1, gt.g2, 0.f);
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              v[16 * h + 4 * c] = __fmul_rn(f[c].x, sc); v[16 * h + 4 * c + 1] = __fmul_rn(f[c].y, sc);
              v[16 * h + 4 * c + 2] = __fmul_rn(f[c].z, sc); v[16 * h + 4 * c + 3] = __fmul_rn(f[c].w, sc);
            }
          } else if (ch0 < p.C0) {                                    // C0 % 16 == 0
#pragma unroll
            for (int c = 0; c < 16; ++c) {
              const float4 w = w1s[ch0 + c];
              const float a = __fmaf_rn(w.z, gt.g2, __fmaf_rn(w.y, gt.g1, __fmul_rn(w.x, gt.g0)));
              v[16 * h + c] = __fmul_rn(max_nan(__fadd_rn(a, w.w), 0.f), sc);
            }
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) v[16 * h + c] = 0.f;
          }
        }
        put_slab(p, sm, v, row, seq);
      }
    };
    // slabs [s0, s1) of the epilogue of tensor-core layer l (< last) of tile it_ -> K-slabs of layer l + 1.
    // sexp_in: scale of the layer's input; nexp: scale of its output; tcol: its accumulator.
    auto emit_mid = [&](int l, int it_, int s0, int s1, int sexp_in, int nexp, uint32_t tcol) {
      const Layer &L = LL[l];
      const float inv = pow2f(-sexp_in - L.wexp), sc = pow2f(nexp);
      const float floor_v = L.relu ? 0.f : __int_as_float(0xff800000);     // max(y, floor): ReLU, or the identity
      const int N = L.N, cout = L.cout;
      const float4 *bl4 = reinterpret_cast<const float4 *>(biases + l * MAXC);
      float *yrow = nullptr;
      if (L.y) {
        int bi, j; bool valid;
        dense_row(it_, bi, j, valid);
        if (valid) yrow = L.y + (long long)bi * L.ybs + j;
      }
      for (int s = s0; s < s1; ++s, ++seq) {
        if ((int)(seq & 1u) != g) continue;
        float v[32];
        SPROF(seq, 0);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int ch0 = s * KS + 16 * h;
          if (ch0 < N) {
            float a[16];
            tc::tmem_ld16(tcol + ch0, a);
            if (h == 0) SPROF(seq, 1);
#pragma unroll
            for (int c4 = 0; c4 < 4; ++c4) {
              const float4 b = bl4[(ch0 >> 2) + c4];
              a[4 * c4] = max_nan(__fmaf_rn(a[4 * c4], inv, b.x), floor_v);
              a[4 * c4 + 1] = max_nan(__fmaf_rn(a[4 * c4 + 1], inv, b.y), floor_v);
              a[4 * c4 + 2] = max_nan(__fmaf_rn(a[4 * c4 + 2], inv, b.z), floor_v);
              a[4 * c4 + 3] = max_nan(__fmaf_rn(a[4 * c4 + 3], inv, b.w), floor_v);
            }
            if (yrow) {
              float *q = yrow + (long long)ch0 * p.n;
              if (ch0 + 16 <= cout) {
#pragma unroll
                for (int c = 0; c < 16; ++c) q[(long long)c * p.n] = a[c];
              } else {
#pragma unroll
                for (int c = 0; c < 16; ++c)
                  if (ch0 + c < cout) q[(long long)c * p.n] = a[c];
              }
            }
#pragma unroll
            for (int c = 0; c < 16; ++c) v[16 * h + c] = __fmul_rn(a[c], sc);
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) v[16 * h + c] = 0.f;
          }
        }
        SPROF(seq, 2);
        put_slab(p, sm, v, row, seq);
      }
    };
    // the last layer of tile it_ (its accumulator is complete): pooled maxima or the dense output; the two groups
    // take alternate blocks of 16 columns.  Ends with the arrival that lets the tensor core overwrite the accumulator.
    auto last_layer = [&](int it_, int sexp_in) {
      const Layer &L = LL[last];
      const float inv = pow2f(-sexp_in - L.wexp);
      const float floor_v = L.relu ? 0.f : __int_as_float(0xff800000);
      const int N = L.N, cout = L.cout;
      const uint32_t tcol = lane_base + L.col;
      if (SINK == SINK_DENSE) {
        int bi, j; bool valid;
        dense_row(it_, bi, j, valid);
        float *yrow = valid ? L.y + (long long)bi * L.ybs + j : nullptr;
        const float4 *bl4 = reinterpret_cast<const float4 *>(biases + last * MAXC);
        for (int ch0 = 16 * g; ch0 < N; ch0 += 32) {
          float a[16];
          tc::tmem_ld16(tcol + ch0, a);
          if (yrow) {
#pragma unroll
            for (int c4 = 0; c4 < 4; ++c4) {
              const float4 b = bl4[(ch0 >> 2) + c4];
              a[4 * c4] = max_nan(__fmaf_rn(a[4 * c4], inv, b.x), floor_v);
              a[4 * c4 + 1] = max_nan(__fmaf_rn(a[4 * c4 + 1], inv, b.y), floor_v);
              a[4 * c4 + 2] = max_nan(__fmaf_rn(a[4 * c4 + 2], inv, b.z), floor_v);
              a[4 * c4 + 3] = max_nan(__fmaf_rn(a[4 * c4 + 3], inv, b.w), floor_v);
            }
            float *q = yrow + (long long)ch0 * p.n;
            if (ch0 + 16 <= cout) {
#pragma unroll
              for (int c = 0; c < 16; ++c) q[(long long)c * p.n] = a[c];
            } else {
#pragma unroll
              for (int c = 0; c < 16; ++c)
                if (ch0 + c < cout) q[(long long)c * p.n] = a[c];
            }
          }
        }
      } else {
        // max of the RAW accumulators over the rows of a pooling entry (a warp's 32 rows, or 16 of them); the
        // entry's scale (a positive power of two: max commutes with it) is undone in the combine
        float *pool = pools + (it_ & 1) * (p.pool_entries * MAXC);
        float *pinv = pool_inv + (it_ & 1) * p.pool_entries;
        if (p.ns >= 32) {
          if (lane == 0 && g == 0) pinv[warp & 3] = inv;
          for (int ch0 = 16 * g; ch0 < N; ch0 += 32) {
            float v[16];
            tc::tmem_ld16(tcol + ch0, v);
            float mine = 0.f;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
              const float r = warp_max_nan(v[c]);
              if (lane == c) mine = r;
            }
            if (lane < 16) pool[(warp & 3) * MAXC + ch0 + lane] = mine;
          }
        } else {                                                      // 16-row lists: two entries per warp
          if ((lane & 15) == 0 && g == 0) pinv[row >> 4] = inv;
          for (int ch0 = 16 * g; ch0 < N; ch0 += 32) {
            float v[16];
            tc::tmem_ld16(tcol + ch0, v);
            float mine = 0.f;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
              float r = v[c];
#pragma unroll
              for (int o = 8; o >= 1; o >>= 1) r = max_nan(r, __shfl_xor_sync(0xffffffffu, r, o));
              if ((lane & 15) == c) mine = r;
            }
            pool[(row >> 4) * MAXC + ch0 + (lane & 15)] = mine;
          }
        }
      }
      tc::fence_before_sync();
      tc::mbar_arrive(bar_acc_free(p, sm));
    };
    // pooled sinks: combine the entries of each list of tile it_, unscale, bias + ReLU, write (B, C3, m, G) [and
    // channels-last].  Entered through a barrier of all row threads (the pool is complete); the pool of this tile
    // parity is rewritten two tiles later, after the next tile's barrier.
    auto combine = [&](int it_) {
      if (SINK != SINK_POOL) return;
      tc::named_barrier(1, ROWS);
      const Layer &L = LL[last];
      const float floor_v = L.relu ? 0.f : __int_as_float(0xff800000);
      const int cout = L.cout;
      const float *bl = biases + last * MAXC;
      const float *pool = pools + (it_ & 1) * (p.pool_entries * MAXC);
      const float *pinv = pool_inv + (it_ & 1) * p.pool_entries;
      const int tile_ = blockIdx.x + it_ * gridDim.x;
      const int epl = p.ns >= 32 ? p.ns >> 5 : 1;                     // entries per list
      const unsigned q0 = (unsigned)tile_ * (unsigned)lpt;
      int lgi = (int)(q0 % (unsigned)p.G);
      const unsigned t0 = q0 / (unsigned)p.G;
      int lseed = (int)(t0 % (unsigned)p.m), lbi = (int)(t0 / (unsigned)p.m);
      for (int l = 0; l < lpt && (long long)q0 + l < p.Q; ++l) {
        for (int c = g * TM + row; c < cout; c += ROWS) {
          float u = __fmul_rn(pool[(l * epl) * MAXC + c], pinv[l * epl]);
          for (int e = 1; e < epl; ++e) u = max_nan(u, __fmul_rn(pool[(l * epl + e) * MAXC + c], pinv[l * epl + e]));
          const float y = max_nan(__fadd_rn(u, bl[c]), floor_v);
          p.out[(((long long)lbi * cout + c) * p.m + lseed) * p.G + lgi] = y;
          if (p.out_t) p.out_t[((long long)q0 + l) * cout + c] = y;
        }
        if (++lgi == p.G) {
          lgi = 0;
          if (++lseed == p.m) { lseed = 0; ++lbi; }
        }
      }
    };

    if (PIPE) {
      // ---- two tensor-core layers, layer-2 accumulators double-buffered (tile parity): software pipeline
      //   prologue: S(0), S(1), E2(0)[0, PS)
      //   tile t:   E2(t)[PS, S1) | E2(t+1)[0, PS) | pool(t) | S(t+2) | combine(t)
      // so that layer 3 of tile t + 1 restarts on staged slabs the moment pool(t) has read the accumulator.
      const Layer &L0 = LL[0];
      float b0[3] = {0.f, 0.f, 0.f};                                  // bound of the source slabs of tiles t, t+1, t+2
      int e0[3] = {0, 0, 0};                                          // and their scale exponents
      gather_a(0); gather_b();
      // t = -2, -1 are the prologue: only the parts that exist run (S(0) | E2(0)[0,PS), S(1))
#pragma unroll 1
      for (int t = -2; t < ntile; ++t) {
        if (pr && t >= 0) PROF(g, t, 0);
        int n1 = 0;
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {                                 // E2(t)[PS, S1), then E2(t+1)[0, PS)
          const int tt = t + h;
          if (tt < 0 || tt >= ntile) continue;
          const int nx = warp_scale_exp(__fmaf_rn(b0[h], L0.wnorm, L0.bmax));
          if (h == 0) n1 = nx;
          if (h == 1) {
            tc::mbar_wait(bar_acc_full(p, sm, 0), (uint32_t)tt & 1u);
            tc::fence_after_sync();
            if (pr && t >= 0) PROF(g, t, 2);
          }
          emit_mid(0, tt, h == 0 ? PS : 0, h == 0 ? p.S[1] : PS, e0[h], nx,
                   lane_base + L0.col + (uint32_t)((tt & 1) * L0.N));
          if (h == 0) {
            if (pr && t >= 0) PROF(g, t, 1);
          }
        }
        gather_b();                                                   // tile t + 2: stage (B)
        if (pr && t >= 0) PROF(g, t, 3);
        if (t >= 0) {
          tc::mbar_wait(bar_acc_full(p, sm, 1), (uint32_t)t & 1u);
          tc::fence_after_sync();
          if (pr) PROF(g, t, 4);
          last_layer(t, n1);
          if (pr) PROF(g, t, 5);
        }
        if (t + 2 < ntile) {
          prepare(t + 2, b0[2]); e0[2] = warp_scale_exp(b0[2]);
          emit_source(e0[2]);
        }
        if (pr && t >= 0) PROF(g, t, 6);
        gather_a(t + 3);
        if (t >= 0) combine(t);
        if (pr && t >= 0) PROF(g, t, 7);
        b0[0] = b0[1]; b0[1] = b0[2]; e0[0] = e0[1]; e0[1] = e0[2];
      }
    } else {
      // ---- any depth, one accumulator per layer:  S(0);  tile t:  E_0(t) .. E_{last-1}(t) | S(t+1) | last(t) | combine(t)
      // (with aliased accumulators -- layer 2 in layer 0's columns -- S(t+1) follows last(t))
      float bn = 0.f;
      int en = 0;
      gather_a(0); gather_b();
      // t = -1 is the prologue (S(0) only)
#pragma unroll 1
      for (int t = -1; t < ntile; ++t) {
        float bound = bn;
        int sexp = en;
        if (pr && t >= 0) PROF(g, t, 0);
        if (t >= 0) {
#pragma unroll 1
          for (int l = 0; l < last; ++l) {
            const Layer &L = LL[l];
            const float nbound = __fmaf_rn(bound, L.wnorm, L.bmax);
            const int nexp = warp_scale_exp(nbound);
            tc::mbar_wait(bar_acc_full(p, sm, l), (uint32_t)t & 1u);
            tc::fence_after_sync();
            if (pr && l == last - 1) PROF(g, t, 1);
            emit_mid(l, t, 0, p.S[l + 1], sexp, nexp, lane_base + L.col);
            bound = nbound; sexp = nexp;
          }
        }
        if (pr && t >= 0) PROF(g, t, 2);
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {                                 // S(t+1) before last(t), or after it (aliased)
          if (h == (p.alias ? 1 : 0) && t + 1 < ntile) {
            prepare(t + 1, bn); en = warp_scale_exp(bn);
            emit_source(en);
            gather_a(t + 2); gather_b();
          }
          if (h == 0 && t >= 0) {
            if (pr) PROF(g, t, 3);
            tc::mbar_wait(bar_acc_full(p, sm, last), (uint32_t)t & 1u);
            tc::fence_after_sync();
            if (pr) PROF(g, t, 4);
            last_layer(t, sexp);
            if (pr) PROF(g, t, 5);
          }
        }
        if (t >= 0) combine(t);
        if (pr && t >= 0) PROF(g, t, 7);
      }
    }
  } else if (warp == MMA_WARP) {
    // ------------------------------------------------------------------ MMA issuer: the same slab sequence
    if (lane == 0) {
      uint32_t na = 0, nw = 0;
      // all slabs [s0, s1) of layer l into the accumulator at column dcol; `first`: slab s0 starts the accumulation
      auto run = [&](int l, int s0, int s1, uint32_t dcol, int it_) {
        const Layer &L = LL[l];
        const uint32_t d = tmem + dcol;
        const uint32_t idesc = idesc_f16_m128(L.N);
        const uint32_t b_lbo = (uint32_t)L.N * 16u, b_part = (uint32_t)L.N * 64u, w_slab = (uint32_t)L.N * 128u;
        for (int s = s0; s < s1; ++s) {
          uint32_t ws, w_base;
          if (p.w_resident) {
            ws = (uint32_t)p.wslab0[l] + (uint32_t)s;                 // every slab has its own slot
            w_base = sm + p.off_w + p.woff[l] + (uint32_t)s * w_slab;
            tc::mbar_wait(bar_w_full(p, sm, ws), 0u);
          } else {
            ws = nw % (uint32_t)p.NWS;
            w_base = sm + p.off_w + ws * p.w_slot;
            tc::mbar_wait(bar_w_full(p, sm, ws), (nw / (uint32_t)p.NWS) & 1u);
          }
          const uint32_t as = na % (uint32_t)p.NA;
          tc::mbar_wait(bar_a_full(p, sm, as), (na / (uint32_t)p.NA) & 1u);
          tc::fence_after_sync();
          if (s == 0 && l >= p.NG - 2) PROF(NGRP, it_, 2 * (l - (p.NG - 2)));
          if (s == p.S[l] - 1 && l >= p.NG - 2) PROF(NGRP, it_, 2 * (l - (p.NG - 2)) + 1);
          const uint32_t a_base = sm + as * A_SLOT;
#pragma unroll
          for (int kk = 0; kk < KS / 16; ++kk) {
            const uint64_t ah = tc::smem_desc(a_base + kk * 2 * (TM * 16), TM * 16, 128);
            const uint64_t al = tc::smem_desc(a_base + A_PART + kk * 2 * (TM * 16), TM * 16, 128);
            const uint64_t bh = tc::smem_desc(w_base + kk * 2 * b_lbo, b_lbo, 128);
            const uint64_t bl = tc::smem_desc(w_base + b_part + kk * 2 * b_lbo, b_lbo, 128);
            const uint32_t acc = (s > 0 || kk > 0) ? 1u : 0u;
            mma_f16(d, al, bh, idesc, acc);
            mma_f16(d, ah, bl, idesc, 1u);
            mma_f16(d, ah, bh, idesc, 1u);
          }
          tc::mma_commit(bar_a_empty(p, sm, as));                    // A slot free once these MMAs have read it
          if (!p.w_resident) tc::mma_commit(bar_w_empty(p, sm, ws));
          ++na; ++nw;
        }
      };
      if (PIPE) {
        // the row threads' order:  t = -2: S(0);  t = -1: E2(0)[0,PS), S(1);  t >= 0: E2(t)[PS,S1), E2(t+1)[0,PS), S(t+2)
        const Layer &L0 = LL[0], &L1 = LL[1];
        for (int t = -2; t < ntile; ++t) {
          if (t >= 0) {
            run(1, PS, p.S[1], L1.col, t);
            tc::mma_commit(bar_acc_full(p, sm, 1));
          }
          if (t + 1 >= 0 && t + 1 < ntile) {
            if (t >= 0) {
              tc::mbar_wait(bar_acc_free(p, sm), (uint32_t)t & 1u);  // pool(t) has read the accumulator
              tc::fence_after_sync();
            }
            run(1, 0, PS, L1.col, t + 1);
          }
          if (t + 2 < ntile) {
            run(0, 0, p.S[0], L0.col + (uint32_t)(((t + 2) & 1) * L0.N), t + 2);
            tc::mma_commit(bar_acc_full(p, sm, 0));
          }
        }
      } else {
        for (int t = 0; t < ntile; ++t) {
          for (int l = 0; l < p.NG; ++l) {
            if (t > 0 && l == (p.alias ? 0 : last)) {                 // last(t - 1) has read the accumulator
              tc::mbar_wait(bar_acc_free(p, sm), (uint32_t)(t - 1) & 1u);
              tc::fence_after_sync();
            }
            run(l, 0, p.S[l], LL[l].col, t);
            tc::mma_commit(bar_acc_full(p, sm, l));
          }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ weight loader
    if (lane == 0) {
      uint32_t nw = 0;
      const int rounds = p.w_resident ? (ntile > 0 ? 1 : 0) : ntile;
      for (int it = 0; it < rounds; ++it) {
        for (int l = 0; l < p.NG; ++l) {
          const Layer &L = LL[l];
          const uint32_t bytes = (uint32_t)L.N * 128u;
          const uint8_t *src = reinterpret_cast<const uint8_t *>(L.wp);
          for (int s = 0; s < p.S[l]; ++s) {
            uint32_t ws, dst;
            if (p.w_resident) {
              ws = nw;
              dst = sm + p.off_w + p.woff[l] + (uint32_t)s * bytes;
            } else {
              ws = nw % (uint32_t)p.NWS;
              dst = sm + p.off_w + ws * p.w_slot;
              tc::mbar_wait(bar_w_empty(p, sm, ws), ((nw / (uint32_t)p.NWS) & 1u) ^ 1u);
            }
            const uint32_t full = bar_w_full(p, sm, ws);
            tc::mbar_arrive_expect_tx(full, bytes);
            tc::bulk_g2s(dst, src + (size_t)s * bytes, bytes, full);
            ++nw;
          }
        }
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == MMA_WARP) {
    __syncwarp();
    tc::tmem_dealloc_512(tmem);
  }
}

// (cout, cin) fp32 weights, row stride ldw -> per K-slab s of 32 input channels:
// [hi | lo][k-chunk kc (4)][n (n_pad)][e (8)] fp16 of w * 2^wexp, k = 32 s + 8 kc + e; rows >= cout and columns >= cin
// are zero.  `perm_c` > 0 permutes the input columns [x y z | C features] -> [features | x y z] (SRC_FEAT K order).
__global__ void mlp_f16_pack_kernel(const float *__restrict__ w, int cout, int cin, int ldw, int n_pad, int k_pad,
                                    int wexp, int perm_c, __half *__restrict__ packed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pad * k_pad) return;
  const int n = i / k_pad, k = i % k_pad;
  float v = 0.f;
  if (n < cout && k < cin) {
    int src = k;
    if (perm_c > 0) src = k < perm_c ? k + 3 : k - perm_c;
    v = __ldg(w + (size_t)n * ldw + src) * pow2f(wexp);
  }
  const __half hi = __float2half_rn(v);
  const __half lo = __float2half_rn(v - __half2float(hi));
  const int s = k / KS, kc = (k % KS) / 8, e = k % 8;
  const size_t base = (size_t)s * n_pad * KS * 2;
  packed[base + ((size_t)kc * n_pad + n) * 8 + e] = hi;
  packed[base + (size_t)n_pad * KS + ((size_t)kc * n_pad + n) * 8 + e] = lo;
}

int ilog2(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}

int fill_layers(Params &p, const b200_mlp_layer *layers, int nl) {
  if (nl < 1 || nl > MAXL || !layers) return B200_EINVAL;
  p.NG = nl;
  int P = 0;
  for (int l = 0; l < nl; ++l) {
    const b200_mlp_layer &h = layers[l];
    if (!h.w_packed || !h.bias) return B200_EINVAL;
    if (h.k_pad < KS || h.k_pad % KS || h.n_pad < 16 || h.n_pad % 16 || h.n_pad > MAXC || h.cout < 1 || h.cout > h.n_pad)
      return B200_ECAPACITY;
    if (l > 0 && h.k_pad != (layers[l - 1].n_pad + KS - 1) / KS * KS) return B200_EINVAL;
    if (h.w_exp < -40 || h.w_exp > 40) return B200_ECAPACITY;
    Layer &L = p.L[l];
    L.wp = reinterpret_cast<const __half *>(h.w_packed); L.bias = h.bias; L.K = h.k_pad; L.N = h.n_pad;
    L.cout = h.cout; L.relu = h.relu; L.wexp = h.w_exp; L.wnorm = h.w_norm1; L.bmax = h.b_max;
    L.y = h.y; L.ybs = h.y_batch_stride;
    p.S[l] = h.k_pad / KS;
    P += p.S[l];
  }
  p.P = P;
  // TMEM columns: consecutive if they fit, else layer 1 at the top and layers 0 / 2 share the bottom (layer 2 starts
  // only after layer 0's accumulator has been read completely)
  int total = 0;
  for (int l = 0; l < nl; ++l) total += p.L[l].N;
  if (total <= 512) {
    uint32_t c = 0;
    for (int l = 0; l < nl; ++l) { p.L[l].col = c; c += (uint32_t)p.L[l].N; }
  } else {
    if (nl != 3 || p.L[0].N + p.L[1].N > 512 || p.L[1].N + p.L[2].N > 512) return B200_ECAPACITY;
    p.L[0].col = 0; p.L[2].col = 0; p.L[1].col = 512u - (uint32_t)p.L[1].N;
    p.alias = 1;
  }
  return 0;
}

int plan_smem(Params &p) {
  // weights resident if they leave room for >= 3 A slots; else rings of 4 + 3 (A / W)
  uint32_t wbytes = 0, wmax = 0;
  int slabs = 0;
  for (int l = 0; l < p.NG; ++l) {
    p.woff[l] = wbytes;
    p.wslab0[l] = slabs;
    wbytes += (uint32_t)p.S[l] * (uint32_t)p.L[l].N * 128u;
    wmax = wmax > (uint32_t)p.L[l].N * 128u ? wmax : (uint32_t)p.L[l].N * 128u;
    slabs += p.S[l];
  }
  const uint32_t consts = ((uint32_t)p.C0 * 4 + (uint32_t)p.NG * MAXC) * 4;
  p.pool_entries = p.sink == SINK_POOL ? TM / (p.ns < 32 ? p.ns : 32) : 0;
  const uint32_t pool = p.sink == SINK_POOL ? 2u * (uint32_t)p.pool_entries * (MAXC + 1) * 4 : 2u * ROWS * 4;
  const uint32_t fixed = consts + pool + (2 * 8 + NBARS_FIXED) * 8 + 16 + 128;
  int na = 0;
  if (slabs <= MAX_WSLOTS && fixed + wbytes + 3 * A_SLOT <= (uint32_t)SMEM_LIMIT) {
    p.w_resident = 1; p.NWS = slabs; p.w_slot = 0;
    na = (int)((SMEM_LIMIT - fixed - wbytes) / A_SLOT);
    if (na > 8) na = 8;
  } else {
    p.w_resident = 0; p.w_slot = wmax;
    na = 4;
    int nws = (int)((SMEM_LIMIT - fixed - (uint32_t)na * A_SLOT) / wmax);
    if (nws > 6) nws = 6;
    if (nws < 2) return B200_ECAPACITY;
    p.NWS = nws;
    wbytes = (uint32_t)nws * wmax;
  }
  p.NA = na;
  p.off_w = (uint32_t)na * A_SLOT;
  p.off_consts = p.off_w + ((wbytes + 127u) & ~127u);
  p.off_pool = p.off_consts + consts;
  p.off_bars = p.off_pool + pool;
  p.off_holder = p.off_bars + (2 * (uint32_t)na + NBARS_FIXED) * 8;
  return (int)(p.off_holder + 16);
}

int launch(Params &p, cudaStream_t stream) {
  const int smem = plan_smem(p);
  if (smem < 0) return smem;
  // the software-pipelined schedule needs resident weights (its MMA order differs from the loader's stream order)
  // and room for a second layer-2 accumulator
  p.pipelined = 0;
  if (p.NG == 2 && p.w_resident && p.src == SRC_XYZ && 2 * p.L[0].N + p.L[1].N <= 512) {
    p.pipelined = 1;
    p.L[0].col = 0;
    p.L[1].col = 2u * (uint32_t)p.L[0].N;
  }
  static B200DeviceOnce attr_set;
  const int dev = b200_device();
  if (attr_set.need(dev)) {
    B200_CUDA(cudaFuncSetAttribute(mlp_f16_kernel<SRC_XYZ, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200_CUDA(cudaFuncSetAttribute(mlp_f16_kernel<SRC_XYZ, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200_CUDA(cudaFuncSetAttribute(mlp_f16_kernel<SRC_FEAT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200_CUDA(cudaFuncSetAttribute(mlp_f16_kernel<SRC_DENSE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    attr_set.mark(dev);
  }
  const int grid = p.tiles < b200_sm_count() ? p.tiles : b200_sm_count();
  if (p.src == SRC_XYZ && p.pipelined) mlp_f16_kernel<SRC_XYZ, true><<<grid, NTHREADS, smem, stream>>>(p);
  else if (p.src == SRC_XYZ) mlp_f16_kernel<SRC_XYZ, false><<<grid, NTHREADS, smem, stream>>>(p);
  else if (p.src == SRC_FEAT) mlp_f16_kernel<SRC_FEAT, false><<<grid, NTHREADS, smem, stream>>>(p);
  else mlp_f16_kernel<SRC_DENSE, false><<<dim3(grid, p.heads > 1 ? 2 : 1), NTHREADS, smem, stream>>>(p);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // namespace

#ifdef MLP_PROFILE
extern "C" int b200_mlp_f16_read_profile(unsigned long long *host) {
  return (int)cudaMemcpyFromSymbol(host, g_mlp_prof, sizeof(g_mlp_prof));
}
extern "C" int b200_mlp_f16_read_slab_profile(unsigned long long *host) {
  return (int)cudaMemcpyFromSymbol(host, g_slab_prof, sizeof(g_slab_prof));
}
#endif

extern "C" size_t b200_mlp_f16_packed_bytes(int n_pad, int k_pad) { return (size_t)n_pad * k_pad * 4; }

extern "C" int b200_mlp_f16_pack(const float *w, int cout, int cin, int ldw, int n_pad, int k_pad, int w_exp,
                                 int feat_channels, void *packed, void *stream) {
  if (!w || !packed || cout < 1 || cin < 1 || ldw < cin || n_pad < cout || k_pad < cin) return B200_EINVAL;
  if (n_pad % 16 || n_pad > MAXC || k_pad % KS || w_exp < -40 || w_exp > 40) return B200_ECAPACITY;
  if (feat_channels < 0 || (feat_channels > 0 && feat_channels + 3 != cin)) return B200_EINVAL;
  const int n = n_pad * k_pad;
  mlp_f16_pack_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(w, cout, cin, ldw, n_pad, k_pad, w_exp,
                                                                         feat_channels, (__half *)packed);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_group_mlp_maxpool_f16(const float *xyz, const float *new_xyz, const float *rot, const int32_t *idx,
                                          int g, int b, int n, int m, int ns, float radius, const float *w1,
                                          const float *b1, int c1, float w1_norm1, float b1_max,
                                          const b200_mlp_layer *layers, float *out, float *out_t, void *stream) {
  if (g < 1 || b < 0 || n < 1 || m < 0 || ns < 1) return B200_EINVAL;
  if (ns != 16 && ns != 32 && ns != 64 && ns != 128) return B200_ECAPACITY;
  if (c1 < 1 || c1 > MAXC) return B200_ECAPACITY;
  const long long lists = (long long)b * m * g;
  if (lists == 0) return 0;
  if (!xyz || !new_xyz || !idx || !w1 || !b1 || !out) return B200_EINVAL;
  const long long tiles = ceil_div_ll(lists * ns, TM);
  if (lists > 0x7fffffffLL || tiles > 0x7fffffffLL) return B200_ECAPACITY;
  Params p = {};
  p.src = SRC_XYZ; p.sink = SINK_POOL;
  p.xyz = xyz; p.new_xyz = new_xyz; p.rot = rot; p.idx = idx;
  p.G = g; p.B = b; p.N = n; p.m = m; p.ns = ns; p.log_ns = ilog2(ns); p.radius = radius;
  p.w1 = w1; p.b1 = b1; p.C0 = c1; p.w1norm = w1_norm1; p.b1max = b1_max;
  const int rc = fill_layers(p, layers, 2);
  if (rc) return rc;
  if (p.L[0].K != (c1 + KS - 1) / KS * KS || p.L[1].cout != p.L[1].N) return B200_EINVAL;
  p.Q = lists; p.tiles = (int)tiles; p.out = out; p.out_t = out_t;
  return launch(p, (cudaStream_t)stream);
}

extern "C" int b200_group_feat_mlp_maxpool_f16(const float *xyz, const float *new_xyz, const float *feat_t,
                                               const int32_t *idx, int b, int n, int m, int ns, int c, float radius,
                                               const b200_mlp_layer *layers, float *out, float *out_t, void *stream) {
  if (b < 0 || n < 1 || m < 0 || ns < 1 || c < 4) return B200_EINVAL;
  if (ns != 16 && ns != 32 && ns != 64 && ns != 128) return B200_ECAPACITY;
  if (c % 4) return B200_ECAPACITY;
  const long long lists = (long long)b * m;
  if (lists == 0) return 0;
  if (!xyz || !new_xyz || !feat_t || !idx || !out) return B200_EINVAL;
  if ((reinterpret_cast<uintptr_t>(feat_t) & 15u) != 0) return B200_EINVAL;
  const long long tiles = ceil_div_ll(lists * ns, TM);
  if (lists > 0x7fffffffLL || tiles > 0x7fffffffLL) return B200_ECAPACITY;
  Params p = {};
  p.src = SRC_FEAT; p.sink = SINK_POOL;
  p.xyz = xyz; p.new_xyz = new_xyz; p.rot = nullptr; p.idx = idx;
  p.G = 1; p.B = b; p.N = n; p.m = m; p.ns = ns; p.log_ns = ilog2(ns); p.radius = radius;
  p.feat_t = feat_t; p.C = c;
  const int rc = fill_layers(p, layers, 3);
  if (rc) return rc;
  if (p.L[0].K != (c + 3 + KS - 1) / KS * KS || p.L[2].cout != p.L[2].N) return B200_EINVAL;
  p.Q = lists; p.tiles = (int)tiles; p.out = out; p.out_t = out_t;
  return launch(p, (cudaStream_t)stream);
}

extern "C" int b200_pointwise_stack_f16(const float *x1, int c1, const float *x2, int c2, int b, int n,
                                        const b200_mlp_layer *layers, int nlayers, int nheads, void *stream) {
  if (b < 0 || n < 0 || c1 < 1 || c2 < 0 || nheads < 1 || nheads > 2) return B200_EINVAL;
  if (b == 0 || n == 0) return 0;
  if (!x1 || (c2 > 0 && !x2)) return B200_EINVAL;
  Params p = {};
  p.src = SRC_DENSE; p.sink = SINK_DENSE;
  p.x1 = x1; p.x2 = c2 > 0 ? x2 : nullptr; p.c1 = c1; p.c2 = c2; p.n = n; p.B = b;
  p.ns = 1; p.log_ns = 0; p.G = 1; p.m = 1;
  p.heads = nheads;
  if (nheads == 2) {                                                  // head 1 first (fill_layers works on p.L)
    const int rc1 = fill_layers(p, layers + nlayers, nlayers);
    if (rc1) return rc1;
    for (int l = 0; l < nlayers; ++l) p.L1[l] = p.L[l];
  }
  const int rc = fill_layers(p, layers, nlayers);
  if (rc) return rc;
  if (p.L[0].K != (c1 + c2 + KS - 1) / KS * KS) return B200_EINVAL;
  for (int l = 0; l < nlayers; ++l) {
    if (nheads == 2 && (p.L1[l].K != p.L[l].K || p.L1[l].N != p.L[l].N)) return B200_EINVAL;   // same shapes
    if (l == nlayers - 1 && (!p.L[l].y || (nheads == 2 && !p.L1[l].y))) return B200_EINVAL;
  }
  p.tiles_per_batch = ceil_div(n, TM);
  const long long tiles = (long long)p.tiles_per_batch * b;
  if (tiles > 0x7fffffffLL) return B200_ECAPACITY;
  p.tiles = (int)tiles; p.Q = tiles * TM;
  return launch(p, (cudaStream_t)stream);
}
