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
constexpr int NQ = 2;                          // groups ("quarters") of 128 row threads; slab `seq` is produced by the
                                               // NQ / 2 quarters of pair seq % 2, each a 32 / (NQ / 2)-channel part
constexpr int HALVES = NQ / 2;
constexpr int CH = 32 / HALVES;                // channels of a slab per thread
constexpr int ROWS = NQ * TM;                  // row threads
constexpr int MMA_WARP = 4 * NQ;               // then the weight-loader warp
constexpr int NTHREADS = ROWS + 64;
static_assert(ROWS / 32 == 8, "the row-bound rider holds one partial maximum per row warp: 8");
constexpr int SMEM_LIMIT = 227 * 1024;

enum { SRC_XYZ = 0, SRC_FEAT = 1, SRC_DENSE = 2 };
enum { SINK_POOL = 0, SINK_DENSE = 1 };

struct FastDiv {                                                      // n / d for n < 2^31: (umulhi(mul, n) + n) >> shr
  uint32_t mul, shr;
};
__device__ __forceinline__ uint32_t fdiv(uint32_t n, FastDiv f) { return (__umulhi(f.mul, n) + n) >> f.shr; }

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
  FastDiv fd_G, fd_m;                          // list number -> (batch, seed, depth slice) without divisions
  int tiles;
  float *out, *out_t;
  float *out_rb;                               // [list][8]: per-warp partial maxima of |y| over the list's channels, or NULL
  const float *feat_rb;                        // SRC_FEAT: the same rider of feat_t (row bound without scanning the row), or NULL
  // shared memory geometry
  int pool_entries;                            // pooling entries per tile (one per 32 rows, or per 16-row list)
  int NA;                                      // A ring slots
  int NAs, NAa;                                // xyz_pipe_kernel: slots of ring S (source slabs) and ring A (epilogue slabs)
  int w_resident, NWS;                         // resident: NWS = slabs of all layers, each at its own offset
  uint32_t w_slot;                             // ring slot bytes
  uint32_t woff[MAXL];                         // resident: byte offset of layer l's first slab in the weight region
  int wslab0[MAXL];                            // resident: index of layer l's first slab (its w_full barrier)
  int pipelined;                               // two tensor-core layers, layer-2 accumulators double-buffered
  int alias;                                   // layer 2's accumulator shares layer 0's columns
  uint32_t off_w, off_consts, off_pool, off_xchg, off_bars, off_holder;
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
__device__ unsigned long long g_mlp_prof[5][PROF_TILES][PROF_EV];    // xyz_pipe_kernel roles: A0, A1, P1, P2, issuer
#define PROF(role, it, ev)                                                                   \
  do {                                                                                       \
    if (blockIdx.x == 0 && (it) < PROF_TILES) g_mlp_prof[role][it][ev] = (unsigned long long)clock64(); \
  } while (0)
// inside one slab (thread 0, slabs SLAB_SEQ and SLAB_SEQ + 2): entry, TMEM read, math, slot wait, stores, proxy fence, arrive
__device__ unsigned long long g_slab_prof[2][8];
__device__ unsigned long long g_kernel_prof[8];                        // CTA 0 thread 0: entry, set-up done, exit (clock64)
#ifndef SLAB_SEQ_VALUE
#define SLAB_SEQ_VALUE 24
#endif
constexpr uint32_t SLAB_SEQ = SLAB_SEQ_VALUE;
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
// (redux.sync over half-warp member masks is far slower than four shuffle steps: SA3's pooling went 5.4k -> 23k cycles)
// 32 bytes per lane in one request (LDG.E.256): the feature rows are gathered, every lane of a load touches a line of
// its own and the L1 takes them one line per cycle -- half as many instructions is half as many line look-ups
__device__ __forceinline__ void ldg256(const float4 *p, float4 &a, float4 &b) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
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

// The ring position of the slab sequence, advanced once per slab by every row thread (no division by the runtime
// ring length): slot = seq % NA, phase = (seq / NA) & 1.
struct Cursor {
  uint32_t seq, slot, phase;
  __device__ __forceinline__ void next(uint32_t na) {
    ++seq;
    if (++slot == na) { slot = 0; phase ^= 1u; }
  }
};

// CH channels (part `half` of a 32-channel slab) of this row, already multiplied by the warp's scale, into the ring
// slot of cursor c: waits for the slot, splits hi / lo, publishes.  hi = fp16(v) (round to nearest even),
// lo = fp16(v - hi).
__device__ __forceinline__ void put_part(const Params &p, uint32_t sm, const float (&v)[CH], int row, const Cursor &c, int half) {
  const uint32_t bars = sm + p.off_bars + 8 * c.slot;
  SPROF(c.seq, 3);
  tc::mbar_wait(bars + 8 * (uint32_t)p.NA, c.phase ^ 1u);             // a_empty[slot]
  SPROF(c.seq, 4);
  const uint32_t base = sm + c.slot * A_SLOT + (uint32_t)half * ((CH / 8) * TM * 16) + row * 16;
#pragma unroll
  for (int kc = 0; kc < CH / 8; ++kc) {
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
  SPROF(c.seq, 5);
  tc::fence_proxy_async();
  SPROF(c.seq, 6);
  tc::fence_before_sync();
  tc::mbar_arrive(bars);                                              // a_full[slot]
  SPROF(c.seq, 7);
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
#ifdef MLP_PROFILE
  if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) g_kernel_prof[0] = (unsigned long long)clock64();
#endif
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
      tc::mbar_init(bar_a_full(p, sm, s), HALVES * TM);              // the quarters that produce a slab
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
#ifdef MLP_PROFILE
  if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) g_kernel_prof[1] = (unsigned long long)clock64();
#endif
  const int last = p.NG - 1;
  const int ntile = ((int)blockIdx.x < p.tiles) ? (p.tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;   // this CTA's tiles
  const int PS = p.S[1] < 2 ? p.S[1] : 2;                             // pipelined schedule: layer-3 slabs staged ahead

  if (warp < 4 * NQ) {
    // ------------------------------------------------------------------ row threads
    // NQ QUARTERS of 128 threads (thread = row = TMEM lane; warp w may read lanes 32 (w % 4) .. + 31).  All walk the
    // same global slab sequence; slab `seq` is produced by the quarters of PAIR seq % 2, each quarter one part (CH of
    // the 32 channels).  NQ = 2 (one quarter per pair, 32 channels per thread) executes the fewest instructions --
    // the kernel is issue-bound on per-warp overheads, measured: NQ = 4 ran 24.2k warp instructions per CloudCrop
    // tile against 21.5k (profiles/r02_mlp_f16_ncu.txt).  The tensor core consumes the slabs in `seq` order, so when
    // a thread that has produced slab seq - 2 waits for the slot of slab seq, every slab up to seq - 2 - NA has been
    // consumed: the slot is at most one occupant behind and the mbarrier parity wait is meaningful (NA >= 2).
    const int q = warp >> 2, pair = q / HALVES, half = q % HALVES;
    const int row = (warp & 3) * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const int lpt = TM >> p.log_ns;                                   // lists per tile (pooled sinks)
    const bool pr = lane == 0 && (warp & 3) == 0 && half == 0;
    (void)pr;
    Cursor cur = {0u, 0u, 0u};
    const uint32_t NA = (uint32_t)p.NA;
    float *xchg = reinterpret_cast<float *>(smem_raw + p.off_xchg);   // [2][ROWS] row maxima of the quarters

    // ---- the rows of a tile.  Grouped sources gather in three stages so that the two dependent global round trips
    // hide under other work: (A) the neighbour's index, (B) the point, the centre and the rotation, (C) the arithmetic.
    // (All four quarters gather the same rows: 16 loads per thread that hit L1 after the first quarter.)
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
      const unsigned ql = (unsigned)tile_ * (unsigned)lpt + (unsigned)(row >> p.log_ns);
      gt.valid = (long long)ql < p.Q;
      if (gt.valid) {
        const unsigned t = fdiv(ql, p.fd_G);
        const int gi = (int)(ql - t * (unsigned)p.G);
        gt.bi = (int)fdiv(t, p.fd_m);
        gt.seed = (int)(t - (unsigned)gt.bi * (unsigned)p.m);
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
        const float *s_ = gt.r1 + (long long)ch0 * p.n;
#pragma unroll
        for (int c = 0; c < 16; ++c) t[c] = __ldg(s_ + (long long)c * p.n);
      } else if (ch0 >= p.c1 && ch0 + 16 <= cin) {
        const float *s_ = gt.r2 + (long long)(ch0 - p.c1) * p.n;
#pragma unroll
        for (int c = 0; c < 16; ++c) t[c] = __ldg(s_ + (long long)c * p.n);
      } else if (ch0 >= cin) {
#pragma unroll
        for (int c = 0; c < 16; ++c) t[c] = 0.f;
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
    // the largest |input| of this row: every quarter scans a quarter of the channels, they meet in shared memory
    auto row_max_exchange = [&](int it_, float amax) {
      float *x = xchg + (it_ & 1) * ROWS;
      x[q * TM + row] = amax;
      tc::named_barrier(1, ROWS);
#pragma unroll
      for (int o = 1; o < NQ; ++o) amax = max_nan(amax, x[((q + o) & (NQ - 1)) * TM + row]);
      return amax;
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
        float amax = 0.f;
        for (int c0 = 16 * q; c0 < cin; c0 += 16 * NQ) {
          float t[16];
          load16(c0, t);
#pragma unroll
          for (int c = 0; c < 16; ++c) amax = max_nan(amax, fabsf(t[c]));
        }
        amax = row_max_exchange(it_, amax);
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
        // the row's largest |feature| (a first pass over the 16-byte loads the slabs repeat, from L1 / L2)
        // (sixteen 16-byte loads in flight per step: with one block of four per iteration the pass cost one L2 round
        // trip per block, 6k cycles for 256 channels -- profiles/r02_mlp_timeline.txt)
        const int nb = p.C >> 4;
        float amax = gmax;
        if (p.feat_rb) {
          // the producer of feat_t left each row's bound beside it (eight per-warp partial maxima): 32 bytes instead
          // of a pass over the row's C channels -- the pass was 6k (C = 128) to 8.5k cycles (C = 256) per tile
          if (gt.valid) {
            float4 r0, r1;
            ldg256(reinterpret_cast<const float4 *>(p.feat_rb + ((long long)gt.bi * p.N + gt.i) * 8), r0, r1);
            amax = max_nan(max_nan(max_nan(amax, max_nan(r0.x, r0.y)), max_nan(r0.z, r0.w)),
                           max_nan(max_nan(r1.x, r1.y), max_nan(r1.z, r1.w)));
          }
          bound0 = gt.valid ? amax : 0.f;
          return;
        }
#pragma unroll 1
        for (int b0 = q; b0 < nb; b0 += 4 * NQ) {
          float4 f[4][4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int b16 = b0 + u * NQ;
            if (b16 < nb) {
              ldg256(gt.frow + 4 * b16, f[u][0], f[u][1]);
              ldg256(gt.frow + 4 * b16 + 2, f[u][2], f[u][3]);
            } else {
#pragma unroll
              for (int c = 0; c < 4; ++c) f[u][c] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int c = 0; c < 4; ++c)
              amax = max_nan(max_nan(amax, max_nan(fabsf(f[u][c].x), fabsf(f[u][c].y))),
                             max_nan(fabsf(f[u][c].z), fabsf(f[u][c].w)));
        }
        amax = row_max_exchange(it_, amax);
        bound0 = gt.valid ? amax : 0.f;
      } else {
        // the K = 3 layer on the CUDA cores: |y_c| <= max|g| * max_c (|w_c0| + |w_c1| + |w_c2|) + max |b|
        bound0 = __fmaf_rn(gmax, p.w1norm, p.b1max);
      }
    };
    // this quarter's part of the A slabs of the first tensor-core layer (blocks of 16 channels with no per-channel
    // conditions: a runtime test per element costs a convergence barrier each)
    auto emit_source = [&](int sexp0) {
      const float sc = (SRC == SRC_XYZ || gt.valid) ? pow2f(sexp0) : 0.f;
      const int S0 = p.S[0];
      if (SRC == SRC_DENSE) {
        // the pair's slabs are every other one; the next one's loads are in flight while this one is split and stored
        auto loadp = [&](int s, float (&t)[CH]) {
#pragma unroll
          for (int h16 = 0; h16 < CH / 16; ++h16) {
            float u[16];
            load16(s * KS + half * CH + 16 * h16, u);
#pragma unroll
            for (int c = 0; c < 16; ++c) t[16 * h16 + c] = u[c];
          }
        };
        int s = (int)((cur.seq ^ (uint32_t)pair) & 1u);               // first slab with (seq + s) % 2 == pair
        float nxt[CH];
        if (s < S0) loadp(s, nxt);
        if (s == 1) cur.next(NA);
#pragma unroll 1
        for (; s < S0; s += 2) {
          float v[CH];
#pragma unroll
          for (int c = 0; c < CH; ++c) v[c] = __fmul_rn(nxt[c], sc);
          if (s + 2 < S0) loadp(s + 2, nxt);
          put_part(p, sm, v, row, cur, half);
          cur.next(NA);
          if (s + 1 < S0) cur.next(NA);
        }
        return;
      }
      if (SRC == SRC_FEAT) {
        // K order of layer 1: [features (C) | x y z | zero pad], C % 16 == 0.  As for the dense source, the pair's next
        // slab is loaded while this one is split and stored (a slab cost one L2 round trip before its first store)
        const int nb = p.C >> 4;
        auto loadf = [&](int s, float4 (&f)[CH / 4]) {
#pragma unroll
          for (int h16 = 0; h16 < CH / 16; ++h16) {
            const int b16 = (s * KS + half * CH + 16 * h16) >> 4;
            if (b16 < nb) {
              ldg256(gt.frow + 4 * b16, f[4 * h16], f[4 * h16 + 1]);
              ldg256(gt.frow + 4 * b16 + 2, f[4 * h16 + 2], f[4 * h16 + 3]);
            } else {
#pragma unroll
              for (int c = 0; c < 4; ++c) f[4 * h16 + c] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (b16 == nb) f[4 * h16] = make_float4(gt.g0, gt.g1, gt.g2, 0.f);
            }
          }
        };
        int s = (int)((cur.seq ^ (uint32_t)pair) & 1u);               // first slab with (seq + s) % 2 == pair
        float4 nxt[CH / 4];
        if (s < S0) loadf(s, nxt);
        if (s == 1) cur.next(NA);
#pragma unroll 1
        for (; s < S0; s += 2) {
          float v[CH];
#pragma unroll
          for (int c = 0; c < CH / 4; ++c) {
            v[4 * c] = __fmul_rn(nxt[c].x, sc); v[4 * c + 1] = __fmul_rn(nxt[c].y, sc);
            v[4 * c + 2] = __fmul_rn(nxt[c].z, sc); v[4 * c + 3] = __fmul_rn(nxt[c].w, sc);
          }
          if (s + 2 < S0) loadf(s + 2, nxt);
          put_part(p, sm, v, row, cur, half);
          cur.next(NA);
          if (s + 1 < S0) cur.next(NA);
        }
        return;
      }
#pragma unroll 1
      for (int s = 0; s < S0; ++s, cur.next(NA)) {
        if ((int)(cur.seq & 1u) != pair) continue;
        float v[CH];
#pragma unroll
        for (int h16 = 0; h16 < CH / 16; ++h16) {
          const int ch0 = s * KS + half * CH + 16 * h16;
          if (ch0 < p.C0) {                                           // xyz source; C0 % 16 == 0
#pragma unroll
            for (int c = 0; c < 16; ++c) {
              const float4 w = w1s[ch0 + c];
              const float a = __fmaf_rn(w.z, gt.g2, __fmaf_rn(w.y, gt.g1, __fmul_rn(w.x, gt.g0)));
              v[16 * h16 + c] = __fmul_rn(max_nan(__fadd_rn(a, w.w), 0.f), sc);
            }
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) v[16 * h16 + c] = 0.f;
          }
        }
        put_part(p, sm, v, row, cur, half);
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
      if (SRC == SRC_DENSE && L.y) {
        int bi, j; bool valid;
        dense_row(it_, bi, j, valid);
        if (valid) yrow = L.y + (long long)bi * L.ybs + j;
      }
#pragma unroll 1
      for (int s = s0; s < s1; ++s, cur.next(NA)) {
        if ((int)(cur.seq & 1u) != pair) continue;
        float v[CH];
        SPROF(cur.seq, 0);
#pragma unroll
        for (int h16 = 0; h16 < CH / 16; ++h16) {
          const int ch0 = s * KS + half * CH + 16 * h16;
          if (ch0 < N) {
            float a[16];
            tc::tmem_ld16(tcol + ch0, a);
            if (h16 == 0) SPROF(cur.seq, 1);
#pragma unroll
            for (int c4 = 0; c4 < 4; ++c4) {
              const float4 b = bl4[(ch0 >> 2) + c4];
              a[4 * c4] = max_nan(__fmaf_rn(a[4 * c4], inv, b.x), floor_v);
              a[4 * c4 + 1] = max_nan(__fmaf_rn(a[4 * c4 + 1], inv, b.y), floor_v);
              a[4 * c4 + 2] = max_nan(__fmaf_rn(a[4 * c4 + 2], inv, b.z), floor_v);
              a[4 * c4 + 3] = max_nan(__fmaf_rn(a[4 * c4 + 3], inv, b.w), floor_v);
            }
            if (SRC == SRC_DENSE && yrow) {
              float *dst = yrow + (long long)ch0 * p.n;
              if (ch0 + 16 <= cout) {
#pragma unroll
                for (int c = 0; c < 16; ++c) dst[(long long)c * p.n] = a[c];
              } else {
#pragma unroll
                for (int c = 0; c < 16; ++c)
                  if (ch0 + c < cout) dst[(long long)c * p.n] = a[c];
              }
            }
#pragma unroll
            for (int c = 0; c < 16; ++c) v[16 * h16 + c] = __fmul_rn(a[c], sc);
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) v[16 * h16 + c] = 0.f;
          }
        }
        SPROF(cur.seq, 2);
        put_part(p, sm, v, row, cur, half);
      }
    };
    // the last layer of tile it_ (its accumulator is complete): pooled maxima or the dense output; the quarters take
    // every fourth block of 16 columns.  Ends with the arrival that lets the tensor core overwrite the accumulator.
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
#pragma unroll 1
        for (int ch0 = 16 * q; ch0 < N; ch0 += 16 * NQ) {
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
            float *dst = yrow + (long long)ch0 * p.n;
            if (ch0 + 16 <= cout) {
#pragma unroll
              for (int c = 0; c < 16; ++c) dst[(long long)c * p.n] = a[c];
            } else {
#pragma unroll
              for (int c = 0; c < 16; ++c)
                if (ch0 + c < cout) dst[(long long)c * p.n] = a[c];
            }
          }
        }
      } else {
        // max of the RAW accumulators over the rows of a pooling entry (a warp's 32 rows, or 16 of them); the
        // entry's scale (a positive power of two: max commutes with it) is undone in the combine
        float *pool = pools + (it_ & 1) * (p.pool_entries * MAXC);
        float *pinv = pool_inv + (it_ & 1) * p.pool_entries;
        if (p.ns >= 32) {
          if (lane == 0 && q == 0) pinv[warp & 3] = inv;
#pragma unroll 1
          for (int ch0 = 16 * q; ch0 < N; ch0 += 16 * NQ) {
            float v[16];
            tc::tmem_ld16(tcol + ch0, v);
#pragma unroll
            for (int c = 0; c < 16; ++c) v[c] = warp_max_nan(v[c]);
            if (lane == 0) {
              float4 *dst = reinterpret_cast<float4 *>(pool + (warp & 3) * MAXC + ch0);
#pragma unroll
              for (int c4 = 0; c4 < 4; ++c4) dst[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
            }
          }
        } else {                                                      // 16-row lists: two entries per warp
          if ((lane & 15) == 0 && q == 0) pinv[row >> 4] = inv;
#pragma unroll 1
          for (int ch0 = 16 * q; ch0 < N; ch0 += 16 * NQ) {
            float v[16];
            tc::tmem_ld16(tcol + ch0, v);
#pragma unroll
            for (int c = 0; c < 16; ++c) {
#pragma unroll
              for (int o = 8; o >= 1; o >>= 1) v[c] = max_nan(v[c], __shfl_xor_sync(0xffffffffu, v[c], o));
            }
            if ((lane & 15) == 0) {
              float4 *dst = reinterpret_cast<float4 *>(pool + (row >> 4) * MAXC + ch0);
#pragma unroll
              for (int c4 = 0; c4 < 4; ++c4) dst[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
            }
          }
        }
      }
      tc::fence_before_sync();
      tc::mbar_arrive(bar_acc_free(p, sm));
    };
    // pooled sinks: combine the entries of each list of tile it_, unscale, bias + ReLU, write (B, C3, m, G) [and
    // channels-last]; one (list, channel) item per thread and step.  Entered through a barrier of all row threads
    // (the pool is complete); the pool of this tile parity is rewritten two tiles later, after the next barrier.
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
      const unsigned t0 = fdiv(q0, p.fd_G);                           // the tile's first list; the others follow it
      int lgi = (int)(q0 - t0 * (unsigned)p.G), lbi = (int)fdiv(t0, p.fd_m);
      int lseed = (int)(t0 - (unsigned)lbi * (unsigned)p.m);
#pragma unroll 1
      for (int l = 0; l < lpt && (long long)q0 + l < p.Q; ++l) {
        const float *pl = pool + (l * epl) * MAXC;
        float *dst = p.out + (((long long)lbi * cout) * p.m + lseed) * p.G + lgi;
        float ym = 0.f;
#pragma unroll 1
        for (int c = q * TM + row; c < cout; c += ROWS) {
          float u = __fmul_rn(pl[c], pinv[l * epl]);
          for (int e = 1; e < epl; ++e) u = max_nan(u, __fmul_rn(pl[e * MAXC + c], pinv[l * epl + e]));
          const float y = max_nan(__fadd_rn(u, bl[c]), floor_v);
          dst[(long long)c * p.m * p.G] = y;
          if (p.out_t) p.out_t[((long long)q0 + l) * cout + c] = y;
          ym = max_nan(ym, fabsf(y));
        }
        if (p.out_rb) {                                // this warp's share of the list's row bound (see prepare())
          ym = warp_max_nan(ym);
          if (lane == 0) p.out_rb[((long long)q0 + l) * 8 + warp] = ym;
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
      float b0_0 = 0.f, b0_1 = 0.f, b0_2 = 0.f;                       // bound of the source slabs of tiles t, t+1, t+2
      int e0_0 = 0, e0_1 = 0, e0_2 = 0;                               // and their scale exponents
      gather_a(0); gather_b();
      // t = -2, -1 are the prologue: only the parts that exist run (S(0) | E2(0)[0,PS), S(1))
#pragma unroll 1
      for (int t = -2; t < ntile; ++t) {
        if (pr && t >= 0) PROF(pair, t, 0);
        int n1 = 0;
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {                                 // E2(t)[PS, S1), then E2(t+1)[0, PS)
          const int tt = t + h;
          if (tt < 0 || tt >= ntile) continue;
          const int nx = warp_scale_exp(__fmaf_rn(h ? b0_1 : b0_0, L0.wnorm, L0.bmax));
          if (h == 0) n1 = nx;
          if (h == 1) {
            tc::mbar_wait(bar_acc_full(p, sm, 0), (uint32_t)tt & 1u);
            tc::fence_after_sync();
            if (pr && t >= 0) PROF(pair, t, 2);
          }
          emit_mid(0, tt, h == 0 ? PS : 0, h == 0 ? p.S[1] : PS, h ? e0_1 : e0_0, nx,
                   lane_base + L0.col + (uint32_t)((tt & 1) * L0.N));
          if (h == 0) {
            if (pr && t >= 0) PROF(pair, t, 1);
          }
        }
        gather_b();                                                   // tile t + 2: stage (B)
        if (pr && t >= 0) PROF(pair, t, 3);
        if (t >= 0) {
          tc::mbar_wait(bar_acc_full(p, sm, 1), (uint32_t)t & 1u);
          tc::fence_after_sync();
          if (pr) PROF(pair, t, 4);
          last_layer(t, n1);
          if (pr) PROF(pair, t, 5);
        }
        if (t + 2 < ntile) {
          prepare(t + 2, b0_2); e0_2 = warp_scale_exp(b0_2);
          emit_source(e0_2);
        }
        if (pr && t >= 0) PROF(pair, t, 6);
        gather_a(t + 3);
        if (t >= 0) combine(t);
        if (pr && t >= 0) PROF(pair, t, 7);
        b0_0 = b0_1; b0_1 = b0_2; e0_0 = e0_1; e0_1 = e0_2;
      }
    } else {
      // ---- any depth, one accumulator per layer:  S(0);  tile t:  E_0(t) .. E_{last-1}(t) | S(t+1) | last(t) | combine(t)
      // (with aliased accumulators -- layer 2 in layer 0's columns -- or a single layer, S(t+1) follows last(t))
      float bn = 0.f;
      int en = 0;
      gather_a(0); gather_b();
#ifdef MLP_PROFILE
      if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) g_kernel_prof[3] = (unsigned long long)clock64() + (gt.valid ? __float_as_uint(gt.pt[0]) & 0u : 0u);
#endif
      // t = -1 is the prologue (S(0) only)
#pragma unroll 1
      for (int t = -1; t < ntile; ++t) {
        float bound = bn;
        int sexp = en;
        if (pr && t >= 0) PROF(pair, t, 0);
        if (t >= 0) {
#pragma unroll 1
          for (int l = 0; l < last; ++l) {
            const Layer &L = LL[l];
            const float nbound = __fmaf_rn(bound, L.wnorm, L.bmax);
            const int nexp = warp_scale_exp(nbound);
            tc::mbar_wait(bar_acc_full(p, sm, l), (uint32_t)t & 1u);
            tc::fence_after_sync();
            if (pr && l == last - 1) PROF(pair, t, 1);
            emit_mid(l, t, 0, p.S[l + 1], sexp, nexp, lane_base + L.col);
            bound = nbound; sexp = nexp;
          }
        }
        if (pr && t >= 0) PROF(pair, t, 2);
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {                                 // S(t+1) before last(t), or after it (aliased)
          // (the next tile's source slabs go out BEFORE this tile's last layer is read, except when layer 0's
          // accumulator is the one the tensor core is waiting to get back -- aliased columns, or a single-layer
          // stack: then a source longer than the ring would wait for MMAs that wait for this read)
          if (h == ((p.alias || last == 0) ? 1 : 0) && t + 1 < ntile) {
            prepare(t + 1, bn); en = warp_scale_exp(bn);
#ifdef MLP_PROFILE
            if (t == -1 && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) g_kernel_prof[4] = (unsigned long long)clock64() + ((unsigned)en & 0u);
#endif
            emit_source(en);
#ifdef MLP_PROFILE
            if (t == -1 && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) g_kernel_prof[5] = (unsigned long long)clock64();
#endif
            gather_a(t + 2); gather_b();
          }
          if (h == 0 && t >= 0) {
            if (pr) PROF(pair, t, 3);
            tc::mbar_wait(bar_acc_full(p, sm, last), (uint32_t)t & 1u);
            tc::fence_after_sync();
            if (pr) PROF(pair, t, 4);
            last_layer(t, sexp);
            if (pr) PROF(pair, t, 5);
          }
        }
        if (t >= 0) combine(t);
        if (pr && t >= 0) PROF(pair, t, 7);
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
          if (s == 0 && l >= p.NG - 2) PROF(2, it_, 2 * (l - (p.NG - 2)));
          if (s == p.S[l] - 1 && l >= p.NG - 2) PROF(2, it_, 2 * (l - (p.NG - 2)) + 1);
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
#ifdef MLP_PROFILE
  if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) g_kernel_prof[2] = (unsigned long long)clock64();
#endif
  if (warp == MMA_WARP) {
    __syncwarp();
    tc::tmem_dealloc_512(tmem);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// xyz_pipe_kernel: the xyz-only grouped modules (SA1, CloudCrop: layer 1 on the CUDA cores, two tensor-core layers,
// weights resident) with the row warps SPECIALISED BY FUNCTION.  In mlp_f16_kernel all row warps run one program --
// source slabs, mid epilogue, pooling, combine -- in lock step, so a tile costs the SUM of the stages' latencies
// (10.9k cycles per CloudCrop tile for 3.8k cycles of MMAs, whatever the number of warps: 8 and 16 row warps both ran
// 95-97 us, profiles/r02_mlp_f16_ncu.txt).  Here the stages run side by side on four groups of four warps (each group
// covers the 128 rows = TMEM lanes):
//   A0, A1   epilogue of tensor layer 1 (TMEM -> bias + ReLU -> scale -> hi | lo) = the K-slabs of layer 2, alternate
//            slabs, into ring A;
//   P1       max-pool of the first half of layer 2's columns, then combine + output of the tile;
//   P2       max-pool of the second half, then gather + layer 1 (CUDA cores) of tile t + 2 = the K-slabs of tensor
//            layer 1, into ring S;
// plus the MMA issuer and the weight loader.  Every ring slot has ONE producer group (ring lengths are multiples of
// the number of groups that fill them), so the mbarrier parity waits are race-free although the groups drift apart.
// Layer 1's accumulators are double-buffered by tile parity (acc0_full / acc0_free per buffer), layer 2's single
// accumulator is handed back by the poolers (acc1_free).  A tile's per-warp input bound travels from P2 to the other
// groups through a 4-deep ring in shared memory; everything else about the scaling is derived from it.
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void named_barrier_arrive(int id, int nthreads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

struct XyzBars {
  uint32_t s_full, s_empty, a_full, a_empty, w_full, acc0_full, acc0_free, acc1_full, acc1_free;
};
__device__ __forceinline__ XyzBars xyz_bars(const Params &p, uint32_t sm) {
  XyzBars b;
  uint32_t at = sm + p.off_bars;
  b.s_full = at; at += 8 * p.NAs;
  b.s_empty = at; at += 16 * p.NAs;                                  // [slot][parity of the slab number], see ring S
  b.a_full = at; at += 8 * p.NAa;
  b.a_empty = at; at += 8 * p.NAa;
  b.w_full = at; at += 8 * MAX_WSLOTS;
  b.acc0_full = at; at += 16;
  b.acc0_free = at; at += 16;
  b.acc1_full = at; at += 8;
  b.acc1_free = at;
  return b;
}

// The rings of xyz_pipe_kernel move UNITS of 16 channels (one K = 16 step: hi 4 KB | lo 4 KB), half a K-slab: with
// CloudCrop's 160 KB of resident weights 48 KB are left, and a whole tile of source units (4) must fit beside two
// epilogue units -- a slot is released only when the tensor pipe, which executes in issue order, has worked off
// everything queued before it, so a source ring shorter than a tile stalls its producer for a whole layer.
// (Measured: UNIT = 16 ran CloudCrop in 88 us against 76 us with UNIT = 32 -- twice the fences, waits and arrivals,
// and two 16-channel slots buffer only 768 cycles of MMA work.  Kept as a compile-time choice.)
constexpr int UNIT = 32;
constexpr int U_PART = TM * UNIT * 2;          // hi (or lo) part of a unit
constexpr int U_SLOT = 2 * U_PART;
// 16 channels of a row (already scaled) -> hi | lo fp16 at K-chunks 2 h16, 2 h16 + 1 of a unit (base = unit + row * 16)
__device__ __forceinline__ void split_store16(uint32_t base0, int h16, const float (&v)[16]) {
  const uint32_t base = base0 + (uint32_t)h16 * 2u * (TM * 16);
#pragma unroll
  for (int kc = 0; kc < 2; ++kc) {
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
    const uint32_t at = base + (uint32_t)kc * (TM * 16);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(at), "r"(hi[0]), "r"(hi[1]), "r"(hi[2]), "r"(hi[3]) : "memory");
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(at + U_PART), "r"(lo[0]), "r"(lo[1]), "r"(lo[2]), "r"(lo[3]) : "memory");
  }
}
// the same in two steps, so that a slab can be COMPUTED before its ring slot is free and only STORED after:
// 16 scaled channels -> 8 + 8 packed fp16 pairs (hi | lo)
__device__ __forceinline__ void split16(const float (&v)[16], uint32_t (&hi)[8], uint32_t (&lo)[8]) {
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float a = v[2 * e], b = v[2 * e + 1];
    const __half2 h = __floats2half2_rn(a, b);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(__fsub_rn(a, hf.x), __fsub_rn(b, hf.y));
    hi[e] = *reinterpret_cast<const uint32_t *>(&h);
    lo[e] = *reinterpret_cast<const uint32_t *>(&l);
  }
}
__device__ __forceinline__ void store16(uint32_t base0, int h16, const uint32_t (&hi)[8], const uint32_t (&lo)[8]) {
  const uint32_t base = base0 + (uint32_t)h16 * 2u * (TM * 16);
#pragma unroll
  for (int kc = 0; kc < 2; ++kc) {
    const uint32_t at = base + (uint32_t)kc * (TM * 16);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(at), "r"(hi[4 * kc]), "r"(hi[4 * kc + 1]),
                 "r"(hi[4 * kc + 2]), "r"(hi[4 * kc + 3]) : "memory");
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(at + U_PART), "r"(lo[4 * kc]), "r"(lo[4 * kc + 1]),
                 "r"(lo[4 * kc + 2]), "r"(lo[4 * kc + 3]) : "memory");
  }
}
// the warp-uniform exponent of a warp-uniform bound (no reduction needed)
__device__ __forceinline__ int scale_exp_uniform(float bound) {
  const int e = (int)((__float_as_uint(bound) & 0x7fffffffu) >> 23) - 127;
  int s = 13 - e;
  s = s > 60 ? 60 : s;
  return s < -60 ? -60 : s;
}

constexpr int XYZ_GROUPS = 4;
constexpr int XYZ_THREADS = XYZ_GROUPS * TM + 64;

__global__ void __launch_bounds__(XYZ_THREADS, 1) xyz_pipe_kernel(const Params p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t sm = tc::smem_u32(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float4 *w1s = reinterpret_cast<float4 *>(smem_raw + p.off_consts);
  float *biases = reinterpret_cast<float *>(w1s + p.C0);              // [layer][MAXC]
  float *pools = reinterpret_cast<float *>(smem_raw + p.off_pool);    // [tile parity][entries][MAXC] raw maxima
  volatile float *scal = reinterpret_cast<volatile float *>(smem_raw + p.off_xchg);   // [tile & 3][warp & 3] input bound
  volatile uint32_t *holder = reinterpret_cast<volatile uint32_t *>(smem_raw + p.off_holder);
  const XyzBars B = xyz_bars(p, sm);
  const Layer &L0 = p.L[0], &L1 = p.L[1];

  for (int i = tid; i < p.C0; i += XYZ_THREADS)
    w1s[i] = make_float4(__ldg(p.w1 + 3 * i), __ldg(p.w1 + 3 * i + 1), __ldg(p.w1 + 3 * i + 2), __ldg(p.b1 + i));
  for (int l = 0; l < 2; ++l)
    for (int i = tid; i < p.L[l].N; i += XYZ_THREADS) biases[l * MAXC + i] = __ldg(p.L[l].bias + i);
  if (tid == 0) {
    for (int s = 0; s < p.NAs; ++s) tc::mbar_init(B.s_full + 8 * s, TM);
    for (int s = 0; s < 2 * p.NAs; ++s) tc::mbar_init(B.s_empty + 8 * s, 1);
    for (int s = 0; s < p.NAa; ++s) { tc::mbar_init(B.a_full + 8 * s, TM); tc::mbar_init(B.a_empty + 8 * s, 1); }
    for (int s = 0; s < p.NWS; ++s) tc::mbar_init(B.w_full + 8 * s, 1);
    for (int b = 0; b < 2; ++b) { tc::mbar_init(B.acc0_full + 8 * b, 1); tc::mbar_init(B.acc0_free + 8 * b, 2 * TM); }
    tc::mbar_init(B.acc1_full, 1);
    tc::mbar_init(B.acc1_free, XYZ_GROUPS * TM);
    tc::fence_barrier_init();
  }
  if (warp == 4 * XYZ_GROUPS) tc::tmem_alloc_512(sm + p.off_holder);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *holder;
  const int ntile = ((int)blockIdx.x < p.tiles) ? (p.tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int S0 = L0.K / UNIT, S1 = L1.K / UNIT;                       // units per tile feeding tensor layers 1, 2
  const int row = (warp & 3) * 32 + lane;
  const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const uint32_t ringS = sm, ringA = sm + (uint32_t)p.NAs * U_SLOT;
  // scaling chain of a tile from its per-warp input bound b0: exponent of layer 1's output (= tensor layer 1's input),
  // of tensor layer 1's output, ...
  // Ring S is filled by P1 and P2 alternately (slab k by group k % 2).  With an even number of slots a slot always
  // has the same producer and one empty-barrier per slot does; with ONE slot (CloudCrop: 160 KB of weights leave three
  // slots in all) the producers share it, and a parity wait on a single barrier could be two phases off -- so the
  // slot has one empty-barrier per slab parity: slab k is released on barrier k % 2 (its phase k / 2), and the
  // producer of slab k waits for slab k - 1 there, which cannot complete a second time before slab k exists.
  const bool s_shared = (p.NAs & 1) != 0;                             // (then NAs == 1)
  auto s_empty_bar = [&](uint32_t k, uint32_t slot) { return B.s_empty + 16 * slot + (s_shared ? 8 * (k & 1u) : 0u); };
  auto exp_in0 = [&](float b0) { return scale_exp_uniform(b0); };
  auto exp_in1 = [&](float b0) { return scale_exp_uniform(__fmaf_rn(b0, L0.wnorm, L0.bmax)); };

  // Every group pools a quarter of tensor layer 2's columns the moment the accumulator is complete (the A groups have
  // nothing urgent then: their next slot frees only after the next tile's first MMAs): max of the RAW accumulators
  // over the rows of a pooling entry (a warp's 32 rows, or 16 of them); the entry's scale (a positive power of two:
  // max commutes with it) is undone in the combine.  Ends with the arrival that hands the accumulator back.
  float *pool_inv = pools + 2 * p.pool_entries * MAXC;                // [tile parity][entries] the entries' 1 / scale
  auto pool_quarter = [&](int t, int gq) {
    const int nblk = L1.N >> 4, blk0 = (gq * nblk) >> 2, blk1 = ((gq + 1) * nblk) >> 2;
    const uint32_t tcol = lane_base + L1.col;
    float *pool = pools + (t & 1) * (p.pool_entries * MAXC);
    tc::mbar_wait(B.acc1_full, (uint32_t)t & 1u);
    tc::fence_after_sync();
    if (lane == 0 && (warp & 3) == 0) PROF(gq, t, 6);
    if (p.ns >= 32) {
#pragma unroll 1
      for (int blk = blk0; blk < blk1; ++blk) {
        float v[16];
        tc::tmem_ld16(tcol + 16 * blk, v);
#pragma unroll
        for (int c = 0; c < 16; ++c) v[c] = warp_max_nan(v[c]);
        if (lane == 0) {
          float4 *dst = reinterpret_cast<float4 *>(pool + (warp & 3) * MAXC + 16 * blk);
#pragma unroll
          for (int c4 = 0; c4 < 4; ++c4) dst[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
        }
      }
    } else {                                                          // 16-row lists: two entries per warp
#pragma unroll 1
      for (int blk = blk0; blk < blk1; ++blk) {
        float v[16];
        tc::tmem_ld16(tcol + 16 * blk, v);
#pragma unroll
        for (int c = 0; c < 16; ++c) {
#pragma unroll
          for (int o = 8; o >= 1; o >>= 1) v[c] = max_nan(v[c], __shfl_xor_sync(0xffffffffu, v[c], o));
        }
        if ((lane & 15) == 0) {
          float4 *dst = reinterpret_cast<float4 *>(pool + (row >> 4) * MAXC + 16 * blk);
#pragma unroll
          for (int c4 = 0; c4 < 4; ++c4) dst[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
        }
      }
    }
    tc::fence_before_sync();
    tc::mbar_arrive(B.acc1_free);
    if (lane == 0 && (warp & 3) == 0) PROF(gq, t, 7);
  };

  if (warp < 8) {
    // ------------------------------------------------------------------ A0 / A1: tensor layer 1's epilogue
    const int g = warp >> 2;
    Cursor cur = {0u, 0u, 0u};
    const uint32_t NAa = (uint32_t)p.NAa;
    const float4 *bl4 = reinterpret_cast<const float4 *>(biases);
    const float floor_v = L0.relu ? 0.f : __int_as_float(0xff800000);
    const int N = L0.N;
#pragma unroll 1
    for (int t = 0; t < ntile; ++t) {
      if (lane == 0 && (warp & 3) == 0) PROF(g, t, 0);
      bool pooled = t == 0;                                           // tile t - 1 is pooled after this tile's first slab,
      // or before it if its accumulator is complete first (a late source must not delay the pool: the next tile's
      // layer 2 waits for it)
      while (!mbar_test(B.acc0_full + 8 * (t & 1), (uint32_t)(t >> 1) & 1u)) {
        if (!pooled && mbar_test(B.acc1_full, (uint32_t)(t - 1) & 1u)) {
          pool_quarter(t - 1, g);
          named_barrier_arrive(1, XYZ_GROUPS * TM);
          pooled = true;
        } else if (pooled) {                                          // only one thing left to wait for: sleep on it
          tc::mbar_wait(B.acc0_full + 8 * (t & 1), (uint32_t)(t >> 1) & 1u);
        } else {
          __nanosleep(100);                                           // (a hot poll loop steals the working warps' issue slots)
        }
      }
      tc::fence_after_sync();
      if (lane == 0 && (warp & 3) == 0) PROF(g, t, 1);
      const float b0 = scal[(t & 3) * 4 + (warp & 3)];
      const float inv = pow2f(-exp_in0(b0) - L0.wexp), sc = pow2f(exp_in1(b0));
      const uint32_t tcol = lane_base + L0.col + (uint32_t)((t & 1) * N);
#pragma unroll 1
      for (int s = 0; s < S1; ++s, cur.next(NAa)) {
        if ((int)(cur.seq & 1u) != g) continue;
        if (lane == 0 && (warp & 3) == 0 && s < 4) PROF(g, t, 2 + (s >> 1) * 2);
        // the slab is computed into registers first (its accumulator has been complete since acc0_full) and stored
        // once the slot is free (measured: neutral, 74.8 vs 75.8 us -- the issuer is paced by the in-order tensor pipe,
        // not by late slabs; kept because it takes the slot wait out of the compute)
        uint32_t hi[UNIT / 16][8], lo[UNIT / 16][8];
#pragma unroll
        for (int h16 = 0; h16 < UNIT / 16; ++h16) {
          // (checking for the previous tile's completed accumulator here, between the halves, to pool earlier was
          // measured slower: 77.8 us)
          const int ch0 = s * UNIT + 16 * h16;
          float a[16];
          if (ch0 < N) {
            tc::tmem_ld16(tcol + ch0, a);
#pragma unroll
            for (int c4 = 0; c4 < 4; ++c4) {
              const float4 b = bl4[(ch0 >> 2) + c4];
              a[4 * c4] = __fmul_rn(max_nan(__fmaf_rn(a[4 * c4], inv, b.x), floor_v), sc);
              a[4 * c4 + 1] = __fmul_rn(max_nan(__fmaf_rn(a[4 * c4 + 1], inv, b.y), floor_v), sc);
              a[4 * c4 + 2] = __fmul_rn(max_nan(__fmaf_rn(a[4 * c4 + 2], inv, b.z), floor_v), sc);
              a[4 * c4 + 3] = __fmul_rn(max_nan(__fmaf_rn(a[4 * c4 + 3], inv, b.w), floor_v), sc);
            }
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) a[c] = 0.f;
          }
          split16(a, hi[h16], lo[h16]);
        }
        tc::mbar_wait(B.a_empty + 8 * cur.slot, cur.phase ^ 1u);
#pragma unroll
        for (int h16 = 0; h16 < UNIT / 16; ++h16) store16(ringA + cur.slot * U_SLOT + row * 16, h16, hi[h16], lo[h16]);
        tc::fence_proxy_async();
        tc::fence_before_sync();
        tc::mbar_arrive(B.a_full + 8 * cur.slot);
        if (lane == 0 && (warp & 3) == 0 && s < 4) PROF(g, t, 3 + (s >> 1) * 2);
        if (!pooled) {
          pool_quarter(t - 1, g);
          named_barrier_arrive(1, XYZ_GROUPS * TM);
          pooled = true;
        }
      }
      if (!pooled) {                                                  // (a tile without a slab of this group)
        pool_quarter(t - 1, g);
        named_barrier_arrive(1, XYZ_GROUPS * TM);
      }
      tc::fence_before_sync();
      tc::mbar_arrive(B.acc0_free + 8 * (t & 1));                     // this buffer's accumulators have been read
    }
    if (ntile > 0) {
      pool_quarter(ntile - 1, g);
      named_barrier_arrive(1, XYZ_GROUPS * TM);
    }
  } else if (warp < 16) {
    // ------------------------------------------------------------------ P1 / P2: pooling (+ combine | + source)
    const int pg = (warp >> 2) - 2;                                   // 0 = P1, 1 = P2
    const int lpt = TM >> p.log_ns;                                   // lists per tile
    const int cout = L1.cout;
    const float floor_v = L1.relu ? 0.f : __int_as_float(0xff800000);
    // ---- P2 only: the source of a tile (gather in three stages, see mlp_f16_kernel)
    Cursor cur = {0u, 0u, 0u};
    const uint32_t NAs = (uint32_t)p.NAs;
    bool gvalid = false;
    int gbi = 0, gseed = 0, gi_ = 0;
    float gpt[3], gct[3], gR[9];
    auto gather_a = [&](int it_) {
      gvalid = false;
      if (it_ >= ntile) return;
      const int tile_ = blockIdx.x + it_ * gridDim.x;
      const unsigned ql = (unsigned)tile_ * (unsigned)lpt + (unsigned)(row >> p.log_ns);
      gvalid = (long long)ql < p.Q;
      if (gvalid) {
        const unsigned t = fdiv(ql, p.fd_G);
        const int gi = (int)(ql - t * (unsigned)p.G);
        gbi = (int)fdiv(t, p.fd_m);
        gseed = (int)(t - (unsigned)gbi * (unsigned)p.m);
        gi_ = __ldg(p.idx + ((((long long)gi * p.B + gbi) * p.m + gseed) * p.ns + (row & (p.ns - 1))));
      }
    };
    auto gather_b = [&]() {
      if (!gvalid) return;
      const float *pt = p.xyz + ((long long)gbi * p.N + gi_) * 3;
      const float *ct = p.new_xyz + ((long long)gbi * p.m + gseed) * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) { gpt[c] = __ldg(pt + c); gct[c] = __ldg(ct + c); }
      if (p.rot) {
        const float *R = p.rot + ((long long)gbi * p.m + gseed) * 9;
#pragma unroll
        for (int c = 0; c < 9; ++c) gR[c] = __ldg(R + c);
      }
    };
    // stage (C) + layer 1 on the CUDA cores + the K-slabs of tensor layer 1 of tile it_ (gathered before)
    auto source = [&](int it_) {
      float g0 = 0.f, g1 = 0.f, g2 = 0.f;
      if (gvalid) {
        float d0 = __fsub_rn(gpt[0], gct[0]), d1 = __fsub_rn(gpt[1], gct[1]), d2 = __fsub_rn(gpt[2], gct[2]);
        if (p.radius > 0.f) {
          d0 = __fdiv_rn(d0, p.radius); d1 = __fdiv_rn(d1, p.radius); d2 = __fdiv_rn(d2, p.radius);
        }
        if (p.rot) {                                                  // row vector times the seed's 3x3
          g0 = __fmaf_rn(d2, gR[6], __fmaf_rn(d1, gR[3], __fmul_rn(d0, gR[0])));
          g1 = __fmaf_rn(d2, gR[7], __fmaf_rn(d1, gR[4], __fmul_rn(d0, gR[1])));
          g2 = __fmaf_rn(d2, gR[8], __fmaf_rn(d1, gR[5], __fmul_rn(d0, gR[2])));
        } else {
          g0 = d0; g1 = d1; g2 = d2;
        }
      }
      // |y_c| <= max|g| * max_c (|w_c0| + |w_c1| + |w_c2|) + max |b|, the largest of the warp's 32 rows
      const float gmax = max_nan(max_nan(fabsf(g0), fabsf(g1)), fabsf(g2));
      const float b0 = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(__fmaf_rn(gmax, p.w1norm, p.b1max)) & 0x7fffffffu));
      if (lane == 0 && pg == 0) scal[(it_ & 3) * 4 + (warp & 3)] = b0;   // published by the slabs' mbarrier arrivals
      const float sc = pow2f(exp_in0(b0));
#pragma unroll 1
      for (int s = 0; s < S0; ++s, cur.next(NAs)) {
        if ((int)(cur.seq & 1u) != pg) continue;
        if (lane == 0 && (warp & 3) == 0) PROF(2 + pg, it_ - 2, 1);   // slot wait starts
        if (cur.seq >= NAs) {                                         // slab seq - NAs has left the slot
          const uint32_t kp = cur.seq - NAs;
          tc::mbar_wait(s_empty_bar(kp, cur.slot), (s_shared ? (kp >> 1) : (kp / NAs)) & 1u);
        }
        if (lane == 0 && (warp & 3) == 0) PROF(2 + pg, it_ - 2, 2);   // ... slot acquired
#pragma unroll
        for (int h16 = 0; h16 < UNIT / 16; ++h16) {
          const int ch0 = s * UNIT + 16 * h16;
          float v[16];
          if (ch0 < p.C0) {                                           // C0 % 16 == 0
#pragma unroll
            for (int c = 0; c < 16; ++c) {
              const float4 w = w1s[ch0 + c];
              const float a = __fmaf_rn(w.z, g2, __fmaf_rn(w.y, g1, __fmul_rn(w.x, g0)));
              v[c] = __fmul_rn(max_nan(__fadd_rn(a, w.w), 0.f), sc);
            }
          } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) v[c] = 0.f;
          }
          split_store16(ringS + cur.slot * U_SLOT + row * 16, h16, v);
        }
        tc::fence_proxy_async();
        tc::fence_before_sync();
        tc::mbar_arrive(B.s_full + 8 * cur.slot);
      }
    };
    {                                                                 // two tiles ahead of the tensor core
      gather_a(0); gather_b();
      if (ntile > 0) source(0);
      gather_a(1); gather_b();
      if (ntile > 1) source(1);
      gather_a(2);
    }
#pragma unroll 1
    for (int t = 0; t < ntile; ++t) {
      float *pool = pools + (t & 1) * (p.pool_entries * MAXC);
      float *pinv = pool_inv + (t & 1) * p.pool_entries;
      const bool prp = lane == 0 && (warp & 3) == 0;
      if (prp) PROF(2 + pg, t, 0);
      gather_b();                                                     // tile t + 2: stage (B), under the wait and the pool
      pool_quarter(t, 2 + pg);
      if (pg == 0) {
        const float b0 = scal[(t & 3) * 4 + (warp & 3)];
        const float inv = pow2f(-exp_in1(b0) - L1.wexp);
        if (p.ns >= 32) {
          if (lane == 0) pinv[warp & 3] = inv;
        } else if ((lane & 15) == 0) {
          pinv[row >> 4] = inv;
        }
      }
      // all quarters of the pool are written.  (Nothing that can block may stand between the pool and this barrier:
      // the A groups only ARRIVE, and their arrival for tile t + 1 -- a whole layer 2 after this tile's accumulator is
      // handed back -- must not find the poolers of tile t still on their way here.)
      tc::named_barrier(1, XYZ_GROUPS * TM);
      if (prp) PROF(2 + pg, t, 3);
      // P1 produces its source unit of tile t + 2 first and combines after; P2 the other way round -- with a single
      // slot in ring S its unit has to wait until P1's has gone through the tensor pipe, and combining meanwhile
      // keeps it on time for the next pool
      if (pg == 0 && t + 2 < ntile) {
        source(t + 2);
        gather_a(t + 3);
      }
      if (prp) PROF(2 + pg, t, 4);
      {
        // combine the entries of each list (P1 the even ones, P2 the odd ones), unscale, bias + ReLU, write
        // (B, C3, m, G) [and channels-last]
        const float *bl = biases + MAXC;
        const int tile_ = blockIdx.x + t * gridDim.x;
        const int epl = p.ns >= 32 ? p.ns >> 5 : 1;                   // entries per list
        const unsigned q0 = (unsigned)tile_ * (unsigned)lpt;
        const unsigned t0 = fdiv(q0, p.fd_G);                         // the tile's first list; the others follow it
        int lgi = (int)(q0 - t0 * (unsigned)p.G), lbi = (int)fdiv(t0, p.fd_m);
        int lseed = (int)(t0 - (unsigned)lbi * (unsigned)p.m);
#pragma unroll 1
        for (int l = 0; l < lpt && (long long)q0 + l < p.Q; ++l) {
          const float *pl = pool + (l * epl) * MAXC;
          float *dst = p.out + (((long long)lbi * cout) * p.m + lseed) * p.G + lgi;
          float ym = 0.f;
#pragma unroll 1
          for (int c = (l & 1) == pg ? row : cout; c < cout; c += TM) {
            float u = __fmul_rn(pl[c], pinv[l * epl]);
            for (int e = 1; e < epl; ++e) u = max_nan(u, __fmul_rn(pl[e * MAXC + c], pinv[l * epl + e]));
            const float y = max_nan(__fadd_rn(u, bl[c]), floor_v);
            dst[(long long)c * p.m * p.G] = y;
            if (p.out_t) p.out_t[((long long)q0 + l) * cout + c] = y;
            ym = max_nan(ym, fabsf(y));
          }
          if (p.out_rb && (l & 1) == pg) {             // the group's four warps: partial row bounds 0..3, 4..7 unused
            ym = warp_max_nan(ym);
            if (lane == 0) {
              p.out_rb[((long long)q0 + l) * 8 + (warp & 3)] = ym;
              p.out_rb[((long long)q0 + l) * 8 + 4 + (warp & 3)] = 0.f;
            }
          }
          if (++lgi == p.G) {
            lgi = 0;
            if (++lseed == p.m) { lseed = 0; ++lbi; }
          }
        }
      }
      if (pg == 1 && t + 2 < ntile) {
        source(t + 2);
        gather_a(t + 3);
      }
      if (prp) PROF(2 + pg, t, 5);
    }
  } else if (warp == 4 * XYZ_GROUPS) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      Cursor cs = {0u, 0u, 0u}, ca = {0u, 0u, 0u};
      const uint32_t NAs = (uint32_t)p.NAs, NAa = (uint32_t)p.NAa;
      // one unit (16 channels = one K = 16 step, three fp16 products) of a layer into accumulator column dcol;
      // u: the unit's number within the layer (weights: K-slab u / 2, K step u % 2)
      auto unit = [&](const Layer &L, uint32_t w_layer, int u, uint32_t a_base, uint32_t dcol) {
        const uint32_t idesc = idesc_f16_m128(L.N);
        const uint32_t b_lbo = (uint32_t)L.N * 16u, b_part = (uint32_t)L.N * 64u;
#pragma unroll
        for (int kk = 0; kk < UNIT / 16; ++kk) {
          const int ks = u * (UNIT / 16) + kk;                        // K = 16 step of the layer: K-slab ks / 2, step ks % 2
          const uint32_t w_base = w_layer + (uint32_t)(ks >> 1) * (uint32_t)L.N * 128u + (uint32_t)(ks & 1) * 2u * b_lbo;
          const uint64_t ah = tc::smem_desc(a_base + kk * 2 * (TM * 16), TM * 16, 128);
          const uint64_t al = tc::smem_desc(a_base + U_PART + kk * 2 * (TM * 16), TM * 16, 128);
          const uint64_t bh = tc::smem_desc(w_base, b_lbo, 128);
          const uint64_t bl = tc::smem_desc(w_base + b_part, b_lbo, 128);
          mma_f16(tmem + dcol, al, bh, idesc, ks == 0 ? 0u : 1u);
          mma_f16(tmem + dcol, ah, bl, idesc, 1u);
          mma_f16(tmem + dcol, ah, bh, idesc, 1u);
        }
      };
      // Two in-order streams -- tensor layer 1 of tiles tS = 0, 1, ... from ring S and tensor layer 2 of tiles
      // tA = 0, 1, ... from ring A -- merged DYNAMICALLY: whichever stream's next slab is ready is issued (layer 2
      // first: it is the long one), so a late source slab never holds back a layer-2 slab that is ready and vice
      // versa.  (A static order cost ~6k cycles per tile here: the issuer sat in layer 1 of tile t + 2 while tile
      // t + 1's first layer-2 slabs were ready.)  Each accumulator is guarded by its own full / free barriers.
      int tS = 0, sS = 0, tA = 0, sA = 0;
      bool okS = false, okA = false;                                  // the stream's accumulator is known to be free
      for (int s = 0; s < p.NWS; ++s) tc::mbar_wait(B.w_full + 8 * s, 0u);
      while (tA < ntile) {
        bool did = false;
        // ---- layer 2 of tile tA
        if (!okA) okA = tA == 0 || mbar_test(B.acc1_free, (uint32_t)(tA - 1) & 1u);   // the poolers have read tile tA - 1
        if (okA && mbar_test(B.a_full + 8 * ca.slot, ca.phase)) {
          tc::fence_after_sync();
          if (sA < 4) PROF(4, tA, 2 + sA);
          unit(L1, sm + p.off_w + p.woff[1], sA, ringA + ca.slot * U_SLOT, L1.col);
          tc::mma_commit(B.a_empty + 8 * ca.slot);
          ca.next(NAa);
          if (++sA == S1) {
            tc::mma_commit(B.acc1_full);
            PROF(4, tA, 6);
            sA = 0; ++tA; okA = false;
          }
          did = true;
        }
        // ---- layer 1 of tile tS (at most two tiles ahead: two accumulator buffers)
        if (tS < ntile) {
          if (!okS)                                                   // the epilogue of tile tS - 2 has read this buffer
            okS = tS < 2 || mbar_test(B.acc0_free + 8 * (tS & 1), (uint32_t)((tS - 2) >> 1) & 1u);
          if (okS && mbar_test(B.s_full + 8 * cs.slot, cs.phase)) {
            tc::fence_after_sync();
            unit(L0, sm + p.off_w + p.woff[0], sS, ringS + cs.slot * U_SLOT, L0.col + (uint32_t)((tS & 1) * L0.N));
            tc::mma_commit(s_empty_bar(cs.seq, cs.slot));
            cs.next(NAs);
            if (++sS == S0) {
              tc::mma_commit(B.acc0_full + 8 * (tS & 1));
              PROF(4, tS, 7);
              sS = 0; ++tS; okS = false;
            }
            did = true;
          }
        }
        if (!did) __nanosleep(40);
      }
    }
  } else {
    // ------------------------------------------------------------------ weight loader: everything, once
    if (lane == 0 && ntile > 0) {
      for (int l = 0; l < 2; ++l) {
        const uint32_t bytes = (uint32_t)p.L[l].N * 128u;
        const uint8_t *src = reinterpret_cast<const uint8_t *>(p.L[l].wp);
        for (int s = 0; s < p.S[l]; ++s) {
          const uint32_t full = B.w_full + 8 * (p.wslab0[l] + s);
          tc::mbar_arrive_expect_tx(full, bytes);
          tc::bulk_g2s(sm + p.off_w + p.woff[l] + (uint32_t)s * bytes, src + (size_t)s * bytes, bytes, full);
        }
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 4 * XYZ_GROUPS) {
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

FastDiv make_fastdiv(uint32_t d) {                                   // d >= 1; exact for n < 2^31
  uint32_t l = 0;
  while ((1ull << l) < d) ++l;
  FastDiv f;
  f.mul = (uint32_t)(((1ull << 32) * ((1ull << l) - d)) / d + 1);
  f.shr = l;
  return f;
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
  const uint32_t pool = p.sink == SINK_POOL ? 2u * (uint32_t)p.pool_entries * (MAXC + 1) * 4 : 0u;
  const uint32_t xchg = p.src == SRC_XYZ ? 64u : 2u * ROWS * 4;      // xyz: the tiles' input bounds [4][4]
  const uint32_t fixed = consts + pool + xchg + (2 * 8 + NBARS_FIXED) * 8 + 16 + 128;
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
  p.off_xchg = p.off_pool + ((pool + 15u) & ~15u);
  p.off_bars = p.off_xchg + xchg;
  p.off_holder = p.off_bars + (2 * (uint32_t)na + NBARS_FIXED) * 8;
  return (int)(p.off_holder + 16);
}

int g_tiles_per_cta = 1;

int launch(Params &p, cudaStream_t stream) {
  p.fd_G = make_fastdiv((uint32_t)(p.G > 0 ? p.G : 1));
  p.fd_m = make_fastdiv((uint32_t)(p.m > 0 ? p.m : 1));
  const int smem = plan_smem(p);
  if (smem < 0) return smem;
  // xyz-only modules with two tensor-core layers, resident weights, room for a second layer-1 accumulator and at
  // least 3 A slots run the function-specialised kernel
  p.pipelined = 0;
  if (p.NG == 2 && p.w_resident && p.src == SRC_XYZ && 2 * p.L[0].N + p.L[1].N <= 512 && p.NA >= 3) {
    p.pipelined = 1;
    p.L[0].col = 0;
    p.L[1].col = 2u * (uint32_t)p.L[0].N;
    // ring units of UNIT channels: a whole tile of source units if possible, the rest (even, <= 8) for ring A
    const int units = p.NA * (KS / UNIT), u0 = p.L[0].K / UNIT;
    p.NAs = u0 < units - 2 ? u0 : units - 2;
    if (p.NAs > 8) p.NAs = 8;
    p.NAa = (units - p.NAs) & ~1;
    if (p.NAa > 8) p.NAa = 8;
  }
  static B200DeviceOnce attr_set;
  const int dev = b200_device();
  if (attr_set.need(dev)) {
    B200_CUDA(cudaFuncSetAttribute(xyz_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200_CUDA(cudaFuncSetAttribute(mlp_f16_kernel<SRC_XYZ, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200_CUDA(cudaFuncSetAttribute(mlp_f16_kernel<SRC_FEAT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    B200_CUDA(cudaFuncSetAttribute(mlp_f16_kernel<SRC_DENSE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    attr_set.mark(dev);
  }
  // as few CTAs as finish in the same number of tile rounds: 256 tiles on 148 SMs are two rounds either way, on 128
  // CTAs the other 20 SMs stay free for the kernels of the other scenes in flight.
  // (Negative result, round 2: a build of mlp_f16_kernel with 16 row warps -- half a K-slab per thread -- for the
  // launches with ONE tile per CTA, SA3 / SA4 / heads, changed nothing: 32.4 -> 36.4, 30.1 -> 29.5, 28.7 -> 32.8 us.
  // A lone tile's ~30 us is not its row threads' slab work.)
  // b200_mlp_f16_set_tiles_per_cta(k): at least k tile rounds per CTA for the lock-step kernels (SA2-SA4, stacks) -- a
  // lone tile costs its whole dependent chain (SA3: 44k cycles), a second one on the same CTA 16k more because the two
  // row groups overlap tiles: fewer, longer CTAs do the same work in less SM-time (throughput) at a longer launch (latency)
  const int sms = b200_sm_count();
  int rounds = (p.tiles + sms - 1) / sms;
  if (p.src != SRC_XYZ && rounds < g_tiles_per_cta) rounds = g_tiles_per_cta < p.tiles ? g_tiles_per_cta : p.tiles;
  // ... and 4 k for the pipelined xyz kernels (their fill / drain is a few tiles long: SA1's 4 tiles per CTA take 37 us,
  // 16 on a quarter of the CTAs ~3x that; measured with k = 4: steady stream 4151 -> 4201 scenes/s, 28 rounds: 4220)
  if (p.src == SRC_XYZ && g_tiles_per_cta > 1 && rounds < 4 * g_tiles_per_cta)
    rounds = 4 * g_tiles_per_cta < p.tiles ? 4 * g_tiles_per_cta : p.tiles;
  const int grid = rounds > 0 ? (p.tiles + rounds - 1) / rounds : 1;
  if (p.src == SRC_XYZ && p.pipelined) xyz_pipe_kernel<<<grid, XYZ_THREADS, smem, stream>>>(p);
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
extern "C" int b200_mlp_f16_read_kernel_profile(unsigned long long *host) {
  return (int)cudaMemcpyFromSymbol(host, g_kernel_prof, sizeof(g_kernel_prof));
}
extern "C" int b200_mlp_f16_read_slab_profile(unsigned long long *host) {
  return (int)cudaMemcpyFromSymbol(host, g_slab_prof, sizeof(g_slab_prof));
}
#endif

extern "C" void b200_mlp_f16_set_tiles_per_cta(int k) { g_tiles_per_cta = k < 1 ? 1 : (k > 64 ? 64 : k); }

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
                                          const b200_mlp_layer *layers, float *out, float *out_t, float *out_row_bound,
                                          void *stream) {
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
  p.Q = lists; p.tiles = (int)tiles; p.out = out; p.out_t = out_t; p.out_rb = out_row_bound;
  return launch(p, (cudaStream_t)stream);
}

extern "C" int b200_group_feat_mlp_maxpool_f16(const float *xyz, const float *new_xyz, const float *feat_t,
                                               const float *feat_row_bound, const int32_t *idx, int b, int n, int m,
                                               int ns, int c, float radius, const b200_mlp_layer *layers, float *out,
                                               float *out_t, float *out_row_bound, void *stream) {
  if (b < 0 || n < 1 || m < 0 || ns < 1 || c < 4) return B200_EINVAL;
  if (ns != 16 && ns != 32 && ns != 64 && ns != 128) return B200_ECAPACITY;
  if (c % 4) return B200_ECAPACITY;
  const long long lists = (long long)b * m;
  if (lists == 0) return 0;
  if (!xyz || !new_xyz || !feat_t || !idx || !out) return B200_EINVAL;
  if ((reinterpret_cast<uintptr_t>(feat_t) & 31u) != 0) return B200_EINVAL;
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
  if ((reinterpret_cast<uintptr_t>(feat_row_bound) & 31u) != 0) return B200_EINVAL;
  p.Q = lists; p.tiles = (int)tiles; p.out = out; p.out_t = out_t; p.out_rb = out_row_bound; p.feat_rb = feat_row_bound;
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
