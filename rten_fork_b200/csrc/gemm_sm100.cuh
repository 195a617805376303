// Warp-specialised, persistent, TMA-fed tcgen05 GEMM for sm_100a.
//
//   D[M,N] = act( alpha * A[M,K] . B[K,N]  + beta * R[M,N] + bias )
//
// Replaces the reference's BLIS-style loop nest + SIMD micro-kernel
// (reference: src/gemm.rs:733-1050 gemm_impl/gemm_block, src/gemm/kernels.rs:206-316 simd_gemm).  The A/B packing
// step of the reference (src/gemm/packing.rs:20-186) has no equivalent here: TMA writes 128B-swizzled K-major (or
// MN-major) tiles straight into shared memory in the layout the tensor core's matrix descriptors read.
//
// Roles (one persistent CTA per SM, 128 x BN output tile per work unit, BK = 32 floats = one 128-byte swizzle row).
// MODE 0 runs 10 warps (320 threads), the fp32-accurate modes 20 warps (640 threads, setmaxnreg 144 / 64):
//   warps 0-7     epilogue        tcgen05.ld -> alpha / bias / residual / activation -> swizzled staging -> TMA store.
//                                 Warp w may only touch TMEM lanes 32(w%4)..+31; warps 4-7 take the upper half of the
//                                 tile's columns.  MODE 1/2: also fold every 4-k-block accumulation chunk into fp32
//                                 running sums (round-to-nearest adds)
//   warp 8        TMA producer    the whole warp walks the loop, an elected lane issues.  MODE 0: A and B tiles of a
//                                 stage; MODE 1: raw A tiles (own ring: 3 slots at BN = 128, 5 at BN <= 64) and raw B tiles; MODE 2: raw A tiles
//                                 only (tiled, im2col-mode, row-window or batched maps)
//   warp 9        MMA issuer      whole warp walks the loop, an elected lane issues tcgen05.mma + commits; TMEM
//                                 allocation; accumulators double-buffered in TMEM
//   warps 10-17   operand split   (MODE 1/2) two warps per TMEM lane quarter on alternate k-blocks: one A row per
//                                 thread -> hi plane + lo plane (MODE 1) / hi plane (round to nearest TF32) + bf16
//                                 A' plane (MODE 2) into the stage's tensor-memory slot with tcgen05.st
//   warps 18-19   MODE 1: write the lo plane of B next to the raw B tile in shared memory.
//                 MODE 2: warp 18 is the B producer (ONE 3-D TMA load per k-block brings the packed B operand:
//                 [B rounded to TF32][B']; it does not wait for the previous grid when B is a load-time constant);
//                 warp 19 idles (ptxas sizes the register file for whole warpgroups)
//
// MODE 0 (TF32):   one MMA per k-step on the raw fp32 bits (the tensor core keeps the top 19 bits).
// MODE 2 (TF32 + bf16 correction): hi = x rounded to nearest TF32 on BOTH operands (A by the split warps, B at pack
//                  time), lo = x - hi (exact); acc += bf16[lo_A | A] . bf16[hi_B ; lo_B] (one kind::f16 MMA of depth 16
//                  per 8 k) + hi_A * hi_B (one kind::tf32 MMA).  Per-product error <= 2^-18 + 2^-19.
// MODE 1 (3xTF32): hi = raw (the tensor core truncates to the top 19 bits itself; verified on B200),
//                  lo = rna_tf32(x - trunc19(x));
//                  acc += lo_A*hi_B + hi_A*lo_B + hi_A*hi_B.
//   The tensor core adds into the TMEM accumulator with truncation, which for long K drifts past the reference's
//   1e-5 tolerance (measured 2.3e-5 at K=1024).  MODE 1 and 2 therefore accumulate in TMEM only over chunks of
//   kChunkKb k-blocks; the epilogue warps drain each chunk and keep the running sum in registers with
//   round-to-nearest adds (the same two-level scheme the reference gets from its kc=256 depth blocks, src/gemm.rs:546).
#pragma once
#include "sm100_ptx.cuh"
#include "vecmath.cuh"

namespace rtg {

enum BiasMode : int { BIAS_NONE = 0, BIAS_ROW = 1, BIAS_COL = 2 };

// Where output element (m, n) lives:  base(m) + n * col_stride, with
//   row-major   : base = m * ld,                                             col_stride = 1
//   NCHW pixels : base = ((m / hw) * chan_total + chan_off) * hw + (m % hw),  col_stride = hw   (m = image*hw + pixel)
struct OutMap {
  int64_t ld;          // row stride (row-major mode)
  int32_t hw;          // 0 => row-major, >0 => NCHW with hw pixels per image
  int32_t chan_total;  // NCHW: channels of the destination tensor (>= N when writing a group slice)
  int32_t chan_off;    // NCHW / NHWC: first channel (column) written by this launch
};

struct Im2colArgs {  // A operand produced by TMA im2col over NHWC input (A_KIND 1)
  int32_t ow, oh;    // output spatial size
  int32_t stride_w, stride_h;
  int32_t pad_l, pad_t;
  int32_t dil_w, dil_h;
  int32_t kw, kh;
  int32_t c_blocks;  // ceil(C / 32): k-blocks per filter tap
  int32_t c_off;     // first channel of this group
  // A_KIND 2 (row window): an M tile is 128 consecutive pixels of one output row; k-block kb is kernel row kb
  int32_t win_tpr;   // tiles per output row = ceil(ow / 128)
  int32_t win_rows;  // n * oh output rows
  // A_KIND 3 (batched GEMM): batch index z = z1 * bat_n0 + z0; 4-D maps {K, rows, z0, z1}; a broadcast operand keeps
  // coordinate 0 in the batch dimensions it does not have
  int32_t bat_count, bat_n0;
  int32_t bat_a0, bat_a1, bat_b0, bat_b1;  // 1: the operand varies along that batch dimension, 0: broadcast
};

struct GemmArgs {
  int32_t M, N, K;
  int32_t num_kb;  // k-blocks of 32
  float alpha, beta;
  const float* bias;
  int32_t bias_mode;
  const float* R;  // residual / beta source, addressed through `rmap`
  float* D;
  OutMap dmap, rmap;
  int32_t act;
  float act_lo, act_hi;
  int32_t flags;   // bit 0: (debug) MODE 1 also overwrites the raw operand tile with its hi part
  // Split-K (splits > 1): work unit u = split * num_tiles + tile walks k-blocks [split * kb_per_split, +kb_per_split)
  // and leaves its raw partial sums in plane `split` of a {N, M, splits} workspace (tmD); splitk_reduce_kernel adds
  // the planes in a fixed order and applies the epilogue.  Used when a layer has too few output tiles for 148 SMs.
  int32_t splits, kb_per_split;
  int32_t group_m;  // > 0: tiles are walked in groups of group_m m-tiles (L2 reuse for large GEMMs), A_KIND 0 / 1 only
  // N-split GEMM over several GPUs: every finished chunk is also stored (TMA, NVLink) into the same window of each
  // peer's output buffer, so the "all-gather" of the column blocks happens tile by tile inside the GEMM.  The maps
  // live in global memory (host-encoded, copied before the launch).
  const CUtensorMap* peer_maps;
  int32_t n_peers;
  Im2colArgs ic;
  unsigned long long* prof;  // RTG_PROF builds: per-CTA wait / busy cycle counters (see PROF_* below), else unused
  unsigned long long one2;   // (1.0f, 1.0f) as a run-time value: see vm_gelu_x2 (vecmath.cuh)
};

// Developer instrumentation (make PROF=1): cycles each role spends waiting on its barriers.
#ifdef RTG_PROF
#define PROF_DECL long long prof_t0_ = clock64(), prof_acc_[4] = {0, 0, 0, 0}
#define PROF_WAIT(slot, stmt)          \
  do {                                 \
    const long long t_ = clock64();    \
    stmt;                              \
    prof_acc_[slot] += clock64() - t_; \
  } while (0)
// contiguous timeline: every PROF_LINE(slot) charges the time since the previous stamp to `slot`
#define PROF_LINE_DECL long long tl_ = clock64()
#define PROF_LINE(slot)               \
  do {                                \
    const long long u_ = clock64();   \
    prof_acc_[slot] += u_ - tl_;      \
    tl_ = u_;                         \
  } while (0)
#define PROF_MARK(role, slot)                                                                                     \
  do {                                                                                                            \
    if (args.prof) {                                                                                              \
      unsigned long long* p_ = args.prof + (blockIdx.x * 4 + (role)) * 8 + (slot);                                \
      if (*p_ == 0) *p_ = static_cast<unsigned long long>(clock64() - prof_entry_);                               \
    }                                                                                                             \
  } while (0)
#define PROF_MARK_LAST(role, slot)                                                                                \
  do {                                                                                                            \
    if (args.prof)                                                                                                \
      args.prof[(blockIdx.x * 4 + (role)) * 8 + (slot)] = static_cast<unsigned long long>(clock64() - prof_entry_); \
  } while (0)
#define PROF_DUMP(role)                                                              \
  do {                                                                               \
    if (args.prof) {                                                                 \
      unsigned long long* p_ = args.prof + (blockIdx.x * 4 + (role)) * 8;            \
      for (int i_ = 0; i_ < 4; ++i_) p_[i_] = static_cast<unsigned long long>(prof_acc_[i_]); \
      p_[4] = static_cast<unsigned long long>(clock64() - prof_t0_);                 \
    }                                                                                \
  } while (0)
#else
#define PROF_DECL
#define PROF_WAIT(slot, stmt) stmt
#define PROF_LINE_DECL
#define PROF_LINE(slot)
#define PROF_DUMP(role)
#define PROF_MARK(role, slot)
#define PROF_MARK_LAST(role, slot)
#endif

constexpr int kFlagDebugHi = 1;  // (debug) MODE 1 also overwrites the raw operand tile with its hi part
constexpr int kFlagStagedD = 2;  // result leaves through shared memory + TMA store (row-major D, 16-byte aligned)
constexpr int kFlagStagedR = 4;  // residual arrives by TMA load into the same staging buffer (needs kFlagStagedD)
constexpr int kFlagDebugNoB = 32, kFlagDebugNoA = 64;  // (debug, wrong results) MODE 2: skip the B / A TMA stream (timing experiments)
constexpr int kFlagBConst = 512;   // the packed B operand was written at load time: nothing in this stream produces it
constexpr int kFlagAlphaDiv = 16;  // alpha is a divisor: D = acc / alpha (8 is the debug "no staging" switch)

constexpr int kBM = 128;
constexpr int kBK = 32;  // floats per k-block (128 bytes)
constexpr int kABytes = kBM * kBK * 4;
constexpr int kChunkKb = 4;  // MODE 1: k-blocks accumulated in TMEM before the epilogue folds them (K = 128)

// EPI2 (fp32-accurate modes only): two epilogue staging buffers per warp instead of one, paid for with pipeline stages
// (BN = 128: 2 instead of 4).  For launches whose tiles have few k-blocks the epilogue is what a tile waits for: with one
// buffer the TMA store of chunk i has to finish reading it (~2000 cycles) before the residual of chunk i + 1 can be
// loaded into it (~1000 more) -- measured 10 K cycles per 128 x 128 tile on ResNet-50's K = 64 residual layers.
template <int BN, int MODE, bool EPI2 = false>
struct GemmCfg {
  static_assert(MODE == 0 || BN <= 128, "3xTF32 keeps BN/2 running sums per epilogue thread in registers");
  static constexpr int kBBytes = BN * kBK * 4;
  static constexpr int kRawBytes = kABytes + kBBytes;
  // MODE 0: a stage is [A raw][B raw], one ring.
  // MODE 1 (3xTF32): two rings.  The A operand of the MMAs lives in tensor memory: the split warps write hi (= raw
  // bits) and lo of every A row into a 64-column TMEM slot per stage, so the three MMAs of a k-step read only B from
  // shared memory.  A raw tile is therefore dead as soon as it has been split, long before the MMAs that use it have
  // run, and sits in its own short ring of kAStages slots; a pipeline stage proper is [B raw][B lo] plus its TMEM slot.
  // That buys a fourth stage at BN = 128 (sixth at BN = 64) in the same shared memory; this mode is bound by the
  // TMA -> split -> tcgen05.st -> MMA chain, i.e. by how many k-blocks are in flight.
  static constexpr int kAStages = MODE ? (BN <= 64 ? 5 : 3) : 0;
  static constexpr int kStageBytes = MODE ? 2 * kBBytes : kRawBytes;
  static constexpr int kARingBytes = kAStages * kABytes;
  static constexpr int kTmemAStageCols = 2 * kBK;  // hi plane + lo plane, one 32-bit column per k
  // Epilogue staging: every epilogue warp owns kEpiBufs buffers of 32 rows x 128 bytes (one 32-column chunk of its
  // 32 accumulator rows, 128B-swizzled) through which the residual arrives (TMA load) and the result leaves (TMA
  // store), so that global traffic is whole 128-byte rows instead of one strided 16-byte access per lane.
  // Eight epilogue warps in both modes (two per TMEM lane quarter, each owning half of the tile's columns): a chunk
  // is a latency-bound chain (tcgen05.ld -> math -> st.shared -> fence -> TMA store), so throughput comes from
  // running more chunks side by side.
  static constexpr int kEpiWarps = 8;
  // Two staging buffers per warp: the residual of chunk i+1 is already in flight (TMA load) while chunk i is being
  // finished, and the TMA store of chunk i may still read its buffer while chunk i+1 is written.  With one buffer the
  // short-K layers (ResNet layer1/2 pointwise convs, K = 64..128) were bound by this chain, not by HBM.
  // 3xTF32 keeps one buffer: the pipeline stage the extra 32 KB would cost is worth more there.
  static constexpr int kEpiBufs = (MODE != 0 && !EPI2) ? 1 : 2;
  static constexpr int kEpiBufBytes = 4096;
  // Column bias of the whole launch (zero padded to n_tiles * BN entries), loaded once per CTA; launches with more
  // output columns than this fall back to a per-tile slice per warp at the start of the same region.
  static constexpr int kBiasTable = 2048;
  static constexpr int kEpiBiasBytes = kBiasTable * 4;
  static constexpr int kEpiBytes = kEpiWarps * kEpiBufs * kEpiBufBytes + kEpiBiasBytes;
  static constexpr int kMaxSmem = 227 * 1024 - 2048 - kEpiBytes;  // barriers + alignment slack
  static constexpr int kStagesSmem = (kMaxSmem - kARingBytes) / kStageBytes;
  static constexpr int kStagesTmem = MODE ? (512 - 2 * BN) / kTmemAStageCols : 8;  // accumulators take 2 * BN columns
  static constexpr int kStagesRaw = kStagesSmem < kStagesTmem ? kStagesSmem : kStagesTmem;
  // MODE 2: even, so that a stage is always served by the same one of the two split warps of a lane quarter
  static constexpr int kStages = kStagesRaw > 8 ? 8 : MODE != 0 ? (kStagesRaw & ~1) : kStagesRaw;
  static constexpr int kSmemBytes = kARingBytes + kStages * kStageBytes + kEpiBytes + 1024 /*align*/ + 512 /*barriers*/;
  // Eight epilogue warps: a thread carries only BN/2 running sums in the fp32-accurate modes.
  static constexpr int kColsPerThread = BN / (kEpiWarps / 4);
  // MODE 2 has no B split.  One warp's operand-split loop (wait -> LDS -> split -> tcgen05.st -> wait::st -> arrive) takes
  // ~640 busy cycles per k-block against ~480 for the MMAs, so two warps per TMEM lane quarter take alternate blocks:
  // 8 split warps, plus 2 idle warps because ptxas sizes the launch register file for whole warpgroups (640 threads x 96).
  static constexpr int kXfThreads = MODE ? 320 : 0;
  static constexpr int kThreads = (kEpiWarps + 2) * 32 + kXfThreads;
  static constexpr int kTmemCols = MODE ? 512 : (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
};

__device__ __forceinline__ void out_row_base(const OutMap& om, int m, int64_t& base, int64_t& cs) {
  if (om.hw > 0) {
    const int img = m / om.hw;
    base = (static_cast<int64_t>(img) * om.chan_total + om.chan_off) * om.hw + (m - img * om.hw);
    cs = om.hw;
  } else {
    base = static_cast<int64_t>(m) * om.ld + om.chan_off;
    cs = 1;
  }
}

// ---------------------------------------------------------------- out-of-line epilogue paths
// The epilogue is latency-bound and its code used to be ~40 K instructions (every activation, the IEEE division, the
// strided store paths, each unrolled over 32 columns and inlined per call site); ncu attributed a third of the
// epilogue warps' stall samples on the short-K layers to instruction fetch (stall_no_inst).  Everything that is not on
// the path of a fused Conv / MatMul (+bias, +residual, Relu / Clip) into a row-major tensor now lives in these
// functions, called through a scratch copy so that the chunk itself stays in registers on the hot path.
// Eight elements at a time with the selection outside the element loop: one element's Gelu is a chain of ~60 dependent
// instructions (division, exponential), and a loop over single elements ran a 32-column chunk in ~10 K cycles, twice the
// main loop of BERT's 768 -> 3072 layer (329 us against 145 us for the 3072 -> 768 layer of the same flops); eight
// independent chains interleave.
static __device__ __noinline__ void epi_cold_act(float* v, int act, float lo, float hi, unsigned long long one2) {
  if (act == ACT_GELU) {  // BERT's 768 -> 3072 layers: two elements per instruction (vm_gelu_x2), eight pairs at a time
#pragma unroll 1
    for (int j0 = 0; j0 < 32; j0 += 16) {
      float t[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) t[i] = v[j0 + i];
#pragma unroll
      for (int i = 0; i < 16; i += 2) vm_gelu_x2(t[i], t[i + 1], one2);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[j0 + i] = t[i];
    }
    return;
  }
#pragma unroll 1
  for (int j0 = 0; j0 < 32; j0 += 8) {
    float t[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) t[i] = v[j0 + i];
    apply_act_array(t, act, lo, hi);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[j0 + i] = t[i];
  }
}
struct Float8 {
  float4 a, b;
};
static __device__ __noinline__ Float8 epi_cold_gelu8(Float8 x, unsigned long long one2) {
  vm_gelu_x2(x.a.x, x.a.y, one2);
  vm_gelu_x2(x.a.z, x.a.w, one2);
  vm_gelu_x2(x.b.x, x.b.y, one2);
  vm_gelu_x2(x.b.z, x.b.w, one2);
  return x;
}
// fused `MatMul -> Div(scalar)` with a divisor that is not a power of two: IEEE division, eight values per call in registers
static __device__ __noinline__ Float8 epi_cold_div8(Float8 x, float alpha) {
  x.a.x = __fdiv_rn(x.a.x, alpha), x.a.y = __fdiv_rn(x.a.y, alpha), x.a.z = __fdiv_rn(x.a.z, alpha), x.a.w = __fdiv_rn(x.a.w, alpha);
  x.b.x = __fdiv_rn(x.b.x, alpha), x.b.y = __fdiv_rn(x.b.y, alpha), x.b.z = __fdiv_rn(x.b.z, alpha), x.b.w = __fdiv_rn(x.b.w, alpha);
  return x;
}
static __device__ __noinline__ void epi_cold_residual(float* v, const float* r, int64_t cs, float beta, int ncols) {
  for (int j = 0; j < ncols; ++j) v[j] += beta * r[j * cs];
}
static __device__ __noinline__ void epi_cold_store(const float* v, float* d, int64_t cs, int ncols) {
  for (int j = 0; j < ncols; ++j) d[j * cs] = v[j];
}
#define RTG_EPI_COLD(call_on_t)                  \
  do {                                           \
    float t_[32];                                \
    _Pragma("unroll") for (int j_ = 0; j_ < 32; ++j_) t_[j_] = v[j_]; \
    call_on_t;                                   \
    _Pragma("unroll") for (int j_ = 0; j_ < 32; ++j_) v[j_] = t_[j_]; \
  } while (0)

// two fp32 lanes per instruction (sm_100 FADD2 / FFMA2; lane-wise the same IEEE operations)
__device__ __forceinline__ void epi_add2(float& a0, float& a1, float b0, float b1) {
  asm("{\n\t.reg .b64 a, b;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 b, {%2, %3};\n\tadd.rn.f32x2 a, a, b;\n\tmov.b64 {%0, %1}, a;\n\t}"
      : "+f"(a0), "+f"(a1)
      : "f"(b0), "f"(b1));
}
__device__ __forceinline__ void epi_fma2(float& a0, float& a1, float r0, float r1, float s) {  // a += r * s (fused)
  asm("{\n\t.reg .b64 a, r, m;\n\tmov.b64 a, {%0, %1};\n\tmov.b64 r, {%2, %3};\n\tmov.b64 m, {%4, %4};\n\t"
      "fma.rn.f32x2 a, r, m, a;\n\tmov.b64 {%0, %1}, a;\n\t}"
      : "+f"(a0), "+f"(a1)
      : "f"(r0), "f"(r1), "f"(s));
}

// Relu / Clip inline, everything else out of line
__device__ __forceinline__ void epi_act(float (&v)[32], int act, float lo, float hi, unsigned long long one2) {
  if (act == ACT_NONE) return;
  if (act == ACT_RELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
  } else if (act == ACT_CLIP) {
    // f32::clamp as max.NaN / min.NaN: NaN stays NaN like the reference's clamp (plain max / min would return the bound);
    // the one difference is the sign of a zero (-0.0 clamped at lo = +0.0 leaves as +0.0, equal in value).  The nested ?:
    // compiled to a branch per element, and the MobileNetV2 pointwise layers (K = 16 ... 32, one k-block per tile) were
    // bound by exactly that (ncu: the epilogue's stalls were branch_resolving / wait inside this loop)
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      float r;
      asm("{\n\t.reg .f32 t;\n\tmax.NaN.f32 t, %1, %2;\n\tmin.NaN.f32 %0, t, %3;\n\t}" : "=f"(r) : "f"(v[j]), "f"(lo), "f"(hi));
      v[j] = r;
    }
  } else if (act == ACT_GELU) {
    // Out of line but in registers: eight values per call go and come back in registers.  The array form below goes
    // through local memory, which this kernel's 227 KB of shared memory leave almost no L1 for (ncu: the round trips were
    // a third of the Gelu time of BERT's 768 -> 3072 layer; in the fp32-accurate modes the epilogue warps also drain the
    // accumulator every four k-blocks, so their time is the tile's).  Inlining it (16 pairs x 2 call sites) tripled the
    // kernel's code and spilled on the common path.
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      const Float8 r = epi_cold_gelu8(Float8{make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]),
                                             make_float4(v[j + 4], v[j + 5], v[j + 6], v[j + 7])}, one2);
      v[j] = r.a.x, v[j + 1] = r.a.y, v[j + 2] = r.a.z, v[j + 3] = r.a.w;
      v[j + 4] = r.b.x, v[j + 5] = r.b.y, v[j + 6] = r.b.z, v[j + 7] = r.b.w;
    }
  } else {
    RTG_EPI_COLD(epi_cold_act(t_, act, lo, hi, one2));
  }
}

template <int BN, int MODE, int A_KIND, int B_MAJOR, bool EPI2 = false>
__global__ void __launch_bounds__(GemmCfg<BN, MODE, EPI2>::kThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmB2, const __grid_constant__ CUtensorMap tmD,
                    const __grid_constant__ CUtensorMap tmR, const GemmArgs args) {
  using Cfg = GemmCfg<BN, MODE, EPI2>;
#ifdef RTG_PROF
  const long long prof_entry_ = clock64();
  if (threadIdx.x == 0 && args.prof) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    args.prof[(blockIdx.x * 4 + 3) * 8 + 5] = gt;
  }
#endif
  constexpr int kStages = Cfg::kStages;
  // EPI2 launches have at most kChunkKb k-blocks per unit (gemm_launch.cuh): one accumulation chunk, so the epilogue
  // reads it chunk by chunk like MODE 0 instead of carrying BN/2 running sums (same bits: the chunk is the whole sum)
  constexpr bool kChunked = MODE != 0 && !EPI2;
  constexpr int kTmaWarp = Cfg::kEpiWarps, kMmaWarp = Cfg::kEpiWarps + 1;
  extern __shared__ uint8_t smem_raw[];
  // 128B swizzle atoms need 1024-byte aligned tiles.
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring_base = smem_base + Cfg::kARingBytes;  // MODE 1: the A ring comes first
  uint32_t epi_base = ring_base + kStages * Cfg::kStageBytes;  // (not const: see the epilogue role)
  constexpr int kAStages = Cfg::kAStages;
  uint32_t bar_base = epi_base + Cfg::kEpiBytes;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto xf_bar = [&](int s) { return bar_base + 8u * (2 * kStages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (3 * kStages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (3 * kStages + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (3 * kStages + 4);
  auto rbar = [&](int w, int b) { return bar_base + 8u * (3 * kStages + 5 + w * Cfg::kEpiBufs + b); };
  auto afull_bar = [&](int i) { return bar_base + 8u * (3 * kStages + 5 + Cfg::kEpiWarps * Cfg::kEpiBufs + i); };
  auto aempty_bar = [&](int i) { return bar_base + 8u * (3 * kStages + 5 + Cfg::kEpiWarps * Cfg::kEpiBufs + kAStages + i); };
  static_assert(8 * (3 * kStages + 5 + Cfg::kEpiWarps * Cfg::kEpiBufs + 2 * kAStages) <= 512, "barrier region");
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - ptx::smem_u32(smem_raw)));

  int warp = threadIdx.x >> 5;
  int lane = threadIdx.x & 31;

  const int m_tiles_1 = (args.M + kBM - 1) / kBM;  // per batch entry (A_KIND 3: args.M is the per-entry row count)
  const int m_tiles = A_KIND == 2 ? args.ic.win_rows * args.ic.win_tpr
                                  : A_KIND == 3 ? args.ic.bat_count * m_tiles_1 : m_tiles_1;
  const int n_tiles = (args.N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  // Tile index -> (m-tile, n-tile).  Plain order is n fastest.  With args.group_m > 0 (large GEMMs) tiles are walked
  // in groups of group_m m-tiles, m fastest inside a group: the ~148 tiles in flight then cover a group_m x
  // (148 / group_m) block of the output and share that many A row panels and B column panels in L2, instead of
  // 2 row panels and every column panel (at 16384^3 the plain order re-reads all of B from HBM for every other m-tile).
  const int group_m = args.group_m;
  auto tile_mn = [&](int tile, int& mt, int& nt) {
    if (n_tiles == 1) {  // one column of tiles (N <= BN, every MobileNet / most ResNet layers): no division
      mt = tile;
      nt = 0;
    } else if (group_m <= 0) {
      mt = tile / n_tiles;
      nt = tile - mt * n_tiles;
    } else {
      const int per_group = group_m * n_tiles;
      const int g = tile / per_group, r = tile - g * per_group;
      const int gm = min(group_m, m_tiles - g * group_m);  // rows of the (last, partial) group
      nt = r / gm;
      mt = g * group_m + (r - nt * gm);
    }
  };
  const int num_kb = args.num_kb;
  const int chunk_kb = kChunked ? kChunkKb : num_kb;  // k-blocks per TMEM accumulation chunk
  const int splits = args.splits > 1 ? args.splits : 1;
  const int num_units = num_tiles * splits;  // work units: (split, tile), tile fastest
  const int kb_per_split = splits > 1 ? args.kb_per_split : num_kb;

  if (warp == kTmaWarp && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    if (MODE == 2) ptx::prefetch_tmap(&tmB2);
    if (args.flags & kFlagStagedD) ptx::prefetch_tmap(&tmD);
    if (args.flags & kFlagStagedR) ptx::prefetch_tmap(&tmR);
  }
  if (warp == 0) {
    // one barrier per lane and pass (a single thread needed ~1000 cycles for the ~35 of them, on every launch's critical
    // path in front of the first TMA load)
    for (int s = lane; s < kStages; s += 32) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), 1);
      // one arrival per split warp (MODE 2: the four A warps only; its B' tile arrives by TMA)
      ptx::mbar_init(xf_bar(s), MODE == 2 ? 4 : 6);  // A-split warps of the four lane quarters (+ 2 B-lo warps)
    }
    if (lane >= 8 && lane < 10) {
      ptx::mbar_init(tfull_bar(lane - 8), 1);
      ptx::mbar_init(tempty_bar(lane - 8), Cfg::kEpiWarps);  // one arrival per epilogue warp
    }
    for (int i = lane; i < Cfg::kEpiWarps * Cfg::kEpiBufs; i += 32) ptx::mbar_init(rbar(i / Cfg::kEpiBufs, i % Cfg::kEpiBufs), 1);
    if (lane >= 16 && lane < 16 + kAStages) {
      ptx::mbar_init(afull_bar(lane - 16), 1);
      ptx::mbar_init(aempty_bar(lane - 16), 4);  // one arrival per A-split warp
    }
    ptx::fence_mbar_init();
  }
  if (warp == kMmaWarp) {
    ptx::tmem_alloc(tmem_slot, Cfg::kTmemCols);
    ptx::tmem_relinquish();
  }
  // Two setup barriers instead of one __syncthreads: the TMA producers (A stream; MODE 2: also the B stream's warp) need
  // the mbarriers, not tensor memory, so they only meet warp 0 (named barrier 2) and issue their first loads while the
  // TMEM allocation (~1 400 cycles after the CTA starts) is still under way; every other warp meets on named barrier 3
  // as before.  The first loads' latency then overlaps the allocation instead of following it.
  constexpr int kProducerWarps = MODE == 2 ? 2 : 1;
  const bool is_producer = warp == kTmaWarp || (MODE == 2 && warp == kMmaWarp + 9);
  uint32_t tmem_base = 0;
  if (is_producer) {
    asm volatile("bar.sync 2, %0;" ::"n"(32 * (1 + kProducerWarps)) : "memory");
  } else {
    if (warp == 0) asm volatile("bar.arrive 2, %0;" ::"n"(32 * (1 + kProducerWarps)) : "memory");
    ptx::tc_fence_before();
    asm volatile("bar.sync 3, %0;" ::"n"(Cfg::kThreads - 32 * kProducerWarps) : "memory");
    ptx::tc_fence_after();
    tmem_base = *tmem_slot_ptr;
  }
  if (threadIdx.x == 0) PROF_MARK(2, 7);  // setup (barriers, TMEM allocation) done
  // Programmatic dependent launch: everything above touched only this CTA's shared / tensor memory, so it may run
  // while the previous kernel in the stream drains.  From here on global memory is read and written: wait for the
  // predecessor's results (and its last reads of buffers this launch overwrites), then let the successor start its
  // own prologue early.  Both are no-ops when the launch did not ask for programmatic serialisation.
  // The B producer of a launch whose B operand is a load-time constant (weights) skips the wait: its first kStages
  // loads overlap the previous kernel's tail.
  if (!(MODE == 2 && warp == kMmaWarp + 9 && (args.flags & kFlagBConst))) {
    ptx::grid_dependency_wait();
    ptx::grid_launch_dependents();
  }
  // MODE 1: 512 threads start with 128 registers each.  An epilogue thread carries BN/2 running sums plus a 32-column
  // chunk and spilled at that size; the TMA / MMA / operand-split warpgroups need fewer, so every role branch starts
  // by re-balancing the register file: 8 x 32 x kRegsEpi + 8 x 32 x kRegsOther = 64 K.
  // MODE 2 (20 warps x 96 at launch = 61440): 8 x 32 x 144 + 12 x 32 x 64 = 61440.  setmaxnreg is a warpgroup-wide
  // instruction: the four warps of a group (8-11: TMA, MMA, split 0-1; 12-15: split 2-5; 16-19: split 6-7, idle) must
  // ask for the same count.
  constexpr int kRegsEpi = 144, kRegsOther = 64;
  const bool bias_table = args.bias_mode == BIAS_COL && n_tiles * BN <= Cfg::kBiasTable;

  if (warp == kTmaWarp) {
    // ===================================================== TMA producer
    if (MODE != 0) ptx::reg_dealloc<kRegsOther>();
    // Like the MMA warp, the whole warp walks the loop and an elected lane issues: tile coordinates, ring slots and
    // phases are warp-uniform values.
    {
      PROF_DECL;
      int s = 0;
      uint32_t ph = 0;
      int ia = 0;  // MODE 1: slot / phase of the A ring
      uint32_t pa = 0;
      const int ic_dil_h = args.ic.dil_h, ic_dil_w = args.ic.dil_w, ic_c_off = args.ic.c_off;
      const int ic_c_blocks = args.ic.c_blocks, ic_kw = args.ic.kw;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
        const int split = splits > 1 ? unit / num_tiles : 0, tile = unit - split * num_tiles;
        const int kb_begin = split * kb_per_split, kb_end = min(num_kb, kb_begin + kb_per_split);
        int tmt, tnt;
        tile_mn(tile, tmt, tnt);
        int m0 = tmt * kBM;
        const int n0 = tnt * BN;
        int pix_w = 0, pix_h = 0, pix_n = 0;
        int z0 = 0, z1 = 0;
        if (A_KIND == 3) {
          const int z = tmt / m_tiles_1;
          m0 = (tmt - z * m_tiles_1) * kBM;
          z1 = z / args.ic.bat_n0;
          z0 = z - z1 * args.ic.bat_n0;
        }
        if (A_KIND == 2) {
          const int rowidx = tmt / args.ic.win_tpr;
          pix_w = (tmt % args.ic.win_tpr) * kBM;  // first output column of the tile
          pix_n = rowidx / args.ic.oh;
          pix_h = (rowidx - pix_n * args.ic.oh) * args.ic.stride_h;
        }
        if (A_KIND == 1) {
          const int ohw = args.ic.oh * args.ic.ow;
          pix_n = m0 / ohw;
          const int r = m0 - pix_n * ohw;
          const int oy = r / args.ic.ow;
          const int ox = r - oy * args.ic.ow;
          pix_w = ox * args.ic.stride_w - args.ic.pad_l;
          pix_h = oy * args.ic.stride_h - args.ic.pad_t;
        }
        int tap_x = 0, tap_y = 0, cb = 0;
        if (A_KIND == 1 && kb_begin != 0) {  // k-block -> (tap row, tap column, channel block)
          const int tap = kb_begin / ic_c_blocks;
          cb = kb_begin - tap * ic_c_blocks;
          tap_y = tap / ic_kw;
          tap_x = tap - tap_y * ic_kw;
        }
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          uint32_t a_dst, b_dst, fa;
          const uint32_t fb = full_bar(s);
          if (MODE != 0) {
            PROF_WAIT(0, ptx::mbar_wait(aempty_bar(ia), pa ^ 1u));
            a_dst = smem_base + ia * kABytes;
            fa = afull_bar(ia);
            b_dst = ring_base + s * Cfg::kStageBytes;
          } else {
            PROF_WAIT(0, ptx::mbar_wait(empty_bar(s), ph ^ 1u));
            a_dst = ring_base + s * Cfg::kStageBytes;
            fa = fb;
            b_dst = a_dst + kABytes;
          }
#ifdef RTG_PROF
          const long long ti_ = clock64();
#endif
          PROF_MARK(0, 5);  // first TMA issue
          const bool leader = ptx::elect_one() != 0;
          if (leader && MODE == 2 && (args.flags & kFlagDebugNoA)) {  // (debug, wrong results) timing without the A stream
            ptx::mbar_arrive(fa);
          } else if (leader) {
            ptx::mbar_expect_tx(fa, MODE != 0 ? kABytes : Cfg::kRawBytes);
            if (A_KIND == 0) {
              ptx::tma_load_2d(a_dst, &tmA, fa, kb * kBK, m0);
            } else if (A_KIND == 3) {
              ptx::tma_load_4d(a_dst, &tmA, fa, kb * kBK, m0, z0 * args.ic.bat_a0, z1 * args.ic.bat_a1);
            } else if (A_KIND == 2) {
              // {window floats, output column, padded input row, image}: columns past ow are zero-filled by TMA
              ptx::tma_load_4d(a_dst, &tmA, fa, 0, pix_w, pix_h + kb * ic_dil_h, pix_n);
            } else {
              ptx::tma_load_im2col_4d(a_dst, &tmA, fa, ic_c_off + cb * kBK, pix_w, pix_h, pix_n,
                                      static_cast<uint16_t>(tap_x * ic_dil_w), static_cast<uint16_t>(tap_y * ic_dil_h));
            }
          }
          if (A_KIND == 1) {
            if (++cb == ic_c_blocks) {
              cb = 0;
              if (++tap_x == ic_kw) {
                tap_x = 0;
                ++tap_y;
              }
            }
          }
          if (MODE != 0) {  // the B slot (and its TMEM slot) is free once the MMAs of its previous use have completed
            if (++ia == kAStages) {
              ia = 0;
              pa ^= 1u;
            }
            if (MODE == 2) continue;  // the B stream has its own producer warp (below)
            PROF_WAIT(0, ptx::mbar_wait(empty_bar(s), ph ^ 1u));
            if (leader) ptx::mbar_expect_tx(fb, Cfg::kBBytes);
          }
          if (!leader) {
            // only the elected lane issues the B loads
          } else if (A_KIND == 3) {
            const int bz0 = z0 * args.ic.bat_b0, bz1 = z1 * args.ic.bat_b1;
            if (B_MAJOR == 0) {
              ptx::tma_load_4d(b_dst, &tmB, fb, kb * kBK, n0, bz0, bz1);
            } else {
#pragma unroll
              for (int c = 0; c < BN / 32; ++c) ptx::tma_load_4d(b_dst + c * 4096, &tmB, fb, n0 + c * 32, kb * kBK, bz0, bz1);
            }
          } else if (B_MAJOR == 0) {
            // B stored [N][K] (K contiguous): one box of BN rows x 128 bytes
            ptx::tma_load_2d(b_dst, &tmB, fb, kb * kBK, n0);
          } else {
            // B stored [K][N] (N contiguous): BN/32 boxes of 32 k-rows x 128 bytes, 4 KB apart
#pragma unroll
            for (int c = 0; c < BN / 32; ++c) ptx::tma_load_2d(b_dst + c * 4096, &tmB, fb, n0 + c * 32, kb * kBK);
          }
#ifdef RTG_PROF
          prof_acc_[1] += clock64() - ti_;
#endif
          if (++s == kStages) {
            s = 0;
            ph ^= 1u;
          }
        }
      }
      if (lane == 0) PROF_DUMP(0);
    }
  } else if (warp == kMmaWarp) {
    // ===================================================== MMA issuer
    if (MODE != 0) ptx::reg_dealloc<kRegsOther>();
    // The whole warp walks the loop (uniform control flow: stage / descriptor arithmetic stays in uniform registers and
    // the waits are not behind a divergent branch); one elected lane issues the MMAs and their commits.
    {
      constexpr uint32_t idesc = ptx::make_idesc_tf32(kBM, BN, 0, B_MAJOR);
      // K-major operands: 128B swizzle, 8-row groups 1024 B apart.  MN-major fp32 operands must use the
      // 128B-swizzle-with-32B-atoms layout: 4 k-rows (512 B) per atom, 32-column chunks 4096 B apart.
      constexpr uint32_t b_layout = B_MAJOR ? ptx::kSwizzle128Base32 : ptx::kSwizzle128;
      constexpr uint32_t b_lbo = B_MAJOR ? 4096 : 16;
      constexpr uint32_t b_sbo = B_MAJOR ? 512 : 1024;
      // descriptor words of stage 0; a stage / k-step only moves the 14-bit "address >> 4" field of the low word
      const uint64_t a_d0 = ptx::make_smem_desc(ring_base, 16, 1024);  // MODE 0 only (MODE 1 reads A from TMEM)
      const uint64_t b_d0 = ptx::make_smem_desc(ring_base + (MODE ? 0 : kABytes), b_lbo, b_sbo, b_layout);
      // MODE 2: the bf16 correction tile B' is always K-major (written by the split warps), whatever B's own layout
      constexpr uint32_t idesc_c = ptx::make_idesc_bf16(kBM, BN);
      const uint64_t bc_d0 = ptx::make_smem_desc(ring_base, 16, 1024);
      const uint32_t bc_hi = static_cast<uint32_t>(bc_d0 >> 32), bc_lo0 = static_cast<uint32_t>(bc_d0);
      const uint32_t a_hi = static_cast<uint32_t>(a_d0 >> 32), b_hi = static_cast<uint32_t>(b_d0 >> 32);
      const uint32_t a_lo0 = static_cast<uint32_t>(a_d0), b_lo0 = static_cast<uint32_t>(b_d0);
      constexpr uint32_t kStageStep = Cfg::kStageBytes >> 4;
      constexpr uint32_t kBLoStep = Cfg::kBBytes >> 4;  // MODE 1: the lo plane of B follows the raw B tile
      constexpr uint32_t kAStep = 32 >> 4;                       // 8 floats along the swizzled 128-byte row
      constexpr uint32_t kBStep = (B_MAJOR ? 1024 : 32) >> 4;    // MN-major: 8 k-rows = 1024 bytes further
      int s = 0;
      uint32_t so = 0;  // s * kStageStep
      uint32_t ph = 0;
      int as = 0;
      uint32_t aph = 0;
      PROF_DECL;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
        const int kb_begin = (splits > 1 ? unit / num_tiles : 0) * kb_per_split, kb_end = min(num_kb, kb_begin + kb_per_split);
        for (int kb0 = kb_begin; kb0 < kb_end; kb0 += chunk_kb) {
          PROF_WAIT(0, ptx::mbar_wait(tempty_bar(as), aph ^ 1u));
          ptx::tc_fence_after();
          const uint32_t d_tmem = tmem_base + as * BN;
          const int kb1 = min(kb_end, kb0 + chunk_kb);
          for (int kb = kb0; kb < kb1; ++kb) {
            PROF_WAIT(1, ptx::mbar_wait(MODE ? xf_bar(s) : full_bar(s), ph));
            if (MODE == 2) PROF_WAIT(1, ptx::mbar_wait(full_bar(s), ph));  // B and B' have landed
            ptx::tc_fence_after();
#ifdef RTG_PROF
            const long long ti_ = clock64();
#endif
            PROF_MARK(1, 5);  // first operands ready
            const uint32_t a_lo = a_lo0 + so, b_lo = b_lo0 + so;
            if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < kBK / 8; ++k) {
              const uint32_t acc = (kb != kb0 || k != 0) ? 1u : 0u;
              if (MODE == 0) {
                ptx::mma_tf32_ss_w(d_tmem, a_lo + k * kAStep, a_hi, b_lo + k * kBStep, b_hi, idesc, acc);
              } else if (MODE == 1) {
                // lo(A) * B, A * lo(B), A * B: smallest terms first.  A planes: TMEM columns [0,32) hi, [32,64) lo
                const uint32_t a_t = tmem_base + 2 * BN + s * Cfg::kTmemAStageCols + k * 8;
                ptx::mma_tf32_ts_w(d_tmem, a_t + kBK, b_lo + k * kBStep, b_hi, idesc, acc);
                ptx::mma_tf32_ts_w(d_tmem, a_t, b_lo + kBLoStep + k * kBStep, b_hi, idesc, 1);
                ptx::mma_tf32_ts_w(d_tmem, a_t, b_lo + k * kBStep, b_hi, idesc, 1);
              } else {
                // MODE 2: both correction products in one bf16 MMA of depth 16 = [lo(A) 8 | hi(A) 8] x [hi(B) 8 ; lo(B) 8]
                // (A' in TMEM columns [32,64) of the stage, B' K-major behind the raw B tile), then A * B in TF32
                const uint32_t a_t = tmem_base + 2 * BN + s * Cfg::kTmemAStageCols + k * 8;
                ptx::mma_f16_ts_w(d_tmem, a_t + kBK, bc_lo0 + so + kBLoStep + k * 2, bc_hi, idesc_c, acc);
                ptx::mma_tf32_ts_w(d_tmem, a_t, b_lo + k * kBStep, b_hi, idesc, 1);
              }
            }
            ptx::mma_commit(empty_bar(s));
            }
#ifdef RTG_PROF
            prof_acc_[2] += clock64() - ti_;
#endif
            so += kStageStep;
            if (++s == kStages) {
              s = 0;
              so = 0;
              ph ^= 1u;
            }
          }
          if (ptx::elect_one()) ptx::mma_commit(tfull_bar(as));
          PROF_MARK_LAST(1, 6);  // last accumulator commit
          if (++as == 2) {
            as = 0;
            aph ^= 1u;
          }
        }
      }
      if (lane == 0) PROF_DUMP(1);
    }
  } else if (warp < Cfg::kEpiWarps) {
    // ===================================================== epilogue
    if (MODE != 0) ptx::reg_alloc<kRegsEpi>();
    // Opaque to the compiler inside this role: it otherwise rematerialises the shared-window base (S2R SR_CgaCtaId + 3 ALU)
    // and the thread ids (S2R SR_TID.X) in front of nearly every shared-memory access, ~100 instructions per tile on
    // this role's latency-bound path.  Only here, where there are registers to keep them in: done for the whole kernel
    // the values were spilled and re-loaded from local memory inside the split warps' loop.
    asm volatile("" : "+r"(epi_base), "+r"(bar_base), "+r"(warp), "+r"(lane));
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    constexpr int kCols = Cfg::kColsPerThread;
    const int col0 = (warp >> 2) * kCols;  // first tile column this warp handles
    const bool staged_d = (args.flags & kFlagStagedD) != 0;
    const bool staged_r = (args.flags & kFlagStagedR) != 0;
    const uint32_t ebuf0 = epi_base + warp * Cfg::kEpiBufs * Cfg::kEpiBufBytes;
    const uint32_t ebias_tab = epi_base + Cfg::kEpiWarps * Cfg::kEpiBufs * Cfg::kEpiBufBytes;
    if (bias_table) {
      for (int n = threadIdx.x; n < n_tiles * BN; n += Cfg::kEpiWarps * 32)
        ptx::sts_f32(ebias_tab + n * 4, n < args.N ? __ldg(args.bias + n) : 0.f);
      asm volatile("bar.sync 1, %0;" ::"n"(Cfg::kEpiWarps * 32) : "memory");  // the epilogue warps only
    }
    int eb = 0;            // staging buffer of the next chunk
    uint32_t rph = 0;      // phase bit per staging buffer of this warp's residual barriers
    int as = 0;
    uint32_t aph = 0;
    PROF_DECL;
    PROF_LINE_DECL;
    for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
      const int split = splits > 1 ? unit / num_tiles : 0, tile = unit - split * num_tiles;
      const int kb_begin = split * kb_per_split, kb_end = min(num_kb, kb_begin + kb_per_split);
      int tmt, tnt;
      tile_mn(tile, tmt, tnt);
      const int n0 = tnt * BN;
      // store coordinates of this warp's 32 rows: (st_r, st_z) in the {columns, rows, planes} view of D / R
      int m, st_r, st_z = 0;
      bool row_ok;
      if (A_KIND == 2) {
        st_z = tmt / args.ic.win_tpr;
        st_r = (tmt % args.ic.win_tpr) * kBM + q * 32;
        m = st_z * args.ic.ow + st_r + lane;
        row_ok = st_r + lane < args.ic.ow;
      } else if (A_KIND == 3) {
        st_z = tmt / m_tiles_1;
        st_r = (tmt - st_z * m_tiles_1) * kBM + q * 32;
        m = st_z * args.M + st_r + lane;
        row_ok = st_r + lane < args.M;
      } else {
        st_r = tmt * kBM + q * 32;
        st_z = split;  // split-K: plane of the partial-sum workspace (0 otherwise)
        m = st_r + lane;
        row_ok = m < args.M;
      }
      // element offsets of this row for the strided (non-TMA) output / residual paths; not needed when both are staged
      int64_t d_base = 0, r_base = 0, d_cs = 1, r_cs = 1;
      if (!staged_d) out_row_base(args.dmap, m, d_base, d_cs);
      if (args.R != nullptr && !staged_r) out_row_base(args.rmap, m, r_base, r_cs);
      const float row_bias = (args.bias_mode == BIAS_ROW && row_ok) ? args.bias[m] : 0.f;
      const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);

      // this warp's slice of the per-column bias -> shared memory (read back as broadcast loads in finish), issued
      // before the wait for the accumulators so that the global-load latency is hidden
      uint32_t ebias = ebias_tab + (n0 + col0) * 4;
      if (args.bias_mode == BIAS_COL && !bias_table) {
        ebias = ebias_tab + warp * (BN / 2) * 4;
#pragma unroll
        for (int i = 0; i < kCols / 32; ++i) {
          const int n = n0 + col0 + i * 32 + lane;
          ptx::sts_f32(ebias + (i * 32 + lane) * 4, n < args.N ? __ldg(args.bias + n) : 0.f);
        }
        __syncwarp();
      }

      // residual chunk (32 rows x 32 columns at column c0 of the tile) -> staging buffer b, by TMA
      // `newer` = stores issued after the last one that used buffer b (they may still be reading the other buffer)
      auto fetch_residual = [&](int c0, int b, int newer) {
        if (lane == 0) {
          // earlier stores of this warp no longer read the buffer
          if (newer > 0) {
            ptx::bulk_wait_read<1>();
          } else {
            ptx::bulk_wait_read<0>();
          }
          ptx::mbar_expect_tx(rbar(warp, b), Cfg::kEpiBufBytes);
          ptx::tma_load_3d(ebuf0 + b * Cfg::kEpiBufBytes, &tmR, rbar(warp, b), n0 + c0, st_r, st_z);
        }
      };

      // alpha / bias / residual / activation on 32 consecutive columns starting at n0 + c0, then store.  Every
      // selection is warp-uniform and sits outside the element loops, so the executed path stays a few hundred
      // instructions per chunk.
      auto finish = [&](float (&v)[32], int c0) {
        const int ncols = min(32, args.N - (n0 + c0));
        if (args.flags & kFlagAlphaDiv) {  // fused `MatMul -> Div(scalar)`: IEEE division, as the unfused operator
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            const Float8 r = epi_cold_div8(Float8{make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]),
                                                  make_float4(v[j + 4], v[j + 5], v[j + 6], v[j + 7])}, args.alpha);
            v[j] = r.a.x, v[j + 1] = r.a.y, v[j + 2] = r.a.z, v[j + 3] = r.a.w;
            v[j + 4] = r.b.x, v[j + 5] = r.b.y, v[j + 6] = r.b.z, v[j + 7] = r.b.w;
          }
        } else if (args.alpha != 1.f) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] *= args.alpha;
        }
        // bias before the residual: a fused Conv -> Add computes (acc + bias) + r exactly as the unfused pair does
        if (args.bias_mode == BIAS_ROW) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += row_bias;
        } else if (args.bias_mode == BIAS_COL) {
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 b4 = ptx::lds_v4(ebias + (c0 - col0 + 4 * j4) * 4);  // same address in every lane: broadcast
            epi_add2(v[4 * j4 + 0], v[4 * j4 + 1], b4.x, b4.y);
            epi_add2(v[4 * j4 + 2], v[4 * j4 + 3], b4.z, b4.w);
          }
        }
        if (staged_d) {
          const uint32_t buf = ebuf0 + eb * Cfg::kEpiBufBytes;
          const uint32_t row = buf + lane * 128;
          const uint32_t sw = lane & 7;
          if (staged_r) {
            PROF_LINE(2);
            ptx::mbar_wait_parked(rbar(warp, eb), (rph >> eb) & 1u);
            PROF_LINE(0);
            rph ^= 1u << eb;
#pragma unroll
            for (int j4 = 0; j4 < 8; ++j4) {
              const float4 r = ptx::lds_v4(row + ((j4 ^ sw) << 4));
              epi_fma2(v[4 * j4 + 0], v[4 * j4 + 1], r.x, r.y, args.beta);
              epi_fma2(v[4 * j4 + 2], v[4 * j4 + 3], r.z, r.w, args.beta);
            }
          } else {
            PROF_LINE(2);
            if (lane == 0) ptx::bulk_wait_read<Cfg::kEpiBufs - 1>();
            __syncwarp();
            PROF_LINE(0);
            if (args.R != nullptr && row_ok)  // residual that TMA cannot address (unaligned / NCHW): direct loads
              RTG_EPI_COLD(epi_cold_residual(t_, args.R + r_base + static_cast<int64_t>(n0 + c0) * r_cs, r_cs, args.beta, ncols));
          }
          epi_act(v, args.act, args.act_lo, args.act_hi, args.one2);
          PROF_LINE(2);
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4)
            ptx::sts_v4(row + ((j4 ^ sw) << 4), make_float4(v[4 * j4], v[4 * j4 + 1], v[4 * j4 + 2], v[4 * j4 + 3]));
          ptx::fence_proxy_async_smem();
          __syncwarp();
#ifdef RTG_PROF_EPI_FINE
          PROF_LINE(1);
#endif
          if (lane == 0) {
            ptx::tma_store_3d(&tmD, buf, n0 + c0, st_r, st_z);  // clipped to the tensor bounds by the TMA unit
            for (int p = 0; p < args.n_peers; ++p) ptx::tma_store_3d(args.peer_maps + p, buf, n0 + c0, st_r, st_z);
            ptx::bulk_commit();
          }
          if (Cfg::kEpiBufs > 1) eb ^= 1;
          PROF_LINE(3);
          return;
        }
        // outputs TMA cannot address (NCHW for a host caller, unaligned rows): strided stores, out of line
        if (!row_ok) return;
        if (args.R != nullptr)
          RTG_EPI_COLD(epi_cold_residual(t_, args.R + r_base + static_cast<int64_t>(n0 + c0) * r_cs, r_cs, args.beta, ncols));
        epi_act(v, args.act, args.act_lo, args.act_hi, args.one2);
        RTG_EPI_COLD(epi_cold_store(t_, args.D + d_base + static_cast<int64_t>(n0 + c0) * d_cs, d_cs, ncols));
      };

      // chunks of this warp that lie inside the matrix
      const int nch = max(0, min(kCols, args.N - n0 - col0) + 31) / 32;
      // buffers alternate strictly, also across tiles: the last store used eb ^ 1, the one before it eb
      constexpr int kOther = Cfg::kEpiBufs > 1 ? 1 : 0;  // eb ^ kOther: the other staging buffer (or the only one)
      constexpr int kAhead = Cfg::kEpiBufs;              // residual chunks in flight
      if (staged_r && nch > 0) fetch_residual(col0, eb, kOther);
      if (Cfg::kEpiBufs > 1 && staged_r && nch > 1) fetch_residual(col0 + 32, eb ^ 1, 0);

      if (!kChunked) {
        PROF_LINE(3);
          ptx::mbar_wait_parked(tfull_bar(as), aph);
          PROF_LINE(0);
        ptx::tc_fence_after();
#pragma unroll 1
        for (int i = 0; i < nch; ++i) {
          const int c0 = col0 + i * 32;
          float v[32];
          ptx::tmem_ld_32x32(lane_addr + as * BN + c0, v);
          ptx::tmem_ld_wait();
          finish(v, c0);
          // chunk i + 2 reuses the buffer whose store was just issued (eb has toggled: it is eb ^ 1 now)
          if (staged_r && i + kAhead < nch) fetch_residual(c0 + 32 * kAhead, eb ^ kOther, 0);
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(tempty_bar(as));
        if (++as == 2) {
          as = 0;
          aph ^= 1u;
        }
      } else {
        // running sums in registers, one per output column of this thread's row
        float acc[kCols];
        for (int kb0 = kb_begin; kb0 < kb_end; kb0 += chunk_kb) {
          PROF_LINE(3);
          ptx::mbar_wait_parked(tfull_bar(as), aph);
          PROF_LINE(0);
          ptx::tc_fence_after();
#pragma unroll
          for (int c0 = 0; c0 < kCols; c0 += 32) {
            float v[32];
            ptx::tmem_ld_32x32(lane_addr + as * BN + col0 + c0, v);
            ptx::tmem_ld_wait();
            if (kb0 == kb_begin) {
#pragma unroll
              for (int j = 0; j < 32; ++j) acc[c0 + j] = v[j];
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) acc[c0 + j] = __fadd_rn(acc[c0 + j], v[j]);
            }
          }
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(tempty_bar(as));
          if (++as == 2) {
            as = 0;
            aph ^= 1u;
          }
          PROF_LINE(1);
        }
#pragma unroll 1
        for (int ci = 0; ci < nch; ++ci) {  // one call site of `finish`: the chunk is selected, not unrolled
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = (kCols > 32 && ci > 0) ? acc[(32 + j) % kCols] : acc[j];
          finish(v, col0 + ci * 32);
          if (staged_r && ci + kAhead < nch) fetch_residual(col0 + (ci + kAhead) * 32, eb ^ kOther, 0);
        }
      }
    }
    // the stores must have READ their staging buffers before shared memory goes away; their global writes complete with
    // the grid (a dependent launch waits for that in griddepcontrol.wait)
    if (staged_d && lane == 0) ptx::bulk_wait_read<0>();
    PROF_LINE(0);
    if (warp == 0 && lane == 0) {
      PROF_MARK_LAST(2, 5);  // epilogue drained (stores complete)
      PROF_DUMP(2);
    }
  } else if (MODE == 2 && warp == kMmaWarp + 9) {
    // ===================================================== B producer (MODE 2)
    // One 3-D TMA load per k-block brings both planes of the packed B operand (make_b_pack: plane 0 = B rounded to
    // TF32, plane 1 = the bf16 correction operand B', both K-major) into the stage: [B hi][B'].  Its own warp, so that
    // the A stream (warp kTmaWarp) and the B stream do not queue behind each other's waits and issue latencies.
    ptx::reg_dealloc<kRegsOther>();
    int s = 0;
    uint32_t ph = 0;
    for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
      const int split = splits > 1 ? unit / num_tiles : 0, tile = unit - split * num_tiles;
      const int kb_begin = split * kb_per_split, kb_end = min(num_kb, kb_begin + kb_per_split);
      int tmt, tnt;
      tile_mn(tile, tmt, tnt);
      const int n0 = tnt * BN;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        ptx::mbar_wait(empty_bar(s), ph ^ 1u);
        if (ptx::elect_one()) {
          if (args.flags & kFlagDebugNoB) {  // (debug, wrong results) timing without the B stream
            ptx::mbar_arrive(full_bar(s));
          } else {
            ptx::mbar_expect_tx(full_bar(s), (args.flags & 128) ? Cfg::kBBytes : 2 * Cfg::kBBytes);  // 128: debug half box
            ptx::tma_load_3d(ring_base + s * Cfg::kStageBytes, &tmB2, full_bar(s), kb * kBK, n0, 0);
          }
        }
        if (++s == kStages) {
          s = 0;
          ph ^= 1u;
        }
      }
    }
  } else if (MODE != 0 && warp > kMmaWarp) {
    // ===================================================== operand split (fp32-accurate modes)
    // x = hi + lo with hi = x truncated to TF32 (the tensor core reads that from the raw fp32 bits by itself) and
    // lo = x - hi (exact in fp32).  Thread = one row of the A tile; per block the row goes to the stage's tensor-memory
    // slot as its raw bits plus, MODE 1: lo rounded to nearest TF32 (+ half an ulp, the tensor core drops the low 13
    // bits); MODE 2: A' = per group of 8 k four words of bf16 pairs of lo(A) and four of bf16(A).
    // MODE 1 only: warps 8-9 of the role write the lo plane of B next to the raw B tile.
    auto lo_of = [](float x) {
      return __float_as_uint(x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u)) + 0x1000u;
    };
    ptx::reg_dealloc<kRegsOther>();
    const int sw = warp - (kMmaWarp + 1);
    if (sw < 8) {
      PROF_DECL;
      int total = 0;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
        const int kb_begin = (splits > 1 ? unit / num_tiles : 0) * kb_per_split;
        total += min(num_kb, kb_begin + kb_per_split) - kb_begin;
      }
      const int first = sw >> 2;  // the two warps of a lane quarter take alternate blocks of the CTA's flat k-block stream
      const int q = warp & 3;     // TMEM lane quarter this warp may write
      const int row = q * 32 + lane;
      const uint32_t swz = row & 7;
      int ia = first % kAStages, s = first % kStages;
      uint32_t pa = 0, ph = 0;
      static_assert(MODE == 0 || kStages % 2 == 0, "a pipeline stage must always be served by the same one of the two warps");
      PROF_LINE_DECL;
      for (int it = first; it < total; it += 2) {
        // The A ring has 3 slots, so this warp only sees every other phase of a slot's barriers and a parity wait
        // alone could be satisfied by the phase before the one it wants.  The slot's previous round (served by the
        // other warp of the quarter) has to be over first: its `aempty` phase cannot run ahead of this warp.
        ptx::mbar_wait_parked(aempty_bar(ia), pa ^ 1u);
        ptx::mbar_wait_parked(afull_bar(ia), pa);
        PROF_LINE(0);
        const uint32_t src = smem_base + ia * kABytes + row * 128;
        // the TMEM slot of this stage is free once the MMAs of its previous use have completed
        ptx::mbar_wait_parked(empty_bar(s), ph ^ 1u);
        ptx::tc_fence_after();
        PROF_LINE(1);
        const uint32_t dst = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + 2 * BN + s * Cfg::kTmemAStageCols;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t r[16], w[16];
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            const float4 x = ptx::lds_v4(src + (((4 * h + j4) ^ swz) << 4));
            r[4 * j4 + 0] = __float_as_uint(x.x), r[4 * j4 + 1] = __float_as_uint(x.y);
            r[4 * j4 + 2] = __float_as_uint(x.z), r[4 * j4 + 3] = __float_as_uint(x.w);
          }
          if (MODE == 2) {
            // hi = x rounded to nearest TF32 (|lo| <= 2^-12 |x|, half of what truncation leaves), written explicitly:
            // the tensor core would truncate the raw bits.  A' = [bf16(lo) x8 | bf16(x) x8] per group of 8 k.
#pragma unroll
            for (int g = 0; g < 2; ++g) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const int k0 = 8 * g + 2 * i;
                const float x0 = __uint_as_float(r[k0]), x1 = __uint_as_float(r[k0 + 1]);
                const uint32_t h0 = (r[k0] + 0x1000u) & 0xFFFFE000u, h1 = (r[k0 + 1] + 0x1000u) & 0xFFFFE000u;
                w[8 * g + i] = ptx::pack_bf16x2(x0 - __uint_as_float(h0), x1 - __uint_as_float(h1));
                w[8 * g + 4 + i] = ptx::pack_bf16x2(x0, x1);
                r[k0] = h0, r[k0 + 1] = h1;
              }
            }
            ptx::tmem_st_32x16(dst + 16 * h, r);
          } else {
            ptx::tmem_st_32x16(dst + 16 * h, r);
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = lo_of(__uint_as_float(r[j]));
          }
          ptx::tmem_st_32x16(dst + kBK + 16 * h, w);
        }
        PROF_LINE(3);
        // the raw tile has been read: hand its slot back to the TMA producer
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(aempty_bar(ia));
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(xf_bar(s));
        PROF_LINE(2);
        ia += 2;
        if (ia >= kAStages) {
          ia -= kAStages;
          pa ^= 1u;
        }
        s += 2;
        if (s >= kStages) {
          s -= kStages;
          ph ^= 1u;
        }
      }
      if (sw == 0 && lane == 0) PROF_DUMP(3);
    } else if (MODE == 1) {
      // ---- B lo plane (MODE 2 has no B split: its B' tile arrives by TMA and these two warps idle)
      const int t = threadIdx.x - (kMmaWarp + 1 + 8) * 32;  // 0..63
      int s = 0;
      uint32_t ph = 0;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
        const int kb_begin = (splits > 1 ? unit / num_tiles : 0) * kb_per_split, kb_end = min(num_kb, kb_begin + kb_per_split);
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          const uint32_t stage = ring_base + s * Cfg::kStageBytes;  // [B raw][B lo]
          ptx::mbar_wait_parked(full_bar(s), ph);
          constexpr int kBVecs = Cfg::kBBytes / 16;
          const uint32_t braw = stage, blo = braw + Cfg::kBBytes;
#pragma unroll 4
          for (int i = t; i < kBVecs; i += 64) {
            const float4 x = ptx::lds_v4(braw + i * 16);
            ptx::sts_v4(blo + i * 16, make_float4(__uint_as_float(lo_of(x.x)), __uint_as_float(lo_of(x.y)),
                                                  __uint_as_float(lo_of(x.z)), __uint_as_float(lo_of(x.w))));
          }
          ptx::fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(xf_bar(s));
          if (++s == kStages) {
            s = 0;
            ph ^= 1u;
          }
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
  if (threadIdx.x == 0) PROF_MARK_LAST(2, 6);  // CTA end
#ifdef RTG_PROF
  if (threadIdx.x == 0 && args.prof) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    args.prof[(blockIdx.x * 4 + 3) * 8 + 6] = gt;
  }
#endif
}

}  // namespace rtg
