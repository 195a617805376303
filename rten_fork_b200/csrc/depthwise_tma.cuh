// 3x3 depthwise convolution (dilation 1, stride 1 or 2) over dense NHWC, input rows staged through shared memory by TMA
// (src/ops/conv/depthwise.rs:49-203 is the reference's kernel; same summation order, so results are bit-identical).
//
// HBM-bound: 4 B read per input element and 4 B written per output element.  The register-window kernel in
// conv_kernels.cuh needs 156 registers per thread, fits 12 warps per SM and hides the load latency by occupancy alone:
// ~2.1-2.9 TB/s (ncu: long_scoreboard 35 %, issue slots 42 % busy, half of them address / predicate arithmetic).  Here
// the loads are asynchronous instead: a CTA owns a band of TW output columns x CB channels of one image and walks down
// its rows; one TMA box {CB channels, input columns under the band, 1 row} per input row lands in a ring of kRing row
// buffers (its own mbarrier each), so kRing - 3 rows per CTA are in flight whatever the occupancy.  The padding comes
// from the TMA unit (out-of-bounds elements are zero-filled; coordinates may be negative).  A thread owns 4 channels x
// 4 output columns of an output row and reads the three input rows under it from shared memory (4.5 / 6.75 16-byte
// shared loads per output vector at stride 1 / 2, a sixth of the shared-memory bandwidth at HBM speed); no register
// window, so the kernel is ~100 registers and one copy of the tap loop: the first version kept the window in
// registers with three statically rotated copies of the loop, 170-200 registers and 90 K instructions, and ran at the
// register-window kernel's speed (ncu: 19 % of the stalls were instruction fetch).
//
// Exactness: the reference skips taps that fall into the padding instead of adding zeros.  Adding the zeros the TMA
// unit filled in is the same bits whenever the channel's weights are finite and its bias is not -0.0 (see `plain` in the
// kernel), which a warp checks once; otherwise every tap is predicated (one setp per tap vector, predicated adds; no
// branches).  Multiply and add stay separate instructions in both forms.
#pragma once
#include "conv_kernels.cuh"
#include "sm100_ptx.cuh"
#include <type_traits>

namespace rtg {

struct DwTile {
  int cvs;           // channel vectors (float4) per CTA: CB = 4 * cvs channels
  int tw;            // output columns per CTA (strips of kDwTX columns; 128 threads = (tw / kDwTX) * cvs)
  int iwb;           // input columns per box
  int row_bytes;     // bytes per ring slot (box bytes rounded up to 128)
  int box_bytes;
  int ch_blocks, bands, row_bands, rows_per_job;
  int force_pred;    // (RTG_DW_PRED=1) always skip padding taps by predicate
  unsigned long long one2;  // (1.0f, 1.0f): see dw_tap
};
constexpr int kDwTX = 4;  // dw_taps below is written out for 4
constexpr int kDwRing = 6;  // row buffers per CTA (5 / 4 buffers with 5 / 6 CTAs per SM measured slower)
constexpr int kDwThreads = 128;

// Four channels of one tap: acc += x * w with the multiply and the add rounded separately, as the reference does, on
// sm_100's packed fp32 pipe: FMUL2 for the two channel pairs, then acc = fma(acc, 1.0, product) — exactly RN(acc + product)
// — as FFMA2.  (A packed add after a packed multiply is contracted into one FFMA2 by ptxas even with .rn, which would
// round once; `one` is (1.0f, 1.0f) from the launch parameters so that it cannot simplify the fma back into an add.)
// Half the issue slots of four FMUL + four FADD.  PRED: skipped unless bit BIT of `mask` is set (the validity of the
// input column under this tap, zero for a row outside the image).
template <int BIT, bool PRED>
__device__ __forceinline__ void dw_tap(unsigned long long (&acc)[2], const float4 x, const unsigned long long (&w)[2],
                                       const uint32_t mask, const unsigned long long one) {
  if (!PRED) {  // padding taps add the TMA unit's zeros: see `plain` in the kernel for when that is the same bits
    asm("{\n\t.reg .b64 x0, x1, m0, m1;\n\t"
        "mov.b64 x0, {%2, %3};\n\tmov.b64 x1, {%4, %5};\n\t"
        "mul.rn.f32x2 m0, x0, %6;\n\tmul.rn.f32x2 m1, x1, %7;\n\t"
        "fma.rn.f32x2 %0, %0, %8, m0;\n\tfma.rn.f32x2 %1, %1, %8, m1;\n\t}"
        : "+l"(acc[0]), "+l"(acc[1])
        : "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w), "l"(w[0]), "l"(w[1]), "l"(one));
    return;
  }
  asm("{\n\t.reg .pred p;\n\t.reg .b32 m;\n\t.reg .b64 x0, x1, m0, m1;\n\t"
      "and.b32 m, %9, %10;\n\tsetp.ne.u32 p, m, 0;\n\t"
      "mov.b64 x0, {%2, %3};\n\tmov.b64 x1, {%4, %5};\n\t"
      "mul.rn.f32x2 m0, x0, %6;\n\tmul.rn.f32x2 m1, x1, %7;\n\t"
      "@p fma.rn.f32x2 %0, %0, %8, m0;\n\t@p fma.rn.f32x2 %1, %1, %8, m1;\n\t}"
      : "+l"(acc[0]), "+l"(acc[1])
      : "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w), "l"(w[0]), "l"(w[1]), "l"(one), "r"(mask), "n"(1u << BIT));
}

// f32::clamp of the reference (src/ops/unary_elementwise.rs:325-339): NaN stays NaN, -0.0 stays -0.0; two compares and
// two selects, no branches
// the same for operands where max / min agree with it bit for bit (v is never -0.0, hi is not -0.0, no NaN bounds)
__device__ __forceinline__ float dw_clip_fast(float v, float lo, float hi) {
  float r;
  asm("{\n\t.reg .f32 t;\n\tmax.NaN.f32 t, %1, %2;\n\tmin.NaN.f32 %0, t, %3;\n\t}" : "=f"(r) : "f"(v), "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ float dw_clip(float v, float lo, float hi) {
  float r;
  asm("{\n\t.reg .pred p, q;\n\tsetp.lt.f32 p, %1, %2;\n\tsetp.gt.f32 q, %1, %3;\n\t"
      "selp.f32 %0, %3, %1, q;\n\tselp.f32 %0, %2, %0, p;\n\t}"
      : "=f"(r)
      : "f"(v), "f"(lo), "f"(hi));
  return r;
}

// the 3 taps of one kernel row for the 4 output columns of a strip (compile-time column indices)
template <int STRIDE, int IW, bool PRED>
__device__ __forceinline__ void dw_taps(unsigned long long (&acc)[kDwTX][2], const float4 (&row)[IW],
                                        const unsigned long long (&w0)[2], const unsigned long long (&w1)[2],
                                        const unsigned long long (&w2)[2], const uint32_t rmask,
                                        const unsigned long long one) {
  dw_tap<0 * STRIDE + 0, PRED>(acc[0], row[0 * STRIDE + 0], w0, rmask, one);
  dw_tap<1 * STRIDE + 0, PRED>(acc[1], row[1 * STRIDE + 0], w0, rmask, one);
  dw_tap<2 * STRIDE + 0, PRED>(acc[2], row[2 * STRIDE + 0], w0, rmask, one);
  dw_tap<3 * STRIDE + 0, PRED>(acc[3], row[3 * STRIDE + 0], w0, rmask, one);
  dw_tap<0 * STRIDE + 1, PRED>(acc[0], row[0 * STRIDE + 1], w1, rmask, one);
  dw_tap<1 * STRIDE + 1, PRED>(acc[1], row[1 * STRIDE + 1], w1, rmask, one);
  dw_tap<2 * STRIDE + 1, PRED>(acc[2], row[2 * STRIDE + 1], w1, rmask, one);
  dw_tap<3 * STRIDE + 1, PRED>(acc[3], row[3 * STRIDE + 1], w1, rmask, one);
  dw_tap<0 * STRIDE + 2, PRED>(acc[0], row[0 * STRIDE + 2], w2, rmask, one);
  dw_tap<1 * STRIDE + 2, PRED>(acc[1], row[1 * STRIDE + 2], w2, rmask, one);
  dw_tap<2 * STRIDE + 2, PRED>(acc[2], row[2 * STRIDE + 2], w2, rmask, one);
  dw_tap<3 * STRIDE + 2, PRED>(acc[3], row[3 * STRIDE + 2], w2, rmask, one);
}

// ACT: 0 none, 1 Relu, 2 Clip, 3 any other activation (apply_act's switch).  CVS: channel vectors per CTA (DwTile::cvs), a
// template parameter so that the shared-memory loads of a row use immediate offsets (runtime strides cost ~100 address
// instructions per output row, a fifth of the loop).
template <int STRIDE, int ACT, int CVS>
__global__ void __launch_bounds__(kDwThreads) depthwise3x3_tma_kernel(const __grid_constant__ CUtensorMap tmX,
                                                                       float* __restrict__ out,
                                                                       const float* __restrict__ wt,
                                                                       const float* __restrict__ bias, const DwGeom g,
                                                                       const DwTile t) {
  constexpr int TX = kDwTX;
  constexpr int IW = (TX - 1) * STRIDE + 3;  // input columns under a strip
  extern __shared__ __align__(128) uint8_t dw_smem[];
  __shared__ __align__(8) uint64_t bars[2 * kDwRing];  // full[kDwRing], empty[kDwRing]
  const uint32_t ring = ptx::smem_u32(dw_smem);
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto full_bar = [&](int k) { return bar0 + 8u * (k % kDwRing); };
  auto empty_bar = [&](int k) { return bar0 + 8u * (kDwRing + k % kDwRing); };

  int job = blockIdx.x;
  const int cb = job % t.ch_blocks;
  job /= t.ch_blocks;
  const int band = job % t.bands;
  job /= t.bands;
  const int rb = job % t.row_bands;
  const int n = job / t.row_bands;

  const int tid = threadIdx.x;
  const int cv = tid % CVS, strip = tid / CVS;
  const int c0 = (cb * CVS + cv) << 2;
  const int ox0 = band * t.tw + strip * TX;
  // (a shape whose strips x channel vectors do not fill the 128 threads leaves the last threads without a strip)
  const bool active = c0 < g.c && ox0 < g.ow && strip * TX < t.tw;
  const int oy0 = rb * t.rows_per_job, oy1 = min(g.oh, oy0 + t.rows_per_job);
  const int iy0 = oy0 * STRIDE - g.pad_t;         // first input row of the job
  const int n_in = (oy1 - oy0 - 1) * STRIDE + 3;  // input rows the job reads
  const int bx_c = cb * CVS * 4, bx_w = band * t.tw * STRIDE - g.pad_l;

  if (tid == 0) {
    for (int s = 0; s < kDwRing; ++s) {
      ptx::mbar_init(bar0 + 8u * s, 1);
      ptx::mbar_init(bar0 + 8u * (kDwRing + s), kDwThreads / 32);  // one arrival per warp
    }
    ptx::fence_mbar_init();
  }
  __syncthreads();
  // the input may come from the kernel launched just before (programmatic dependent launch)
  ptx::grid_dependency_wait();
  ptx::grid_launch_dependents();
  auto load_row = [&](int k) {  // thread 0: input row k of the job -> its ring slot
    ptx::mbar_expect_tx(full_bar(k), t.box_bytes);
    ptx::tma_load_4d(ring + (k % kDwRing) * t.row_bytes, &tmX, full_bar(k), bx_c, bx_w, iy0 + k, n);
  };
  if (tid == 0) {
    ptx::prefetch_tmap(&tmX);
    for (int k = 0; k < kDwRing && k < n_in; ++k) load_row(k);
  }

  float4 w[9];
  float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
  if (active) {
#pragma unroll
    for (int k = 0; k < 9; ++k) w[k] = __ldg(reinterpret_cast<const float4*>(wt + static_cast<int64_t>(k) * g.c + c0));
    if (bias != nullptr) b = __ldg(reinterpret_cast<const float4*>(bias + c0));
  }
  // Adding a padding tap's zero-filled input instead of skipping the tap gives the same bits unless the product is not a
  // zero (non-finite weight) or the accumulator is -0.0 (only possible if the bias is -0.0: x + y = -0.0 needs both);
  // max / min give f32::clamp's bits unless v is -0.0 (ditto), `hi` is -0.0 or a bound is NaN.  Decided per warp.
  bool plain_t = true;
  if (active) {
#pragma unroll
    for (int k = 0; k < 9; ++k)
      plain_t = plain_t && isfinite(w[k].x) && isfinite(w[k].y) && isfinite(w[k].z) && isfinite(w[k].w);
    plain_t = plain_t && __float_as_uint(b.x) != 0x80000000u && __float_as_uint(b.y) != 0x80000000u &&
              __float_as_uint(b.z) != 0x80000000u && __float_as_uint(b.w) != 0x80000000u;
    if (ACT == 2)
      plain_t = plain_t && __float_as_uint(g.act_hi) != 0x80000000u && g.act_lo == g.act_lo && g.act_hi == g.act_hi;
  }
  const bool plain = __all_sync(0xffffffffu, plain_t) && !t.force_pred;
  // packed (f32x2) copies of the weights and the bias, and (1.0f, 1.0f) for dw_tap
  unsigned long long w2[9][2], b2[2];
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    asm("mov.b64 %0, {%1, %2};" : "=l"(w2[k][0]) : "f"(w[k].x), "f"(w[k].y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(w2[k][1]) : "f"(w[k].z), "f"(w[k].w));
  }
  asm("mov.b64 %0, {%1, %2};" : "=l"(b2[0]) : "f"(b.x), "f"(b.y));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b2[1]) : "f"(b.z), "f"(b.w));
  const unsigned long long one = t.one2;
  // validity of the input columns under this strip (global input column = bx_w + strip * TX * STRIDE + j)
  const int ix0 = bx_w + strip * TX * STRIDE;
  uint32_t colmask = 0;
#pragma unroll
  for (int j = 0; j < IW; ++j) colmask |= (ix0 + j >= 0 && ix0 + j < g.w) ? 1u << j : 0u;
  const float4* sp0 = reinterpret_cast<const float4*>(dw_smem) + (strip * TX * STRIDE) * CVS + cv;
  const int slot_vecs = t.row_bytes >> 4;

  // Ring positions advance incrementally (the ring is not a power of two): slot / phase of the next row to wait for,
  // slot of input row k0, slot / phase of the next row to release.
  int w_k = 0, w_slot = 0;
  uint32_t w_ph = 0;
  int s0 = 0;
  int e_k = 0, e_slot = 0;
  uint32_t e_ph = 0;
  float* op = out + ((static_cast<int64_t>(n) * g.oh + oy0) * g.ow + ox0) * g.cp_out + c0;
  const int64_t op_row = static_cast<int64_t>(g.ow) * g.cp_out;
  constexpr int cstride = CVS;  // float4 elements between neighbouring input columns
  int iy = oy0 * STRIDE - g.pad_t;
  // one output row of this thread.  PRED: padding taps are skipped by predicate (always exact); otherwise they add the
  // zeros the TMA unit filled in, and Clip is max / min
  auto row_body = [&](auto pred_tag) {
    constexpr bool PRED = decltype(pred_tag)::value;
    unsigned long long acc2[TX][2];
#pragma unroll
    for (int j = 0; j < TX; ++j) acc2[j][0] = b2[0], acc2[j][1] = b2[1];
    int slot = s0;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const float4* sp = sp0 + slot * slot_vecs;
      if (++slot == kDwRing) slot = 0;
      float4 row[IW];
#pragma unroll
      for (int j = 0; j < IW; ++j) row[j] = sp[j * cstride];
      const uint32_t rmask = (PRED && iy + ky >= 0 && iy + ky < g.h) ? colmask : 0u;
      dw_taps<STRIDE, IW, PRED>(acc2, row, w2[ky * 3], w2[ky * 3 + 1], w2[ky * 3 + 2], rmask, one);
    }
#pragma unroll
    for (int j = 0; j < TX; ++j) {
      float acc[4];
      asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[0]), "=f"(acc[1]) : "l"(acc2[j][0]));
      asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[2]), "=f"(acc[3]) : "l"(acc2[j][1]));
      if (ACT == 1) {
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[i] = fmaxf(acc[i], 0.f);
      } else if (ACT == 2) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          acc[i] = PRED ? dw_clip(acc[i], g.act_lo, g.act_hi) : dw_clip_fast(acc[i], g.act_lo, g.act_hi);
      } else if (ACT == 3) {
        apply_act_array(acc, g.act, g.act_lo, g.act_hi);
      }
      if (ox0 + j < g.ow)
        *reinterpret_cast<float4*>(op + static_cast<int64_t>(j) * g.cp_out) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    }
  };
  for (int oy = oy0; oy < oy1; ++oy, iy += STRIDE, op += op_row) {
    const int k0 = (oy - oy0) * STRIDE;  // the output row reads input rows k0, k0 + 1, k0 + 2 of the job
    // thread 0 refills the slots of the rows the previous output row released, once every warp is done with them
    if (tid == 0) {
      for (; e_k < k0; ++e_k) {
        if (e_k + kDwRing < n_in) {
          ptx::mbar_wait(bar0 + 8u * (kDwRing + e_slot), e_ph);
          ptx::mbar_expect_tx(bar0 + 8u * e_slot, t.box_bytes);
          ptx::tma_load_4d(ring + e_slot * t.row_bytes, &tmX, bar0 + 8u * e_slot, bx_c, bx_w, iy0 + e_k + kDwRing, n);
        }
        if (++e_slot == kDwRing) e_slot = 0, e_ph ^= 1u;
      }
    }
    for (; w_k <= k0 + 2; ++w_k) {
      ptx::mbar_wait(bar0 + 8u * w_slot, w_ph);
      if (++w_slot == kDwRing) w_slot = 0, w_ph ^= 1u;
    }
    if (active) {
      if (plain) {
        row_body(std::false_type{});
      } else {
        row_body(std::true_type{});
      }
    }
    // the rows the next output row no longer reads go back to the producer
    __syncwarp();
    if ((tid & 31) == 0) {
#pragma unroll
      for (int i = 0; i < STRIDE; ++i) {
        int sl = s0 + i;
        if (sl >= kDwRing) sl -= kDwRing;
        ptx::mbar_arrive(bar0 + 8u * (kDwRing + sl));
      }
    }
    s0 += STRIDE;
    if (s0 >= kDwRing) s0 -= kDwRing;
  }
}

}  // namespace rtg
