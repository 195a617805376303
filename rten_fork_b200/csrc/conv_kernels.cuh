// CUDA-core kernels around the convolution path: layout changes, weight pre-arrangement, explicit im2col (fallback for
// geometries outside the TMA im2col limits) and the direct depthwise convolution.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "vecmath.cuh"

namespace rtg {

// ---------------------------------------------------------------- NCHW <-> NHWC
// src: channels [c_off, c_off + c) of an NCHW tensor given by element strides (sn, sc, sp) where the H*W plane is
// contiguous-with-stride sp (sp = 1 for dense NCHW).  dst: [n, hw, cp] dense, channels >= c zero-filled.
// One 32(pixels) x 32(channels) tile per block, transposed through shared memory so both sides are coalesced.
__global__ void nchw_to_nhwc_kernel(float* __restrict__ dst, const float* __restrict__ src, int c, int hw, int cp,
                                    int64_t sn, int64_t sc, int64_t sp, int c_off) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
  const float* s = src + n * sn + static_cast<int64_t>(c_off) * sc;
  for (int i = ty; i < 32; i += 8) {
    const int ch = c0 + i, p = p0 + tx;
    tile[i][tx] = (ch < c && p < hw) ? s[ch * sc + p * sp] : 0.f;
  }
  __syncthreads();
  float* d = dst + static_cast<int64_t>(n) * hw * cp;
  for (int i = ty; i < 32; i += 8) {
    const int p = p0 + i, ch = c0 + tx;
    if (p < hw && ch < cp) d[static_cast<int64_t>(p) * cp + ch] = tile[tx][i];
  }
}

// src: [n, hw, cp] dense NHWC (first c channels valid) -> dst NCHW-like with strides (dn, dc, dp), channel offset.
__global__ void nhwc_to_nchw_kernel(float* __restrict__ dst, const float* __restrict__ src, int c, int hw, int cp,
                                    int64_t dn, int64_t dc, int64_t dp, int c_off) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const float* s = src + static_cast<int64_t>(n) * hw * cp;
  for (int i = ty; i < 32; i += 8) {
    const int p = p0 + i, ch = c0 + tx;
    tile[i][tx] = (p < hw && ch < c) ? s[static_cast<int64_t>(p) * cp + ch] : 0.f;
  }
  __syncthreads();
  float* d = dst + n * dn + static_cast<int64_t>(c_off) * dc;
  for (int i = ty; i < 32; i += 8) {
    const int ch = c0 + i, p = p0 + tx;
    if (ch < c && p < hw) d[ch * dc + p * dp] = tile[tx][i];
  }
}

// ---------------------------------------------------------------- weights
// OIHW [O, Cg, Kh, Kw] -> per group g: [Og][Kh*Kw][c_blocks*32], zero padded.  This is the GPU counterpart of
// prepack_a on the reshaped kernel matrix (reference: src/ops/conv.rs:231-241); K order is (ky, kx, c) to match
// the order in which TMA im2col delivers activation columns.
__global__ void arrange_conv_weights_kernel(float* __restrict__ dst, const float* __restrict__ w, int o_total, int cg,
                                            int khw, int kpt /* c_blocks*32 */) {
  const int64_t total = static_cast<int64_t>(o_total) * khw * kpt;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int ci = static_cast<int>(i % kpt);
    const int tap = static_cast<int>((i / kpt) % khw);
    const int o = static_cast<int>(i / (static_cast<int64_t>(kpt) * khw));
    dst[i] = ci < cg ? w[(static_cast<int64_t>(o) * cg + ci) * khw + tap] : 0.f;
  }
}

// ---------------------------------------------------------------- row-window layout (first-layer convolutions)
// Few input channels (C <= 8) make the per-tap channel block of the im2col path almost empty (C = 3 fills 3 of 32
// columns).  The row-window path instead reads, for every output pixel and kernel row, ONE contiguous run of
// kw * cp floats from a zero-padded NHWC image (cp = C rounded up to 4), so a k-block is a whole kernel row.
//
// dst: [n][hp][wp][cp] with the image at (pad_t, pad_l), zeros elsewhere (borders, channel padding, right slack).
// src: x[n*sn + c*sc + y*sy + x*sx] (any layout).  One thread per destination pixel; index-driven (copies or zeros).
__global__ void pad_to_window_layout_kernel(float* __restrict__ dst, const float* __restrict__ src, int n, int c,
                                            int h, int w, int cp, int hp, int wp, int pad_t, int pad_l, int64_t sn,
                                            int64_t sc, int64_t sy, int64_t sx) {
  const int64_t total = static_cast<int64_t>(n) * hp * wp;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int px = static_cast<int>(i % wp);
    const int py = static_cast<int>((i / wp) % hp);
    const int img = static_cast<int>(i / (static_cast<int64_t>(wp) * hp));
    const int x = px - pad_l, y = py - pad_t;
    const bool in = x >= 0 && x < w && y >= 0 && y < h;
    const float* s = src + img * sn + y * sy + x * sx;
    float* d = dst + i * cp;
    for (int ch = 0; ch < cp; ch += 4) {
      float4 v;
      v.x = (in && ch + 0 < c) ? s[(ch + 0) * sc] : 0.f;
      v.y = (in && ch + 1 < c) ? s[(ch + 1) * sc] : 0.f;
      v.z = (in && ch + 2 < c) ? s[(ch + 2) * sc] : 0.f;
      v.w = (in && ch + 3 < c) ? s[(ch + 3) * sc] : 0.f;
      *reinterpret_cast<float4*>(d + ch) = v;
    }
  }
}

// OIHW [O, C, Kh, Kw] -> [O][Kh][32]: column j of kernel row ky holds w[o][j % cp][ky][j / cp], zero where the
// channel or tap does not exist (those columns meet padding channels / the pixels right of the window).
__global__ void arrange_window_weights_kernel(float* __restrict__ dst, const float* __restrict__ w, int o_total, int c,
                                              int kh, int kw, int cp) {
  const int64_t total = static_cast<int64_t>(o_total) * kh * 32;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int j = static_cast<int>(i % 32);
    const int ky = static_cast<int>((i / 32) % kh);
    const int o = static_cast<int>(i / (32 * static_cast<int64_t>(kh)));
    const int kx = j / cp, ch = j % cp;
    dst[i] = (kx < kw && ch < c) ? w[((static_cast<int64_t>(o) * c + ch) * kh + ky) * kw + kx] : 0.f;
  }
}

// ---------------------------------------------------------------- explicit im2col (fallback)
struct Im2colGeom {
  int n, h, w, cp;     // input NHWC dense, channel pitch cp
  int c_off, cg;
  int kh, kw, oh, ow;
  int stride_h, stride_w, pad_t, pad_l, dil_h, dil_w;
  int kpt;             // c_blocks*32 columns per tap
};
// dst: [n*oh*ow][kh*kw*kpt]; same column order as the arranged weights.  Index-driven: values are copies or zeros.
// reference: VirtualIm2Col::pack_b_impl masked gather (src/ops/conv/im2col.rs:190-260).
__global__ void im2col_nhwc_kernel(float* __restrict__ dst, const float* __restrict__ x, Im2colGeom g) {
  const int64_t kmat = static_cast<int64_t>(g.kh) * g.kw * g.kpt;
  const int64_t total = static_cast<int64_t>(g.n) * g.oh * g.ow * kmat;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t col = i % kmat, m = i / kmat;
    const int ci = static_cast<int>(col % g.kpt);
    const int tap = static_cast<int>(col / g.kpt);
    const int ky = tap / g.kw, kx = tap % g.kw;
    const int ox = static_cast<int>(m % g.ow);
    const int oy = static_cast<int>((m / g.ow) % g.oh);
    const int n = static_cast<int>(m / (static_cast<int64_t>(g.ow) * g.oh));
    const int iy = oy * g.stride_h + ky * g.dil_h - g.pad_t;
    const int ix = ox * g.stride_w + kx * g.dil_w - g.pad_l;
    float v = 0.f;
    if (ci < g.cg && iy >= 0 && iy < g.h && ix >= 0 && ix < g.w)
      v = x[((static_cast<int64_t>(n) * g.h + iy) * g.w + ix) * g.cp + g.c_off + ci];
    dst[i] = v;
  }
}

// ---------------------------------------------------------------- depthwise
struct DwGeom {
  int n, h, w, c;      // NHWC dense, channel pitch cp
  int cp_in, cp_out;
  int kh, kw, oh, ow;
  int stride_h, stride_w, pad_t, pad_l, dil_h, dil_w;
  int act;
  float act_lo, act_hi;
};
// Direct depthwise convolution over NHWC, 4 channels per thread (float4), one output pixel row-segment per thread.
// Accumulation order equals the reference's (bias first, then ky-major / kx-minor, separate multiply and add:
// src/ops/conv/depthwise.rs:49-120), so results are bit-identical to it.
// wt: [kh*kw][c] (tap-major, transposed once from the reference's [c,1,kh,kw]).
// HBM-bound: per output element 4 B written and ~4 B/stride^2 read (taps re-read from L1/L2).
template <int VEC>
__global__ void depthwise_nhwc_kernel(float* __restrict__ out, const float* __restrict__ x,
                                      const float* __restrict__ wt, const float* __restrict__ bias, DwGeom g) {
  const int cv = (g.c + VEC - 1) / VEC;
  const int64_t total = static_cast<int64_t>(g.n) * g.oh * g.ow * cv;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % cv) * VEC;
    const int64_t pix = i / cv;
    const int ox = static_cast<int>(pix % g.ow);
    const int oy = static_cast<int>((pix / g.ow) % g.oh);
    const int n = static_cast<int>(pix / (static_cast<int64_t>(g.ow) * g.oh));
    float acc[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) acc[v] = (bias != nullptr && c0 + v < g.c) ? bias[c0 + v] : 0.f;
    for (int ky = 0; ky < g.kh; ++ky) {
      const int iy = oy * g.stride_h + ky * g.dil_h - g.pad_t;
      if (iy < 0 || iy >= g.h) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int ix = ox * g.stride_w + kx * g.dil_w - g.pad_l;
        if (ix < 0 || ix >= g.w) continue;
        const float* xp = x + ((static_cast<int64_t>(n) * g.h + iy) * g.w + ix) * g.cp_in + c0;
        const float* wp = wt + static_cast<int64_t>(ky * g.kw + kx) * g.c + c0;
        if (VEC == 4) {
          const float4 xv = *reinterpret_cast<const float4*>(xp);
          const float4 wv = *reinterpret_cast<const float4*>(wp);
          acc[0] = __fadd_rn(acc[0], __fmul_rn(xv.x, wv.x));
          acc[1 % VEC] = __fadd_rn(acc[1 % VEC], __fmul_rn(xv.y, wv.y));
          acc[2 % VEC] = __fadd_rn(acc[2 % VEC], __fmul_rn(xv.z, wv.z));
          acc[3 % VEC] = __fadd_rn(acc[3 % VEC], __fmul_rn(xv.w, wv.w));
        } else {
#pragma unroll
          for (int v = 0; v < VEC; ++v)
            if (c0 + v < g.c) acc[v] = __fadd_rn(acc[v], __fmul_rn(xp[v], wp[v]));
        }
      }
    }
    float* op = out + pix * g.cp_out + c0;
    if (VEC == 4) {
      *reinterpret_cast<float4*>(op) = make_float4(apply_act(acc[0], g.act, g.act_lo, g.act_hi),
                                                   apply_act(acc[1 % VEC], g.act, g.act_lo, g.act_hi),
                                                   apply_act(acc[2 % VEC], g.act, g.act_lo, g.act_hi),
                                                   apply_act(acc[3 % VEC], g.act, g.act_lo, g.act_hi));
    } else {
#pragma unroll
      for (int v = 0; v < VEC; ++v)
        if (c0 + v < g.c) op[v] = apply_act(acc[v], g.act, g.act_lo, g.act_hi);
    }
  }
}

// 3x3 depthwise convolution (dilation 1, stride 1 or 2), the MobileNetV2 shapes: one thread owns 4 channels x TX
// output columns x `ty` output rows.  The nine weight vectors stay in registers for the whole strip, and the input
// rows slide through a three-row register window (one new input row per output row at stride 1, two at stride 2), so
// an output vector costs ~1.5 (stride 1) / ~4.5 (stride 2) 16-byte loads instead of 18.  The generic kernel above
// issued one load per tap and weight and ran at ~1.2 TB/s, bound by L1 and index arithmetic rather than HBM.
// Same summation order as the reference (bias, then ky-major / kx-minor, separate multiply and add; taps that fall
// into the padding are skipped, not added as zeros), so results stay bit-identical to it.
template <int STRIDE, int TX>
__global__ void __launch_bounds__(128) depthwise3x3_nhwc_kernel(float* __restrict__ out, const float* __restrict__ x,
                                                                const float* __restrict__ wt,
                                                                const float* __restrict__ bias, DwGeom g, int ty) {
  constexpr int IW = (TX - 1) * STRIDE + 3;  // input columns under a strip
  const int cv = g.c >> 2;
  const int xs_n = (g.ow + TX - 1) / TX, ys_n = (g.oh + ty - 1) / ty;
  const int64_t total = static_cast<int64_t>(g.n) * ys_n * xs_n * cv;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int c0 = static_cast<int>(i % cv) << 2;
    int64_t t = i / cv;
    const int xs = static_cast<int>(t % xs_n);
    t /= xs_n;
    const int ys = static_cast<int>(t % ys_n);
    const int n = static_cast<int>(t / ys_n);
    const int ox0 = xs * TX, oy0 = ys * ty, oy1 = min(oy0 + ty, g.oh);
    const int ix0 = ox0 * STRIDE - g.pad_l;

    float4 w[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) w[k] = __ldg(reinterpret_cast<const float4*>(wt + static_cast<int64_t>(k) * g.c + c0));
    const float4 b = bias != nullptr ? __ldg(reinterpret_cast<const float4*>(bias + c0)) : make_float4(0.f, 0.f, 0.f, 0.f);
    bool colv[IW];
#pragma unroll
    for (int j = 0; j < IW; ++j) colv[j] = ix0 + j >= 0 && ix0 + j < g.w;

    float4 r[3][IW];
    bool rowv[3];
    const float* xn = x + static_cast<int64_t>(n) * g.h * g.w * g.cp_in + c0;
    auto load_row = [&](float4 (&dst)[IW], bool& valid, int iy) {
      valid = iy >= 0 && iy < g.h;
      const float* xr = xn + (static_cast<int64_t>(iy) * g.w + ix0) * g.cp_in;
#pragma unroll
      for (int j = 0; j < IW; ++j)
        dst[j] = (valid && colv[j]) ? *reinterpret_cast<const float4*>(xr + static_cast<int64_t>(j) * g.cp_in)
                                    : make_float4(0.f, 0.f, 0.f, 0.f);
    };
    int iy = oy0 * STRIDE - g.pad_t;
    load_row(r[0], rowv[0], iy);
    load_row(r[1], rowv[1], iy + 1);
    for (int oy = oy0; oy < oy1; ++oy, iy += STRIDE) {
      load_row(r[2], rowv[2], iy + 2);
      float* op = out + ((static_cast<int64_t>(n) * g.oh + oy) * g.ow + ox0) * g.cp_out + c0;
#pragma unroll
      for (int j = 0; j < TX; ++j) {
        float4 acc = b;
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {
            const float4 xv = r[ky][j * STRIDE + kx];
            const float4 wv = w[ky * 3 + kx];
            if (rowv[ky] && colv[j * STRIDE + kx]) {
              acc.x = __fadd_rn(acc.x, __fmul_rn(xv.x, wv.x));
              acc.y = __fadd_rn(acc.y, __fmul_rn(xv.y, wv.y));
              acc.z = __fadd_rn(acc.z, __fmul_rn(xv.z, wv.z));
              acc.w = __fadd_rn(acc.w, __fmul_rn(xv.w, wv.w));
            }
          }
        }
        if (ox0 + j < g.ow)
          *reinterpret_cast<float4*>(op + static_cast<int64_t>(j) * g.cp_out) =
              make_float4(apply_act(acc.x, g.act, g.act_lo, g.act_hi), apply_act(acc.y, g.act, g.act_lo, g.act_hi),
                          apply_act(acc.z, g.act, g.act_lo, g.act_hi), apply_act(acc.w, g.act, g.act_lo, g.act_hi));
      }
      // slide the window down by STRIDE input rows
      if (STRIDE == 1) {
#pragma unroll
        for (int j = 0; j < IW; ++j) r[0][j] = r[1][j], r[1][j] = r[2][j];
        rowv[0] = rowv[1], rowv[1] = rowv[2];
      } else {
#pragma unroll
        for (int j = 0; j < IW; ++j) r[0][j] = r[2][j];
        rowv[0] = rowv[2];
        if (oy + 1 < oy1) load_row(r[1], rowv[1], iy + 3);
      }
    }
  }
}

// [c][khw] -> [khw][c]
__global__ void transpose_dw_weights_kernel(float* __restrict__ dst, const float* __restrict__ w, int c, int khw) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < c * khw) {
    const int ch = i % c, tap = i / c;
    dst[i] = w[ch * khw + tap];
  }
}

}  // namespace rtg
