// Internal (device-pointer) entry points of the contraction engine.
#pragma once
#include "gemm_sm100.cuh"
#include "runtime.h"

namespace rtg {

struct MatView {      // [rows, cols] f32 matrix view on the device, element strides
  const float* p;
  int64_t rows, cols, rs, cs;
};

// Packed B operand of RTG_MATH_TF32_BF16X2: two K-major planes of [N][ld] 32-bit words (ld = ceil(K/32)*32, zero beyond
// K), `plane` words apart.  Plane 0: B rounded to nearest TF32 (the hi*hi MMA reads it; the tensor core's own
// truncation is then exact).  Plane 1: the bf16 correction operand B': per group of 8 k, four words of bf16 pairs of
// hi(B) followed by four of lo(B) = B - hi(B).  One 3-D TMA box {32, BN, 2} loads both tiles of a k-block.
struct BPack {
  const float* p = nullptr;
  int64_t ld = 0, plane = 0;
};

struct Epilogue {
  float alpha = 1.f, beta = 0.f;
  bool alpha_div = false;     // D = acc / alpha instead of acc * alpha (fused MatMul -> Div by a scalar)
  const float* bias = nullptr;
  int bias_mode = BIAS_NONE;
  const float* R = nullptr;   // residual / beta source
  OutMap rmap = {0, 0, 0, 0};
  int act = ACT_NONE;
  float act_lo = 0.f, act_hi = 0.f;
  // peer copies of the output (other GPUs' buffers mapped into this process, same layout as D): every tile is stored
  // to D and to each of them.  Needs a row-major, 16-byte aligned D (the TMA-store epilogue).
  // RTG_MATH_TF32_BF16X2: the packed B operand of a constant B, written by make_b_pack at load time; p == NULL =
  // computed by a pre-pass of this call
  BPack b_pack;
  bool b_pack_const = false;  // b_pack was written at load time (not by an earlier launch of this call): loads of it need no stream dependency
  int n_peers = 0;
  float* peer_D[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};

// D = epilogue(A[M,K] . B[K,N]).  A and B may have any strides; operands that TMA cannot address directly (neither
// dimension contiguous, unaligned base or pitch) are first gathered into pool scratch by a copy kernel.
rtg_status gemm_device(Context* ctx, MatView a, MatView b, float* D, OutMap dmap, const Epilogue& ep);

// Writes the packed B operand (BPack) of b [K, N] (any strides): plane 0 at out, plane 1 at out + plane.  Needs
// 2 planes of N * ceil(K/32)*32 words.
rtg_status make_b_pack(Context* ctx, MatView b, float* out, int64_t plane);
inline int64_t b_pack_ld(int64_t k) { return (k + 31) / 32 * 32; }

// M <= 8 (the reference's gemv case, src/gemm.rs:651-704, and batch-1..8 classifier heads): one pass over B at HBM
// speed on the CUDA cores (gemv.cu).  gemm_device takes this path when gemv_supported says so.
bool gemv_supported(const MatView& a, const MatView& b, const OutMap& dmap, const Epilogue& ep);
rtg_status gemv_device(Context* ctx, MatView a, MatView b, float* D, OutMap dmap, const Epilogue& ep);

// Batched product in ONE launch: batch index z = z1 * n0 + z0, entry z of A at a.p + z0*a.bs0 + z1*a.bs1 (a stride of
// 0 broadcasts the operand along that batch dimension), D contiguous [n1*n0][M][N].  The operands must be addressable
// by TMA in place (gemm_batched_supported); callers fall back to one gemm_device per entry otherwise.
struct BatchedView {
  const float* p;
  int64_t rows, cols, rs, cs;  // one matrix
  int64_t bs0, bs1;            // element strides of the two batch dimensions
};
bool gemm_batched_supported(const BatchedView& a, const BatchedView& b, int64_t n0, int64_t n1, int64_t N);
rtg_status gemm_batched_device(Context* ctx, BatchedView a, BatchedView b, int64_t n0, int64_t n1, float* D,
                               const Epilogue& ep);

// Implicit-GEMM convolution over an NHWC activation tensor using TMA im2col loads.
struct ConvGeom {
  int n, h, w, c_total;        // input tensor [n, h, w, c_total] NHWC dense
  int c_off, cg;               // channel slice of this group
  int kh, kw, oh, ow;
  int stride_h, stride_w, pad_t, pad_l, pad_b, pad_r, dil_h, dil_w;
};
// wmat: [Og][kh*kw*c_blocks*32] K-major weight matrix for this group (zero padded per tap), Og output channels.
rtg_status conv_gemm_device(Context* ctx, const float* x_nhwc, const ConvGeom& g, const float* wmat, int og,
                            float* D, OutMap dmap, const Epilogue& ep);  // ep.b_pack: packed wmat (or empty)
bool conv_im2col_supported(const ConvGeom& g);

// Row-window convolution (few input channels): x_pad is the zero-padded [n][hp][wp][cp] image written by
// pad_to_window_layout_kernel, wmat the [Og][kh][32] matrix of arrange_window_weights_kernel.  One k-block per kernel
// row; an M tile is 128 consecutive output pixels of ONE output row.
struct WindowGeom {
  int n, hp, wp, cp;
  int kh, kw, oh, ow;
  int stride_h, stride_w, dil_h;
  int c;  // real input channels (algorithmic flops)
};
rtg_status conv_window_gemm_device(Context* ctx, const float* x_pad, const WindowGeom& g, const float* wmat, int og,
                                   float* D, OutMap dmap, const Epilogue& ep);

// Split-K GEMM whose scratch (partial-sum planes, B') is a file-scope device array: allocates nothing (gemm_host.cu).
// a [M,K] and b [K,N] dense row-major device matrices, D row-major with row stride ldd; bias per row or NULL.
rtg_status gemm_splitk_static(Context* ctx, const float* a, int64_t M, int64_t K, const float* b, int64_t N, float* D,
                              int64_t ldd, float alpha, float beta, const float* bias, int splits_req);

// dst[r*dst_ld + c] = src[r*rs + c*cs]
rtg_status copy_strided_2d(Context* ctx, float* dst, int64_t dst_ld, const float* src, int64_t rows, int64_t cols,
                           int64_t rs, int64_t cs);
rtg_status scale_buffer(Context* ctx, float* p, size_t n, float beta);

}  // namespace rtg
