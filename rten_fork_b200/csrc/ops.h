// Device-level operator entry points (inputs/outputs are rtg_tensor views whose `data` is a DEVICE pointer) and small
// shape helpers.  The C ABI in api_ops.cu stages host tensors and forwards here; the graph executor calls these
// directly.
#pragma once
#include "gemm_host.h"
#include "runtime.h"

namespace rtg {

inline int64_t numel(const rtg_tensor& t) {
  int64_t n = 1;
  for (int i = 0; i < t.ndim; ++i) n *= t.shape[i];
  return n;
}
inline bool is_contiguous(const rtg_tensor& t) {
  for (int i = 0; i < t.ndim; ++i)
    if (t.shape[i] == 0) return true;
  int64_t s = 1;
  for (int i = t.ndim - 1; i >= 0; --i) {
    if (t.shape[i] != 1 && t.strides[i] != s) return false;
    s *= t.shape[i];
  }
  return true;
}
inline void set_contiguous_strides(rtg_tensor& t) {
  int64_t s = 1;
  for (int i = t.ndim - 1; i >= 0; --i) {
    t.strides[i] = s;
    s *= t.shape[i];
  }
}
// 4-D [N,C,H,W]-shaped view whose memory is dense NHWC (channel pitch == C).
inline bool is_nhwc_dense(const rtg_tensor& t) {
  if (t.ndim != 4) return false;
  const int64_t C = t.shape[1], H = t.shape[2], W = t.shape[3];
  return (C == 1 || t.strides[1] == 1) && (W == 1 || t.strides[3] == C) && (H == 1 || t.strides[2] == W * C) &&
         (t.shape[0] == 1 || t.strides[0] == H * W * C);
}
// Span of a view in elements (max offset + 1).
inline int64_t span_of(const rtg_tensor& t) {
  int64_t s = 1;
  for (int i = 0; i < t.ndim; ++i) {
    if (t.shape[i] == 0) return 0;
    s += (t.shape[i] - 1) * t.strides[i];
  }
  return s;
}
inline rtg_tensor make_tensor(void* data, int dtype, int ndim, const int64_t* shape, int device) {
  rtg_tensor t = {};
  t.data = data;
  t.dtype = dtype;
  t.ndim = ndim;
  for (int i = 0; i < ndim; ++i) t.shape[i] = shape[i];
  set_contiguous_strides(t);
  t.device = device;
  return t;
}

// ------------------------------------------------------------------ convolution
struct ConvPlan {            // everything conv() derives from shapes + attributes (src/ops/conv.rs:86-210)
  bool is_1d = false;
  int64_t n, c, h, w;        // input (after the 1-D lift)
  int64_t o, cg, kh, kw;     // kernel [O, C/groups, Kh, Kw]
  int64_t oh, ow;
  int64_t pads[4];           // resolved [top, left, bottom, right]
  int64_t stride_h, stride_w, dil_h, dil_w;
  int64_t groups;
  bool pointwise = false, depthwise = false;
};
rtg_status conv_plan(const rtg_tensor* input, const rtg_tensor* weight, const rtg_tensor* bias,
                     const rtg_conv_attrs* attrs, ConvPlan* plan);

// Pre-arranged device weights of one Conv node (built once per constant weight, or per call for the op-level API).
struct ConvWeights {
  float* wmat = nullptr;     // dense/grouped: [O][Kh*Kw][c_blocks*32]; depthwise: [Kh*Kw][C]
  int kpt = 0;               // c_blocks*32
  bool direct = false;       // pointwise with C % 4 == 0: the OIHW tensor already is the K-major matrix
  bool window = false;       // row-window path (few input channels): wmat is [O][Kh][32]
  // packed copy of wmat for RTG_MATH_TF32_BF16X2 (make_b_pack): two planes [O][wpack.ld], row pitch = wmat's rounded up
  // to 32
  BPack wpack;
  bool persistent = false;   // prepared at load time (Operator::prepare), not by the call that uses it
};
bool conv_window_supported(const ConvPlan& p);
rtg_status conv_prepare_weights(Context* ctx, const ConvPlan& p, const float* w_dev, ConvWeights* out);
void conv_release_weights(Context* ctx, ConvWeights* w);

// x: [N,C,H,W]-shaped device view (NCHW-dense or NHWC-dense fast; anything else is gathered first).
// out / residual: [N,O,Oh,Ow]-shaped device views, NCHW-dense or NHWC-dense.
rtg_status conv_forward(Context* ctx, const ConvPlan& p, const rtg_tensor& x, const ConvWeights& w,
                        const float* bias_dev, const rtg_tensor* residual, int act, float act_lo, float act_hi,
                        const rtg_tensor& out);

// The index-driven layouts the convolution path builds before it contracts, written to `dst` (device) for inspection:
// VirtualIm2Col (src/ops/conv/im2col.rs:75-260) and the layout changes around it.  All are pure gathers: every element
// is a copy of an input element or a zero, so they are compared bit for bit with the oracle's offset tables.
//   kind 0  patch matrix of `group`: [n][oh*ow][kh*kw*kpt], kpt = ceil32(C/groups); column (ky*kw + kx)*kpt + c
//           (im2col_nhwc_kernel over the NHWC copy: the software gather used outside TMA's im2col limits)
//   kind 1  zero-padded NHWC image of the row-window path: [n][hp][wp][cp]   (pad_to_window_layout_kernel)
//   kind 2  NHWC copy: [n][h*w][cp], cp = ceil4(C)                            (nchw_to_nhwc_kernel)
//   kind 3  kind 2 converted back to dense NCHW: [n][c][h][w]                  (nhwc_to_nchw_kernel)
// `dims` receives the shape written (4 entries).
constexpr int kForceIm2colFallback = 1 << 16;  // Context::debug_flags: conv_forward never uses TMA im2col / row windows
rtg_status conv_patches(Context* ctx, int kind, const ConvPlan& p, const rtg_tensor& x, int group, float* dst,
                        size_t dst_floats, int64_t dims[4]);

// ------------------------------------------------------------------ ConvTranspose (conv_transpose.cu)
struct ConvTransposePlan {
  bool is_1d = false;
  int64_t n = 0, c = 0, h = 0, w = 0, kc = 0, o = 0, kh = 0, kw = 0, oh = 0, ow = 0, sh = 1, sw = 1;
  int64_t pads_out[4] = {0, 0, 0, 0};  // the tuple conv_transpose_output_size_and_padding returns
  int64_t pad_t = 0, pad_l = 0;        // what col2im uses
};
rtg_status conv_transpose_plan(const rtg_tensor* x, const rtg_tensor* w, const rtg_tensor* bias,
                               const rtg_conv_transpose_attrs* a, ConvTransposePlan* p);
rtg_status conv_transpose_forward(Context* ctx, const ConvTransposePlan& p, const rtg_tensor& x, const float* w_dev,
                                  const float* bias_dev, const rtg_tensor& out);

// ------------------------------------------------------------------ layout
// dst (any strides) = src (any strides), same shape, f32 or i32 (4-byte elements).
rtg_status copy_view(Context* ctx, const rtg_tensor& dst, const rtg_tensor& src);

// ------------------------------------------------------------------ elementwise
rtg_status unary_forward(Context* ctx, int act, const rtg_tensor& in, float lo, float hi, const rtg_tensor& out);
rtg_status binary_out_shape(const rtg_tensor* a, const rtg_tensor* b, int64_t* shape, int32_t* ndim);
rtg_status binary_forward(Context* ctx, int op, const rtg_tensor& a, const rtg_tensor& b, const rtg_tensor& out);

// ------------------------------------------------------------------ matmul / gemm op
rtg_status matmul_out_shape(const rtg_tensor* a, const rtg_tensor* b, int64_t* shape, int32_t* ndim);
rtg_status matmul_forward(Context* ctx, const rtg_tensor& a, const rtg_tensor& b, const rtg_tensor& out,
                          const Epilogue* fused = nullptr);
rtg_status gemm_op_forward(Context* ctx, const rtg_tensor& a, const rtg_tensor& b, const rtg_tensor* c, float alpha,
                           float beta, bool trans_a, bool trans_b, const rtg_tensor& out,
                           const BPack* b_pack = nullptr);  // b_pack: packed copy of constant weights (GemmOp::prepare)
rtg_status gemm_op_out_shape(const rtg_tensor* a, const rtg_tensor* b, bool trans_a, bool trans_b, int64_t* shape);

// ------------------------------------------------------------------ pooling
rtg_status pool_out_shape(const rtg_tensor* in, const rtg_pool_attrs* attrs, int64_t* shape, int64_t* pads);
rtg_status pool2d_forward(Context* ctx, int kind, const rtg_tensor& in, const rtg_pool_attrs& attrs,
                          const rtg_tensor& out);
rtg_status global_avg_pool_forward(Context* ctx, const rtg_tensor& in, const rtg_tensor& out);

}  // namespace rtg
