// ConvTranspose (reference: src/ops/conv.rs:329-577): the inverse of the im2col convolution.  The reference runs, per
// image, one GemmExecutor product kernel_mat^T [O*Kh*Kw, C] . input [C, H*W] and scatters its columns into the output
// (col2im).  Here the batch is folded into the GEMM's M dimension over the NHWC activations — one launch of the
// tcgen05 contraction: cols[N*H*W, O*Kh*Kw] = x_nhwc[N*H*W, C] . kernel[C, O*Kh*Kw] (the kernel tensor already IS the
// MN-major B operand) — and col2im is a gather: one thread per output element adds the taps that land on it in the
// reference's order (bias, then ky-major / kx-minor; src/ops/conv.rs:341-373), so given the same column values the
// sums are bit-identical.
#include "gemm_host.h"
#include "ops.h"

namespace rtg {

namespace {
struct Col2imGeom {
  int n, o, h, w, kh, kw, oh, ow, sh, sw, pad_t, pad_l;
  int64_t ldc;              // row pitch of the column matrix (O*Kh*Kw)
  int64_t sn, sc, sy, sx;   // element strides of the output
};

__global__ void col2im_gather_kernel(float* __restrict__ out, const float* __restrict__ cols,
                                     const float* __restrict__ bias, Col2imGeom g) {
  const int64_t total = static_cast<int64_t>(g.n) * g.o * g.oh * g.ow;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int ox = static_cast<int>(i % g.ow);
    const int oy = static_cast<int>((i / g.ow) % g.oh);
    const int o = static_cast<int>((i / (static_cast<int64_t>(g.ow) * g.oh)) % g.o);
    const int n = static_cast<int>(i / (static_cast<int64_t>(g.ow) * g.oh * g.o));
    float acc = bias ? bias[o] : 0.f;
    for (int ky = 0; ky < g.kh; ++ky) {
      const int ty = oy + g.pad_t - ky;  // = y * stride_h
      if (ty < 0 || ty % g.sh != 0) continue;
      const int y = ty / g.sh;
      if (y >= g.h) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int tx = ox + g.pad_l - kx;
        if (tx < 0 || tx % g.sw != 0) continue;
        const int x = tx / g.sw;
        if (x >= g.w) continue;
        const int64_t row = (static_cast<int64_t>(n) * g.h + y) * g.w + x;
        acc = __fadd_rn(acc, cols[row * g.ldc + (static_cast<int64_t>(o) * g.kh + ky) * g.kw + kx]);
      }
    }
    out[n * g.sn + o * g.sc + oy * g.sy + ox * g.sx] = acc;
  }
}
}  // namespace

// conv_transpose_output_size_and_padding + the argument checks of conv_transpose()  (src/ops/conv.rs:377-439,441-505)
rtg_status conv_transpose_plan(const rtg_tensor* x, const rtg_tensor* w, const rtg_tensor* bias,
                               const rtg_conv_transpose_attrs* a, ConvTransposePlan* p) {
  if (!x || !w || !a || !p) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  if (x->dtype != RTG_F32 || w->dtype != RTG_F32 || (bias && bias->dtype != RTG_F32))
    return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  int64_t pads[4] = {0, 0, 0, 0}, strides[2] = {1, 1};
  int n_pads = a->n_pads;
  p->is_1d = x->ndim == 3;
  if (p->is_1d) {
    if (w->ndim != 3) return set_error(RTG_ERR_INVALID_VALUE, "kernel must have 3 dims (OCW)");
    if (a->pad_mode == RTG_PAD_FIXED) {  // Padding::expand_1d_to_2d
      if (a->n_pads != 2) return set_error(RTG_ERR_INVALID_VALUE, "expected 2 pad values");
      pads[1] = a->pads[0], pads[3] = a->pads[1];
      n_pads = 4;
    }
    if (a->n_strides != 1) return set_error(RTG_ERR_INVALID_VALUE, "expected 1 stride value");
    strides[1] = a->strides[0];
    p->n = x->shape[0], p->c = x->shape[1], p->h = 1, p->w = x->shape[2];
    p->kc = w->shape[0], p->o = w->shape[1], p->kh = 1, p->kw = w->shape[2];
  } else {
    if (x->ndim != 4) return set_error(RTG_ERR_INVALID_VALUE, "input must have 4 dims (NCHW)");
    if (w->ndim != 4) return set_error(RTG_ERR_INVALID_VALUE, "kernel must have 4 dims (OCHW)");
    for (int i = 0; i < 4; ++i) pads[i] = a->pads[i];
    p->n = x->shape[0], p->c = x->shape[1], p->h = x->shape[2], p->w = x->shape[3];
    p->kc = w->shape[0], p->o = w->shape[1], p->kh = w->shape[2], p->kw = w->shape[3];
  }
  if (bias && bias->ndim != 1) return set_error(RTG_ERR_INVALID_VALUE, "bias must have 1 dims");
  if (p->c != p->kc)
    return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Input channels does not match kernel input channels");
  if (!p->is_1d) {
    if (a->n_strides != 2) return set_error(RTG_ERR_INVALID_VALUE, "expected 2 stride values");
    strides[0] = a->strides[0], strides[1] = a->strides[1];
  }
  p->sh = strides[0], p->sw = strides[1];
  if (p->sh == 0 || p->sw == 0) return set_error(RTG_ERR_INVALID_VALUE, "Strides must be > 0");
  if (p->h == 0 || p->w == 0) return set_error(RTG_ERR_INVALID_VALUE, "Input width and height must be > 0");
  if (a->pad_mode == RTG_PAD_SAME) {
    p->oh = p->h * p->sh, p->ow = p->w * p->sw;
    const int64_t pad_h = (p->h - 1) * p->sh + p->kh - p->oh, pad_w = (p->w - 1) * p->sw + p->kw - p->ow;
    if (pad_h < 0 || pad_w < 0) return set_error(RTG_ERR_INVALID_VALUE, "Input is too small");
    // The reference returns [top, bottom, left, right] here (src/ops/conv.rs:424) and conv_transpose reads the
    // tuple as [top, left, bottom, right] (:511), so col2im's "left" padding is ceil(pad_h / 2).  Results must match
    // the reference, so the same value is used.
    p->pads_out[0] = pad_h / 2, p->pads_out[1] = (pad_h + 1) / 2, p->pads_out[2] = pad_w / 2, p->pads_out[3] = (pad_w + 1) / 2;
  } else {
    if (n_pads != 4) return set_error(RTG_ERR_INVALID_VALUE, "Wrong number of pad values");
    p->oh = (p->h - 1) * p->sh + p->kh - (pads[0] + pads[2]);
    p->ow = (p->w - 1) * p->sw + p->kw - (pads[1] + pads[3]);
    if (p->oh < 0 || p->ow < 0) return set_error(RTG_ERR_INVALID_VALUE, "Input is too small");
    for (int i = 0; i < 4; ++i) p->pads_out[i] = pads[i];
  }
  p->pad_t = p->pads_out[0], p->pad_l = p->pads_out[1];
  return RTG_OK;
}

// x: [N,C,H,W]-shaped device view (any strides), w: dense [C,O,Kh,Kw], out: [N,O,Oh,Ow]-shaped device view.
rtg_status conv_transpose_forward(Context* ctx, const ConvTransposePlan& p, const rtg_tensor& x_in, const float* w_dev,
                                  const float* bias_dev, const rtg_tensor& out_in) {
  rtg_tensor x = x_in, out = out_in;
  if (p.is_1d) {  // lift [N,C,W] -> [N,C,1,W]
    x.ndim = out.ndim = 4;
    x.shape[3] = x_in.shape[2], x.shape[2] = 1, x.strides[3] = x_in.strides[2], x.strides[2] = 0;
    out.shape[3] = out_in.shape[2], out.shape[2] = 1, out.strides[3] = out_in.strides[2], out.strides[2] = 0;
  }
  const int64_t total = p.n * p.o * p.oh * p.ow;
  if (total == 0) return RTG_OK;
  const int64_t M = p.n * p.h * p.w, K = p.c, NC = p.o * p.kh * p.kw;
  if (M > INT32_MAX || NC > INT32_MAX) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "ConvTranspose dims exceed i32");
  // activations as [N*H*W, C]: NHWC-dense inputs are used in place
  Scratch x_nhwc, cols;
  const float* xp = static_cast<const float*>(x.data);
  if (!(is_nhwc_dense(x) && (reinterpret_cast<uintptr_t>(x.data) & 15) == 0 && p.c % 4 == 0)) {
    RTG_TRY(x_nhwc.alloc(ctx, static_cast<size_t>(M) * K * 4));
    rtg_tensor dst = x;
    dst.data = x_nhwc.p;
    dst.strides[0] = p.h * p.w * p.c, dst.strides[1] = 1, dst.strides[2] = p.w * p.c, dst.strides[3] = p.c;
    RTG_TRY(copy_view(ctx, dst, x));
    xp = x_nhwc.f32();
  }
  RTG_TRY(cols.alloc(ctx, static_cast<size_t>(M) * NC * 4));
  Epilogue ep;
  RTG_TRY(gemm_device(ctx, MatView{xp, M, K, K, 1}, MatView{w_dev, K, NC, NC, 1}, cols.f32(), OutMap{NC, 0, 0, 0}, ep));
  Col2imGeom g = {static_cast<int>(p.n), static_cast<int>(p.o), static_cast<int>(p.h), static_cast<int>(p.w),
                  static_cast<int>(p.kh), static_cast<int>(p.kw), static_cast<int>(p.oh), static_cast<int>(p.ow),
                  static_cast<int>(p.sh), static_cast<int>(p.sw), static_cast<int>(p.pad_t), static_cast<int>(p.pad_l),
                  NC, out.strides[0], out.strides[1], out.strides[2], out.strides[3]};
  const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(total, 256), static_cast<int64_t>(ctx->num_sms) * 16));
  col2im_gather_kernel<<<blocks, 256, 0, ctx->stream>>>(static_cast<float*>(out.data), cols.f32(), bias_dev, g);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

}  // namespace rtg
