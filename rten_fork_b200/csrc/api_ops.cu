// C ABI for the operator-level entry points.  Host tensors (device == RTG_HOST) are staged through pool buffers with
// cudaMemcpyAsync on the context stream (pinned host memory makes these truly asynchronous); device tensors are used
// in place.  Each function mirrors one reference operator; see include/rten_gpu.h for the citations.
#include "ops.h"

using namespace rtg;

namespace {

struct Staged {
  rtg_tensor t = {};
  Scratch buf;
  bool host = false;
};

rtg_status check_tensor(const rtg_tensor* t, const char* what) {
  if (!t) return set_errorf(RTG_ERR_MISSING_INPUTS, "Required inputs were missing (%s)", what);
  if (t->ndim < 0 || t->ndim > RTG_MAX_DIMS) return set_errorf(RTG_ERR_INVALID_VALUE, "%s: bad ndim", what);
  for (int i = 0; i < t->ndim; ++i)
    if (t->shape[i] < 0 || t->strides[i] < 0) return set_errorf(RTG_ERR_UNSUPPORTED_VALUE, "%s: negative shape/stride", what);
  return RTG_OK;
}

// Make `src` usable on the device: copy host memory up (the whole span of the view, strides preserved).
rtg_status stage_in(Context* ctx, const rtg_tensor* src, Staged* s) {
  s->t = *src;
  s->host = src->device == RTG_HOST;
  if (!s->host) {
    if (src->device != ctx->device) return set_error(RTG_ERR_INVALID_VALUE, "tensor device does not match context");
    return RTG_OK;
  }
  const int64_t span = span_of(*src);
  RTG_TRY(s->buf.alloc(ctx, static_cast<size_t>(std::max<int64_t>(span, 1)) * 4));
  if (span > 0) RTG_CUDA(cudaMemcpyAsync(s->buf.p, src->data, static_cast<size_t>(span) * 4, cudaMemcpyHostToDevice, ctx->stream));
  s->t.data = s->buf.p;
  s->t.device = ctx->device;
  return RTG_OK;
}

// Output staging: host outputs get a contiguous device buffer that is copied back by finish_out().
rtg_status stage_out(Context* ctx, rtg_tensor* dst, Staged* s) {
  s->t = *dst;
  s->host = dst->device == RTG_HOST;
  if (!s->host) {
    if (dst->device != ctx->device) return set_error(RTG_ERR_INVALID_VALUE, "tensor device does not match context");
    return RTG_OK;
  }
  if (!is_contiguous(*dst)) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "host output tensors must be contiguous");
  RTG_TRY(s->buf.alloc(ctx, static_cast<size_t>(std::max<int64_t>(numel(*dst), 1)) * 4));
  s->t.data = s->buf.p;
  s->t.device = ctx->device;
  return RTG_OK;
}
rtg_status finish_out(Context* ctx, rtg_tensor* dst, Staged* s) {
  if (!s->host) return RTG_OK;
  const int64_t n = numel(*dst);
  if (n > 0) RTG_CUDA(cudaMemcpyAsync(dst->data, s->buf.p, static_cast<size_t>(n) * 4, cudaMemcpyDeviceToHost, ctx->stream));
  RTG_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTG_OK;
}

rtg_status expect_shape(const rtg_tensor* out, const int64_t* shape, int ndim) {
  if (out->ndim != ndim) return set_error(RTG_ERR_INCORRECT_OUTPUT_TYPE, "output tensor has the wrong rank");
  for (int i = 0; i < ndim; ++i)
    if (out->shape[i] != shape[i]) return set_error(RTG_ERR_INCORRECT_OUTPUT_TYPE, "output tensor has the wrong shape");
  return RTG_OK;
}

#define CTX_OR_FAIL(ctx_)                                            \
  Context* ctx = reinterpret_cast<Context*>(ctx_);                   \
  if (!ctx) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");   \
  RTG_CUDA(cudaSetDevice(ctx->device));

}  // namespace

extern "C" {

rtg_status rtg_conv_out_shape(const rtg_tensor* input, const rtg_tensor* weight, const rtg_tensor* bias,
                              const rtg_conv_attrs* attrs, int64_t* out_shape, int32_t* out_ndim) {
  if (!attrs || !out_shape || !out_ndim) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  ConvPlan p;
  RTG_TRY(conv_plan(input, weight, bias, attrs, &p));
  if (p.is_1d) {
    out_shape[0] = p.n, out_shape[1] = p.o, out_shape[2] = p.ow;
    *out_ndim = 3;
  } else {
    out_shape[0] = p.n, out_shape[1] = p.o, out_shape[2] = p.oh, out_shape[3] = p.ow;
    *out_ndim = 4;
  }
  return RTG_OK;
}

rtg_status rtg_conv(rtg_ctx* ctx_, const rtg_tensor* input, const rtg_tensor* weight, const rtg_tensor* bias,
                    const rtg_conv_attrs* attrs, const rtg_tensor* residual, int32_t act, float act_lo, float act_hi,
                    rtg_tensor* output) {
  CTX_OR_FAIL(ctx_);
  if (!attrs || !output) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_TRY(check_tensor(input, "input"));
  RTG_TRY(check_tensor(weight, "kernel"));
  if (bias) RTG_TRY(check_tensor(bias, "bias"));
  ConvPlan p;
  RTG_TRY(conv_plan(input, weight, bias, attrs, &p));
  int64_t oshape[4] = {p.n, p.o, p.is_1d ? p.ow : p.oh, p.ow};
  RTG_TRY(expect_shape(output, oshape, p.is_1d ? 3 : 4));
  Staged x, w, b, r, o;
  RTG_TRY(stage_in(ctx, input, &x));
  RTG_TRY(stage_in(ctx, weight, &w));
  if (bias) RTG_TRY(stage_in(ctx, bias, &b));
  if (residual) RTG_TRY(stage_in(ctx, residual, &r));
  RTG_TRY(stage_out(ctx, output, &o));
  // weights and bias must be dense for the arrangement kernels
  Scratch wc, bc;
  if (!is_contiguous(w.t)) {
    RTG_TRY(wc.alloc(ctx, static_cast<size_t>(numel(w.t)) * 4));
    rtg_tensor t = w.t;
    t.data = wc.p;
    set_contiguous_strides(t);
    RTG_TRY(copy_view(ctx, t, w.t));
    w.t = t;
  }
  if (bias && !is_contiguous(b.t)) {
    RTG_TRY(bc.alloc(ctx, static_cast<size_t>(numel(b.t)) * 4));
    rtg_tensor t = b.t;
    t.data = bc.p;
    set_contiguous_strides(t);
    RTG_TRY(copy_view(ctx, t, b.t));
    b.t = t;
  }
  ConvWeights cw;
  RTG_TRY(conv_prepare_weights(ctx, p, static_cast<const float*>(w.t.data), &cw));
  rtg_status st = conv_forward(ctx, p, x.t, cw, bias ? static_cast<const float*>(b.t.data) : nullptr,
                               residual ? &r.t : nullptr, act, act_lo, act_hi, o.t);
  conv_release_weights(ctx, &cw);
  RTG_TRY(st);
  return finish_out(ctx, output, &o);
}

rtg_status rtg_conv_patches(rtg_ctx* ctx_, int32_t kind, const rtg_tensor* input, const rtg_tensor* weight,
                            const rtg_conv_attrs* attrs, int32_t group, float* out_host, size_t out_floats,
                            int64_t* out_dims) {
  CTX_OR_FAIL(ctx_);
  if (!attrs || !out_host || !out_dims) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_TRY(check_tensor(input, "input"));
  RTG_TRY(check_tensor(weight, "kernel"));
  ConvPlan p;
  RTG_TRY(conv_plan(input, weight, nullptr, attrs, &p));
  Staged x;
  RTG_TRY(stage_in(ctx, input, &x));
  Scratch d;
  RTG_TRY(d.alloc(ctx, std::max<size_t>(out_floats, 1) * 4));
  RTG_TRY(conv_patches(ctx, kind, p, x.t, group, d.f32(), out_floats, out_dims));
  const size_t n = static_cast<size_t>(out_dims[0] * out_dims[1] * out_dims[2] * out_dims[3]);
  if (n > 0) RTG_CUDA(cudaMemcpyAsync(out_host, d.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  RTG_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTG_OK;
}

rtg_status rtg_conv_transpose_out_shape(const rtg_tensor* input, const rtg_tensor* weight, const rtg_tensor* bias,
                                        const rtg_conv_transpose_attrs* attrs, int64_t* out_shape, int32_t* out_ndim) {
  if (!attrs || !out_shape || !out_ndim) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  ConvTransposePlan p;
  RTG_TRY(conv_transpose_plan(input, weight, bias, attrs, &p));
  out_shape[0] = p.n, out_shape[1] = p.o;
  if (p.is_1d) {
    out_shape[2] = p.ow;
    *out_ndim = 3;
  } else {
    out_shape[2] = p.oh, out_shape[3] = p.ow;
    *out_ndim = 4;
  }
  return RTG_OK;
}

rtg_status rtg_conv_transpose(rtg_ctx* ctx_, const rtg_tensor* input, const rtg_tensor* weight, const rtg_tensor* bias,
                              const rtg_conv_transpose_attrs* attrs, rtg_tensor* output) {
  CTX_OR_FAIL(ctx_);
  if (!attrs || !output) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_TRY(check_tensor(input, "input"));
  RTG_TRY(check_tensor(weight, "kernel"));
  if (bias) RTG_TRY(check_tensor(bias, "bias"));
  ConvTransposePlan p;
  RTG_TRY(conv_transpose_plan(input, weight, bias, attrs, &p));
  int64_t oshape[4] = {p.n, p.o, p.is_1d ? p.ow : p.oh, p.ow};
  RTG_TRY(expect_shape(output, oshape, p.is_1d ? 3 : 4));
  Staged x, w, b, o;
  RTG_TRY(stage_in(ctx, input, &x));
  RTG_TRY(stage_in(ctx, weight, &w));
  if (bias) RTG_TRY(stage_in(ctx, bias, &b));
  RTG_TRY(stage_out(ctx, output, &o));
  Scratch wc, bc;
  if (!is_contiguous(w.t) || (reinterpret_cast<uintptr_t>(w.t.data) & 15) != 0) {
    RTG_TRY(wc.alloc(ctx, static_cast<size_t>(std::max<int64_t>(numel(w.t), 1)) * 4));
    rtg_tensor t = w.t;
    t.data = wc.p;
    set_contiguous_strides(t);
    RTG_TRY(copy_view(ctx, t, w.t));
    w.t = t;
  }
  if (bias && !is_contiguous(b.t)) {
    RTG_TRY(bc.alloc(ctx, static_cast<size_t>(std::max<int64_t>(numel(b.t), 1)) * 4));
    rtg_tensor t = b.t;
    t.data = bc.p;
    set_contiguous_strides(t);
    RTG_TRY(copy_view(ctx, t, b.t));
    b.t = t;
  }
  RTG_TRY(conv_transpose_forward(ctx, p, x.t, static_cast<const float*>(w.t.data),
                                 bias ? static_cast<const float*>(b.t.data) : nullptr, o.t));
  return finish_out(ctx, output, &o);
}

rtg_status rtg_matmul_out_shape(const rtg_tensor* a, const rtg_tensor* b, int64_t* out_shape, int32_t* out_ndim) {
  return matmul_out_shape(a, b, out_shape, out_ndim);
}

rtg_status rtg_matmul(rtg_ctx* ctx_, const rtg_tensor* a, const rtg_tensor* b, rtg_tensor* output) {
  CTX_OR_FAIL(ctx_);
  RTG_TRY(check_tensor(a, "a"));
  RTG_TRY(check_tensor(b, "b"));
  if (!output) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  int64_t os[RTG_MAX_DIMS];
  int32_t ond;
  RTG_TRY(matmul_out_shape(a, b, os, &ond));
  RTG_TRY(expect_shape(output, os, ond));
  Staged sa, sb, so;
  RTG_TRY(stage_in(ctx, a, &sa));
  RTG_TRY(stage_in(ctx, b, &sb));
  RTG_TRY(stage_out(ctx, output, &so));
  RTG_TRY(matmul_forward(ctx, sa.t, sb.t, so.t, nullptr));
  return finish_out(ctx, output, &so);
}

rtg_status rtg_gemm_op(rtg_ctx* ctx_, const rtg_tensor* a, const rtg_tensor* b, const rtg_tensor* c, float alpha,
                       float beta, int32_t transpose_a, int32_t transpose_b, rtg_tensor* output) {
  CTX_OR_FAIL(ctx_);
  RTG_TRY(check_tensor(a, "a"));
  RTG_TRY(check_tensor(b, "b"));
  if (c) RTG_TRY(check_tensor(c, "c"));
  if (!output) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  int64_t os[2];
  RTG_TRY(gemm_op_out_shape(a, b, transpose_a != 0, transpose_b != 0, os));
  RTG_TRY(expect_shape(output, os, 2));
  Staged sa, sb, sc, so;
  RTG_TRY(stage_in(ctx, a, &sa));
  RTG_TRY(stage_in(ctx, b, &sb));
  if (c) RTG_TRY(stage_in(ctx, c, &sc));
  RTG_TRY(stage_out(ctx, output, &so));
  RTG_TRY(gemm_op_forward(ctx, sa.t, sb.t, c ? &sc.t : nullptr, alpha, beta, transpose_a != 0, transpose_b != 0, so.t));
  return finish_out(ctx, output, &so);
}

rtg_status rtg_unary(rtg_ctx* ctx_, int32_t act, const rtg_tensor* input, float lo, float hi, rtg_tensor* output) {
  CTX_OR_FAIL(ctx_);
  RTG_TRY(check_tensor(input, "input"));
  if (!output) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  if (input->dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  RTG_TRY(expect_shape(output, input->shape, input->ndim));
  Staged si, so;
  RTG_TRY(stage_in(ctx, input, &si));
  RTG_TRY(stage_out(ctx, output, &so));
  RTG_TRY(unary_forward(ctx, act, si.t, lo, hi, so.t));
  return finish_out(ctx, output, &so);
}

rtg_status rtg_binary_out_shape(const rtg_tensor* a, const rtg_tensor* b, int64_t* out_shape, int32_t* out_ndim) {
  return binary_out_shape(a, b, out_shape, out_ndim);
}

rtg_status rtg_binary(rtg_ctx* ctx_, int32_t op, const rtg_tensor* a, const rtg_tensor* b, rtg_tensor* output) {
  CTX_OR_FAIL(ctx_);
  RTG_TRY(check_tensor(a, "a"));
  RTG_TRY(check_tensor(b, "b"));
  if (!output) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  if (op < RTG_BIN_ADD || op > RTG_BIN_POW) return set_error(RTG_ERR_INVALID_VALUE, "unknown binary op");
  int64_t os[RTG_MAX_DIMS];
  int32_t ond;
  RTG_TRY(binary_out_shape(a, b, os, &ond));
  RTG_TRY(expect_shape(output, os, ond));
  Staged sa, sb, so;
  RTG_TRY(stage_in(ctx, a, &sa));
  RTG_TRY(stage_in(ctx, b, &sb));
  RTG_TRY(stage_out(ctx, output, &so));
  RTG_TRY(binary_forward(ctx, op, sa.t, sb.t, so.t));
  return finish_out(ctx, output, &so);
}

rtg_status rtg_pool2d_out_shape(const rtg_tensor* input, const rtg_pool_attrs* attrs, int64_t* out_shape) {
  if (!attrs || !out_shape) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  return pool_out_shape(input, attrs, out_shape, nullptr);
}

rtg_status rtg_pool2d(rtg_ctx* ctx_, int32_t kind, const rtg_tensor* input, const rtg_pool_attrs* attrs,
                      rtg_tensor* output) {
  CTX_OR_FAIL(ctx_);
  RTG_TRY(check_tensor(input, "input"));
  if (!attrs || !output) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  int64_t os[4];
  RTG_TRY(pool_out_shape(input, attrs, os, nullptr));
  RTG_TRY(expect_shape(output, os, 4));
  Staged si, so;
  RTG_TRY(stage_in(ctx, input, &si));
  RTG_TRY(stage_out(ctx, output, &so));
  RTG_TRY(pool2d_forward(ctx, kind, si.t, *attrs, so.t));
  return finish_out(ctx, output, &so);
}

rtg_status rtg_global_average_pool(rtg_ctx* ctx_, const rtg_tensor* input, rtg_tensor* output) {
  CTX_OR_FAIL(ctx_);
  RTG_TRY(check_tensor(input, "input"));
  if (!output) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  if (input->ndim != 4) return set_error(RTG_ERR_INVALID_VALUE, "input must have 4 dims (NCHW)");
  int64_t os[4] = {input->shape[0], input->shape[1], 1, 1};
  RTG_TRY(expect_shape(output, os, 4));
  Staged si, so;
  RTG_TRY(stage_in(ctx, input, &si));
  RTG_TRY(stage_out(ctx, output, &so));
  RTG_TRY(global_avg_pool_forward(ctx, si.t, so.t));
  return finish_out(ctx, output, &so);
}

}  // extern "C"
