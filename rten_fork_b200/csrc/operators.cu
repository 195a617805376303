// Operator implementations for the graph executor + the OpRegistry readers that build them from `.rten` operator
// nodes (reference: src/op_registry.rs:385-817 impl_read_op!, src/ops/*.rs Operator impls).
#include <cstring>
#include "graph.h"
#include "operators.h"

namespace rtg {

// ---------------------------------------------------------------- helpers
static bool nhwc_like(const rtg_tensor& t) { return t.ndim == 4 && is_nhwc_dense(t) && !is_contiguous(t); }

static rtg_status alloc_like(Context* ctx, const rtg_tensor& like, Value* out) {
  return alloc_value(ctx, like.dtype, like.ndim, like.shape, nhwc_like(like), out);
}

// A contiguous device copy of `v` (or `v` itself when it already is contiguous).  `tmp` keeps the copy alive.
static rtg_status contiguous_view(Context* ctx, const Value& v, Value* tmp, rtg_tensor* view) {
  if (is_contiguous(v.t)) {
    *view = v.t;
    return RTG_OK;
  }
  RTG_TRY(alloc_value(ctx, v.t.dtype, v.t.ndim, v.t.shape, false, tmp));
  RTG_TRY(copy_view(ctx, tmp->t, v.t));
  *view = tmp->t;
  return RTG_OK;
}

// ---------------------------------------------------------------- Conv
rtg_status ConvOp::partial_plan(const rtg_tensor& w, ConvPlan* p) const {
  // What conv() can decide from the kernel shape and attributes alone (used to pre-arrange constant weights).
  if (w.ndim != 4 && w.ndim != 3) return set_error(RTG_ERR_INVALID_VALUE, "kernel must have 4 dims (OCHW)");
  *p = ConvPlan();
  p->o = w.shape[0], p->cg = w.shape[1];
  p->kh = w.ndim == 4 ? w.shape[2] : 1, p->kw = w.ndim == 4 ? w.shape[3] : w.shape[2];
  p->groups = attrs.groups;
  p->c = p->cg * attrs.groups;
  // strides / dilations as conv_plan resolves them (1-D convolutions carry theirs in the W slot)
  const bool one_d = w.ndim == 3;
  p->stride_h = one_d ? 1 : (attrs.n_strides > 0 ? attrs.strides[0] : 1);
  p->stride_w = one_d ? (attrs.n_strides > 0 ? attrs.strides[0] : 1) : (attrs.n_strides > 1 ? attrs.strides[1] : 1);
  p->dil_h = one_d ? 1 : (attrs.n_dilations > 0 ? attrs.dilations[0] : 1);
  p->dil_w = one_d ? (attrs.n_dilations > 0 ? attrs.dilations[0] : 1) : (attrs.n_dilations > 1 ? attrs.dilations[1] : 1);
  bool unit = true, nopad = attrs.pad_mode == RTG_PAD_SAME;
  // the counts come from the file: more entries than the arrays hold are reported when the plan is made (conv_plan)
  for (int i = 0; i < std::min(attrs.n_strides, 2); ++i) unit = unit && attrs.strides[i] == 1;
  for (int i = 0; i < std::min(attrs.n_dilations, 2); ++i) unit = unit && attrs.dilations[i] == 1;
  if (attrs.pad_mode == RTG_PAD_FIXED) {
    nopad = true;
    for (int i = 0; i < std::min(attrs.n_pads, 4); ++i) nopad = nopad && attrs.pads[i] == 0;
  }
  p->pointwise = p->kh == 1 && p->kw == 1 && nopad && attrs.groups == 1 && unit;
  p->depthwise = !p->pointwise && attrs.groups > 0 && p->c == p->o && attrs.groups == p->c;
  return RTG_OK;
}

rtg_status ConvOp::prepare(Context* ctx, const OpInputs& consts) {
  const Value* w = consts.get(1);
  if (!w || !w->t.data || w->t.dtype != RTG_F32 || !is_contiguous(w->t)) return RTG_OK;
  ConvPlan p;
  if (partial_plan(w->t, &p) != RTG_OK) return RTG_OK;  // run() reports the real error
  RTG_TRY(conv_prepare_weights(ctx, p, static_cast<const float*>(w->t.data), &weights));
  prepared = true;
  weights.persistent = true;
  prepared_pointwise = p.pointwise;
  prepared_depthwise = p.depthwise;
  prepared_from = w->t.data;
  return RTG_OK;
}

void ConvOp::release(Context* ctx) {
  if (prepared) conv_release_weights(ctx, &weights);
  prepared = false;
}

rtg_status ConvOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *x, *w;
  RTG_TRY(in.require(0, &x));
  RTG_TRY(in.require(1, &w));
  const Value* b = in.get(2);
  const Value* res = has_residual ? in.get(3) : nullptr;
  if (has_residual && !res) return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  ConvPlan p;
  RTG_TRY(conv_plan(&x->t, &w->t, b ? &b->t : nullptr, &attrs, &p));
  int64_t oshape[4] = {p.n, p.o, p.is_1d ? p.ow : p.oh, p.ow};
  // The load-time fusion Conv -> Add (-> activation) is only an epilogue when the other Add operand has exactly the
  // convolution's output shape and the layer is not depthwise (whose kernel has no residual input).  Anything else —
  // a broadcast operand such as [N, C, 1, 1], a depthwise layer — runs as the reference does: convolution, then the
  // Add operator with numpy broadcasting, then the activation.
  bool unfused_tail = false;
  if (res) {
    const int ond = p.is_1d ? 3 : 4;
    bool same = res->t.ndim == ond && res->t.dtype == RTG_F32;
    for (int i = 0; i < ond && same; ++i) same = res->t.shape[i] == oshape[i];
    unfused_tail = !same || p.depthwise;
  }
  Value y;
  RTG_TRY(alloc_value(ctx, RTG_F32, p.is_1d ? 3 : 4, oshape, !p.is_1d, &y));
  Value wtmp, btmp;
  rtg_tensor wv, bv = {};
  rtg_status st = contiguous_view(ctx, *w, &wtmp, &wv);
  if (st == RTG_OK && b) st = contiguous_view(ctx, *b, &btmp, &bv);
  ConvWeights local;
  const ConvWeights* cw = &weights;
  const bool use_prepared = prepared && prepared_from == w->t.data && prepared_pointwise == p.pointwise &&
                            prepared_depthwise == p.depthwise &&
                            weights.window == (!p.pointwise && !p.depthwise && conv_window_supported(p));
  if (st == RTG_OK && !use_prepared) {
    st = conv_prepare_weights(ctx, p, static_cast<const float*>(wv.data), &local);
    cw = &local;
  }
  if (st == RTG_OK)
    st = conv_forward(ctx, p, x->t, *cw, b ? static_cast<const float*>(bv.data) : nullptr,
                      (res && !unfused_tail) ? &res->t : nullptr, unfused_tail ? ACT_NONE : act, act_lo, act_hi, y.t);
  if (!use_prepared) conv_release_weights(ctx, &local);
  release_value(ctx, &wtmp);
  release_value(ctx, &btmp);
  if (st == RTG_OK && unfused_tail) {
    BinaryOp add(RTG_BIN_ADD, "Add");
    OpInputs ain;
    ain.v = {&y, res};
    std::vector<Value> sum;
    st = add.run(ctx, ain, &sum);
    if (st == RTG_OK) {
      release_value(ctx, &y);
      y = sum[0];
      if (act != ACT_NONE) st = unary_forward(ctx, act, y.t, act_lo, act_hi, y.t);
    }
  }
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- ConvTranspose
rtg_status ConvTransposeOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *x, *w;
  RTG_TRY(in.require(0, &x));
  RTG_TRY(in.require(1, &w));
  const Value* b = in.get(2);
  ConvTransposePlan p;
  RTG_TRY(conv_transpose_plan(&x->t, &w->t, b ? &b->t : nullptr, &attrs, &p));
  int64_t oshape[4] = {p.n, p.o, p.is_1d ? p.ow : p.oh, p.ow};
  Value y;
  RTG_TRY(alloc_value(ctx, RTG_F32, p.is_1d ? 3 : 4, oshape, false, &y));
  Value wtmp, btmp;
  rtg_tensor wv, bv = {};
  rtg_status st = contiguous_view(ctx, *w, &wtmp, &wv);
  if (st == RTG_OK && b) st = contiguous_view(ctx, *b, &btmp, &bv);
  if (st == RTG_OK)
    st = conv_transpose_forward(ctx, p, x->t, static_cast<const float*>(wv.data),
                                b ? static_cast<const float*>(bv.data) : nullptr, y.t);
  release_value(ctx, &wtmp);
  release_value(ctx, &btmp);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- unary
rtg_status UnaryOp::bounds(Context* ctx, const Value* lo_v, const Value* hi_v, float* lo, float* hi) const {
  *lo = act_lo, *hi = act_hi;
  if (act != ACT_CLIP) return RTG_OK;
  // Clip takes min / max as optional scalar inputs (src/ops/unary_elementwise.rs:341-383)
  *lo = -3.40282347e+38f, *hi = 3.40282347e+38f;
  std::vector<float> v;
  if (lo_v) {
    RTG_TRY(host_floats(ctx, *lo_v, &v));
    if (v.size() != 1) return set_error(RTG_ERR_INVALID_VALUE, "Expected scalar");
    *lo = v[0];
  }
  if (hi_v) {
    RTG_TRY(host_floats(ctx, *hi_v, &v));
    if (v.size() != 1) return set_error(RTG_ERR_INVALID_VALUE, "Expected scalar");
    *hi = v[0];
  }
  return RTG_OK;
}

rtg_status UnaryOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  if (x->t.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  float lo, hi;
  RTG_TRY(bounds(ctx, in.get(1), in.get(2), &lo, &hi));
  Value y;
  RTG_TRY(alloc_like(ctx, x->t, &y));
  rtg_status st = unary_forward(ctx, act, x->t, lo, hi, y.t);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

rtg_status UnaryOp::run_in_place(Context* ctx, Value* io, const OpInputs& rest) {
  if (io->t.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  float lo, hi;
  RTG_TRY(bounds(ctx, rest.get(0), rest.get(1), &lo, &hi));
  io->host_f32.reset();
  return unary_forward(ctx, act, io->t, lo, hi, io->t);
}

// ---------------------------------------------------------------- binary
rtg_status BinaryOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *a, *b;
  RTG_TRY(in.require(0, &a));
  RTG_TRY(in.require(1, &b));
  if (a->t.dtype != b->t.dtype) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  int64_t shape[RTG_MAX_DIMS];
  int32_t nd;
  RTG_TRY(binary_out_shape(&a->t, &b->t, shape, &nd));
  if (a->t.dtype == RTG_I32) {
    // i32 operands are shape arithmetic (the reference's operators are generic over f32 / i32,
    // src/ops/binary_elementwise.rs:395-560): small tensors, evaluated on their host mirrors and re-uploaded with one.
    int64_t total = 1;
    for (int i = 0; i < nd; ++i) total *= shape[i];
    if (total > 4096 || op == RTG_BIN_POW)
      return set_error(RTG_ERR_UNSUPPORTED_VALUE, "integer binary operators are implemented for shape-sized tensors only");
    std::vector<int64_t> av, bv;
    RTG_TRY(host_ints(ctx, *a, &av));
    RTG_TRY(host_ints(ctx, *b, &bv));
    auto res = std::make_shared<std::vector<int32_t>>(static_cast<size_t>(total));
    int64_t idx[RTG_MAX_DIMS] = {0};
    for (int64_t e = 0; e < total; ++e) {
      int64_t ia = 0, ib = 0, sa = 1, sb = 1;  // broadcast source offsets (right-aligned shapes, contiguous mirrors)
      for (int d = nd - 1; d >= 0; --d) {
        const int da = d - (nd - a->t.ndim), db = d - (nd - b->t.ndim);
        if (da >= 0) {
          if (a->t.shape[da] != 1) ia += idx[d] * sa;
          sa *= a->t.shape[da];
        }
        if (db >= 0) {
          if (b->t.shape[db] != 1) ib += idx[d] * sb;
          sb *= b->t.shape[db];
        }
      }
      const int32_t x = static_cast<int32_t>(av[ia]), y = static_cast<int32_t>(bv[ib]);
      int32_t r = 0;
      switch (op) {
        case RTG_BIN_ADD: r = static_cast<int32_t>(static_cast<uint32_t>(x) + static_cast<uint32_t>(y)); break;
        case RTG_BIN_SUB: r = static_cast<int32_t>(static_cast<uint32_t>(x) - static_cast<uint32_t>(y)); break;
        case RTG_BIN_MUL: r = static_cast<int32_t>(static_cast<uint32_t>(x) * static_cast<uint32_t>(y)); break;
        case RTG_BIN_DIV:
          if (y == 0) return set_error(RTG_ERR_INVALID_VALUE, "integer division by zero");
          r = (x == INT32_MIN && y == -1) ? INT32_MIN : x / y;
          break;
      }
      (*res)[e] = r;
      for (int d = nd - 1; d >= 0; --d) {
        if (++idx[d] < shape[d]) break;
        idx[d] = 0;
      }
    }
    Value yv;
    RTG_TRY(alloc_value(ctx, RTG_I32, nd, shape, false, &yv));
    if (total > 0) {
      cudaError_t e = cudaMemcpyAsync(yv.t.data, res->data(), static_cast<size_t>(total) * 4, cudaMemcpyHostToDevice, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // `res` may not outlive the copy otherwise
      if (e != cudaSuccess) {
        release_value(ctx, &yv);
        return set_errorf(RTG_ERR_CUDA, "upload failed: %s", cudaGetErrorString(e));
      }
    }
    yv.host_i32 = res;
    out->push_back(yv);
    return RTG_OK;
  }
  // keep the NHWC layout of a full-size operand
  bool nhwc = false;
  if (nd == 4) {
    for (const Value* v : {a, b}) {
      bool full = v->t.ndim == 4;
      for (int i = 0; i < 4 && full; ++i) full = v->t.shape[i] == shape[i];
      if (full && nhwc_like(v->t)) nhwc = true;
    }
  }
  Value y;
  RTG_TRY(alloc_value(ctx, RTG_F32, nd, shape, nhwc, &y));
  rtg_status st = binary_forward(ctx, op, a->t, b->t, y.t);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

rtg_status BinaryOp::run_in_place(Context* ctx, Value* io, const OpInputs& rest) {
  // binary_op_in_place (src/ops/binary_elementwise.rs:261-322) applies only when `io` already has the output shape.
  const Value* other;
  RTG_TRY(rest.require(0, &other));
  int64_t shape[RTG_MAX_DIMS];
  int32_t nd;
  RTG_TRY(binary_out_shape(&io->t, &other->t, shape, &nd));
  bool same = nd == io->t.ndim && io->t.dtype == RTG_F32 && other->t.dtype == RTG_F32;
  for (int i = 0; i < nd && same; ++i) same = shape[i] == io->t.shape[i];
  if (!same) {
    OpInputs in;
    in.v = {io, other};
    std::vector<Value> outs;
    RTG_TRY(run(ctx, in, &outs));
    release_value(ctx, io);
    *io = outs[0];
    return RTG_OK;
  }
  io->host_f32.reset();
  // For commutative operators the executor may hand us the second operand as `io`; the result is the same.
  return binary_forward(ctx, op, io->t, other->t, io->t);
}

// ---------------------------------------------------------------- pooling
rtg_status PoolOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  Value y;
  if (global) {
    if (x->t.ndim != 4) return set_error(RTG_ERR_INVALID_VALUE, "input must have 4 dims (NCHW)");
    int64_t shape[4] = {x->t.shape[0], x->t.shape[1], 1, 1};
    RTG_TRY(alloc_value(ctx, RTG_F32, 4, shape, false, &y));
    rtg_status st = global_avg_pool_forward(ctx, x->t, y.t);
    if (st != RTG_OK) {
      release_value(ctx, &y);
      return st;
    }
  } else {
    int64_t shape[4];
    RTG_TRY(pool_out_shape(&x->t, &attrs, shape, nullptr));
    RTG_TRY(alloc_value(ctx, RTG_F32, 4, shape, nhwc_like(x->t), &y));
    rtg_status st = pool2d_forward(ctx, kind, x->t, attrs, y.t);
    if (st != RTG_OK) {
      release_value(ctx, &y);
      return st;
    }
  }
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- Flatten / Reshape / Identity
static rtg_status resolve_axis(int ndim, int64_t axis, int* out) {
  if (axis < -ndim || axis > ndim) return set_error(RTG_ERR_INVALID_VALUE, "Axis is invalid");
  *out = static_cast<int>(axis < 0 ? axis + ndim : axis);
  return RTG_OK;
}

rtg_status ReshapeOp::target_shape(Context* ctx, const rtg_tensor& x, const Value* shape_v, int64_t* shape,
                                   int* nd) const {
  if (flatten) {
    int ax;
    if (axis < -x.ndim || axis > x.ndim) return set_error(RTG_ERR_INVALID_VALUE, "Axis is invalid");
    ax = static_cast<int>(axis < 0 ? axis + x.ndim : axis);
    int64_t outer = 1, inner = 1;
    for (int i = 0; i < x.ndim; ++i) (i < ax ? outer : inner) *= x.shape[i];
    shape[0] = outer, shape[1] = inner;
    *nd = 2;
    return RTG_OK;
  }
  if (!shape_v) return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  if (shape_v->t.ndim != 1) return set_error(RTG_ERR_INVALID_VALUE, "shape must have 1 dims");
  std::vector<int64_t> req;
  RTG_TRY(host_ints(ctx, *shape_v, &req));
  if (req.size() > RTG_MAX_DIMS) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "too many dimensions");
  // resolve_shape (src/ops/layout.rs:203-273): 0 copies the input dim unless allow_zero, one -1 is inferred
  const int64_t total = numel(x);
  int64_t known = 1;
  int infer = -1;
  for (size_t i = 0; i < req.size(); ++i) {
    int64_t d = req[i];
    if (d == -1) {
      if (infer >= 0) return set_error(RTG_ERR_INVALID_VALUE, "Multiple dimensions in new shape set to -1");
      infer = static_cast<int>(i);
      shape[i] = 1;
      continue;
    }
    if (d < -1) return set_error(RTG_ERR_INVALID_VALUE, "Invalid dimension size in shape");
    if (d == 0 && !allow_zero) {
      if (static_cast<int>(i) >= x.ndim) return set_error(RTG_ERR_INVALID_VALUE, "Zero dim has no corresponding input dim");
      d = x.shape[i];
    }
    shape[i] = d;
    known *= d;
  }
  if (infer >= 0) {
    if (known == 0 || total % known != 0)
      return set_error(RTG_ERR_INVALID_VALUE, "Input length must be a multiple of specified dimensions");
    shape[infer] = total / known;
    known *= shape[infer];
  }
  if (known != total) return set_error(RTG_ERR_INVALID_VALUE, "New shape must have same total elements as current shape");
  *nd = static_cast<int>(req.size());
  return RTG_OK;
}

rtg_status ReshapeOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  int64_t shape[RTG_MAX_DIMS];
  int nd;
  RTG_TRY(target_shape(ctx, x->t, in.get(1), shape, &nd));
  Value y;
  RTG_TRY(alloc_value(ctx, x->t.dtype, x->t.ndim, x->t.shape, false, &y));
  rtg_status st = copy_view(ctx, y.t, x->t);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  y.t.ndim = nd;
  for (int i = 0; i < nd; ++i) y.t.shape[i] = shape[i];
  set_contiguous_strides(y.t);
  y.host_i32 = x->host_i32, y.host_f32 = x->host_f32;
  out->push_back(y);
  return RTG_OK;
}

rtg_status ReshapeOp::run_in_place(Context* ctx, Value* io, const OpInputs& rest) {
  int64_t shape[RTG_MAX_DIMS];
  int nd;
  RTG_TRY(target_shape(ctx, io->t, rest.get(0), shape, &nd));
  if (!is_contiguous(io->t)) {  // e.g. an NHWC activation: re-order once, then the reshape is free
    Value y;
    RTG_TRY(alloc_value(ctx, io->t.dtype, io->t.ndim, io->t.shape, false, &y));
    rtg_status st = copy_view(ctx, y.t, io->t);
    if (st != RTG_OK) {
      release_value(ctx, &y);
      return st;
    }
    y.host_i32 = io->host_i32, y.host_f32 = io->host_f32;
    release_value(ctx, io);
    *io = y;
  }
  io->t.ndim = nd;
  for (int i = 0; i < nd; ++i) io->t.shape[i] = shape[i];
  set_contiguous_strides(io->t);
  return RTG_OK;
}

rtg_status IdentityOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  Value y;
  RTG_TRY(alloc_like(ctx, x->t, &y));
  rtg_status st = copy_view(ctx, y.t, x->t);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  y.host_i32 = x->host_i32, y.host_f32 = x->host_f32;
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- Gemm / MatMul
rtg_status GemmOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *a, *b;
  RTG_TRY(in.require(0, &a));
  RTG_TRY(in.require(1, &b));
  const Value* c = in.get(2);
  int64_t shape[2];
  RTG_TRY(gemm_op_out_shape(&a->t, &b->t, trans_a, trans_b, shape));
  Value y;
  RTG_TRY(alloc_value(ctx, RTG_F32, 2, shape, false, &y));
  rtg_status st = gemm_op_forward(ctx, a->t, b->t, c ? &c->t : nullptr, alpha, beta, trans_a, trans_b, y.t,
                                  b_pack.p && b_pack_from == b->t.data ? &b_pack : nullptr);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

rtg_status GemmOp::prepare(Context* ctx, const OpInputs& consts) {
  release(ctx);
  const Value* b = consts.get(1);
  if (!b || b->t.ndim != 2 || b->t.dtype != RTG_F32 || b->t.device == RTG_HOST) return RTG_OK;
  // the operand as the contraction sees it: [K, N] after the optional transposition (a view change)
  const int64_t K = trans_b ? b->t.shape[1] : b->t.shape[0], N = trans_b ? b->t.shape[0] : b->t.shape[1];
  const int64_t rs = trans_b ? b->t.strides[1] : b->t.strides[0], cs = trans_b ? b->t.strides[0] : b->t.strides[1];
  if (K == 0 || N == 0 || rs < 0 || cs < 0) return RTG_OK;
  void* d = nullptr;
  const int64_t ld = b_pack_ld(K);
  RTG_TRY(ctx->pool.alloc(static_cast<size_t>(N) * ld * 4 * 2, &d));
  b_pack = BPack{static_cast<const float*>(d), ld, N * ld};
  b_pack_from = b->t.data;
  return make_b_pack(ctx, MatView{static_cast<const float*>(b->t.data), K, N, rs, cs}, static_cast<float*>(d), N * ld);
}

void GemmOp::release(Context* ctx) {
  if (b_pack.p) ctx->pool.free(const_cast<float*>(b_pack.p));
  b_pack = BPack{};
  b_pack_from = nullptr;
}

rtg_status MatMulOp::prepare(Context* ctx, const OpInputs& consts) {
  release(ctx);
  const Value* b = consts.get(1);
  if (!b || has_perm[1] || b->t.ndim != 2 || b->t.dtype != RTG_F32 || b->t.device == RTG_HOST) return RTG_OK;
  const int64_t K = b->t.shape[0], N = b->t.shape[1];
  if (K == 0 || N == 0 || b->t.strides[0] < 0 || b->t.strides[1] < 0) return RTG_OK;
  void* d = nullptr;
  const int64_t ld = b_pack_ld(K);
  RTG_TRY(ctx->pool.alloc(static_cast<size_t>(N) * ld * 4 * 2, &d));
  b_pack = BPack{static_cast<const float*>(d), ld, N * ld};
  b_pack_from = b->t.data;
  b_pack_dims[0] = K, b_pack_dims[1] = N, b_pack_dims[2] = b->t.strides[0], b_pack_dims[3] = b->t.strides[1];
  return make_b_pack(ctx, MatView{static_cast<const float*>(b->t.data), K, N, b->t.strides[0], b->t.strides[1]},
                     static_cast<float*>(d), N * ld);
}

void MatMulOp::release(Context* ctx) {
  if (b_pack.p) ctx->pool.free(const_cast<float*>(b_pack.p));
  b_pack = BPack{};
  b_pack_from = nullptr;
}

rtg_status MatMulOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *a, *b;
  RTG_TRY(in.require(0, &a));
  RTG_TRY(in.require(1, &b));
  rtg_tensor ta = a->t, tb = b->t;
  for (int side = 0; side < 2; ++side) {
    if (!has_perm[side]) continue;
    rtg_tensor& t = side == 0 ? ta : tb;
    const rtg_tensor src = t;
    const int r = src.ndim;
    if (!perm_reverse[side] && static_cast<int>(perm[side].size()) != r)
      return set_error(RTG_ERR_INVALID_VALUE, "Permutation is invalid");
    unsigned seen = 0;
    for (int i = 0; i < r; ++i) {
      const int64_t ax = perm_reverse[side] ? r - 1 - i : perm[side][i];
      if (ax < 0 || ax >= r || (seen >> ax) & 1u) return set_error(RTG_ERR_INVALID_VALUE, "Permutation is invalid");
      seen |= 1u << ax;
      t.shape[i] = src.shape[ax], t.strides[i] = src.strides[ax];
    }
  }
  int64_t shape[RTG_MAX_DIMS];
  int32_t nd;
  RTG_TRY(matmul_out_shape(&ta, &tb, shape, &nd));
  Value y;
  RTG_TRY(alloc_value(ctx, RTG_F32, nd, shape, false, &y));
  Epilogue ep;
  const Epilogue* epp = nullptr;
  // fused `MatMul -> Add(constant vector)` (+ activation) is a per-column bias only when the vector has exactly N
  // entries; any other 1-D constant (e.g. shape [1]) is an ordinary broadcast Add, run after the product as the
  // reference does
  const Value* bias = has_bias ? in.get(2) : nullptr;
  if (has_bias && !bias) {
    release_value(ctx, &y);
    return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  }
  const bool bias_tail = bias && !(bias->t.dtype == RTG_F32 && numel(bias->t) == shape[nd - 1] && is_contiguous(bias->t));
  // fused residual: rides in the epilogue when it is a dense f32 tensor of the output's shape and the product is the
  // collapsed [.., M, K] x [K, N] form (one GEMM over all rows); otherwise it is added after the product, as unfused
  const Value* res = has_residual ? in.get(3) : nullptr;
  if (has_residual && !res) {
    release_value(ctx, &y);
    return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  }
  bool res_fused = res && !bias_tail && res->t.dtype == RTG_F32 && res->t.ndim == nd && is_contiguous(res->t) && tb.ndim == 2 &&
                   (reinterpret_cast<uintptr_t>(res->t.data) & 15) == 0;
  for (int i = 0; res_fused && i < nd; ++i) res_fused = res->t.shape[i] == shape[i];
  const bool res_tail = res && !res_fused;
  const bool act_fused = !bias_tail && !res_tail;  // the activation comes last: fused only if everything before it is
  if (bias && !bias_tail) {
    ep.bias = static_cast<const float*>(bias->t.data);
    ep.bias_mode = BIAS_COL;
  }
  if (res_fused) {
    ep.R = static_cast<const float*>(res->t.data);
    ep.beta = 1.f;
    ep.rmap = OutMap{shape[nd - 1], 0, 0, 0};
  }
  if ((bias && !bias_tail) || res_fused || (act_fused && act != ACT_NONE) || scale != 1.f) {
    ep.act = act_fused ? act : ACT_NONE;
    ep.alpha = scale;
    ep.alpha_div = scale_div;
    if (scale_div) {
      // x / 2^k and x * 2^-k are the same real number, hence the same correctly rounded float (BERT's sqrt(64) = 8):
      // the multiply stays on the epilogue's inline path, the IEEE division is an out-of-line loop
      uint32_t bits;
      memcpy(&bits, &scale, 4);
      const uint32_t expo = (bits >> 23) & 0xffu;
      if ((bits & 0x007fffffu) == 0 && expo >= 2 && expo <= 252) {
        ep.alpha = 1.0f / scale;  // exact
        ep.alpha_div = false;
      }
    }
    epp = &ep;
  }
  if (b_pack.p && tb.data == b_pack_from && tb.ndim == 2 && tb.shape[0] == b_pack_dims[0] &&
      tb.shape[1] == b_pack_dims[1] && tb.strides[0] == b_pack_dims[2] && tb.strides[1] == b_pack_dims[3]) {
    ep.b_pack = b_pack, ep.b_pack_const = true;
    epp = &ep;
  }
  rtg_status st = matmul_forward(ctx, ta, tb, y.t, epp);
  // unfused tails, in the order of the original chain: + bias, + residual, activation
  auto add_tail = [&](const Value* operand) {
    BinaryOp add(RTG_BIN_ADD, "Add");
    OpInputs ain;
    ain.v = {&y, operand};
    std::vector<Value> sum;
    rtg_status s2 = add.run(ctx, ain, &sum);
    if (s2 == RTG_OK) {
      release_value(ctx, &y);
      y = sum[0];
    }
    return s2;
  };
  if (st == RTG_OK && bias_tail) st = add_tail(bias);
  if (st == RTG_OK && res_tail) st = add_tail(res);
  if (st == RTG_OK && !act_fused && act != ACT_NONE) st = unary_forward(ctx, act, y.t, 0.f, 0.f, y.t);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

}  // namespace rtg
