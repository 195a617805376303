// Shape-arithmetic operators that BERT-class exports keep in the graph: Concat, Slice, Shape, Expand.
// All of them are index-driven copies (bit-exact): the data moves with copy_view's strided gather kernels, the index
// arithmetic runs on the host from the host mirrors of the small int tensors.
#include "operators.h"

namespace rtg {

namespace {
rtg_status resolve_axis(int ndim, int64_t axis, int* out) {  // src/ops/mod.rs resolve_axis
  if (axis < -ndim || axis >= ndim) return set_error(RTG_ERR_INVALID_VALUE, "Axis is invalid");
  *out = static_cast<int>(axis < 0 ? axis + ndim : axis);
  return RTG_OK;
}
}  // namespace

// ---------------------------------------------------------------- Concat  (src/ops/concat.rs:14-98)
rtg_status ConcatOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* first;
  RTG_TRY(in.require(0, &first));
  int ax;
  RTG_TRY(resolve_axis(first->t.ndim, axis, &ax));
  int64_t shape[RTG_MAX_DIMS];
  for (int i = 0; i < first->t.ndim; ++i) shape[i] = first->t.shape[i];
  shape[ax] = 0;
  for (size_t k = 0; k < in.v.size(); ++k) {
    const Value* x;
    RTG_TRY(in.require(k, &x));
    if (x->t.dtype != first->t.dtype) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
    if (x->t.ndim != first->t.ndim)
      return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Tensors must have the same number of dimensions");
    for (int i = 0; i < x->t.ndim; ++i)
      if (i != ax && x->t.shape[i] != first->t.shape[i])
        return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Dimensions must be the same except for concat axis");
    shape[ax] += x->t.shape[ax];
  }
  Value y;
  RTG_TRY(alloc_value(ctx, first->t.dtype, first->t.ndim, shape, false, &y));
  int64_t off = 0;
  std::shared_ptr<std::vector<int32_t>> host;  // concatenated shape vectors stay readable on the host
  bool all_host = first->t.dtype == RTG_I32 && first->t.ndim == 1;
  if (all_host) host = std::make_shared<std::vector<int32_t>>();
  for (size_t k = 0; k < in.v.size(); ++k) {
    const Value* x = in.v[k];
    if (all_host && x->host_i32) host->insert(host->end(), x->host_i32->begin(), x->host_i32->end());
    else all_host = false;
    if (numel(x->t) == 0) continue;
    rtg_tensor dst = y.t;  // the window of the output this input fills
    dst.shape[ax] = x->t.shape[ax];
    dst.data = static_cast<char*>(y.t.data) + off * y.t.strides[ax] * 4;
    rtg_status st = copy_view(ctx, dst, x->t);
    if (st != RTG_OK) {
      release_value(ctx, &y);
      return st;
    }
    off += x->t.shape[ax];
  }
  if (all_host) y.host_i32 = host;
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- Slice  (src/ops/slice.rs:14-66; SliceRange::clamp /
// resolve, rten-tensor/src/slice_range.rs:230-332: NumPy semantics, endpoints clamped, negative steps walk backwards)
rtg_status SliceOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *x, *starts_v, *ends_v;
  RTG_TRY(in.require(0, &x));
  RTG_TRY(in.require(1, &starts_v));
  RTG_TRY(in.require(2, &ends_v));
  std::vector<int64_t> starts, ends, axes, steps;
  if (starts_v->t.ndim != 1 || ends_v->t.ndim != 1) return set_error(RTG_ERR_INVALID_VALUE, "starts must have 1 dims");
  RTG_TRY(host_ints(ctx, *starts_v, &starts));
  RTG_TRY(host_ints(ctx, *ends_v, &ends));
  const Value* axes_v = in.get(3);
  const Value* steps_v = in.get(4);
  if (axes_v) RTG_TRY(host_ints(ctx, *axes_v, &axes));
  if (steps_v) RTG_TRY(host_ints(ctx, *steps_v, &steps));
  for (int64_t s : steps)
    if (s == 0) return set_error(RTG_ERR_INVALID_VALUE, "steps must be non-zero");
  rtg_tensor view = x->t;
  char* base = static_cast<char*>(x->t.data);
  const size_t n = std::min(starts.size(), ends.size());
  for (size_t i = 0; i < n; ++i) {
    int ax = static_cast<int>(i);
    if (axes_v) {
      if (i >= axes.size()) break;
      RTG_TRY(resolve_axis(x->t.ndim, axes[i], &ax));
    } else if (ax >= x->t.ndim) {
      break;
    }
    const int64_t len = x->t.shape[ax], step = i < steps.size() ? steps[i] : 1;
    auto clamp_index = [&](int64_t v) {
      if (v < 0) {
        v += len;
        if (v < 0) v = step < 0 ? -1 : 0;
      } else if (v >= len) {
        v = step < 0 ? len - 1 : len;
      }
      return v;
    };
    const int64_t s = clamp_index(starts[i]), e = clamp_index(ends[i]);
    int64_t count = step > 0 ? (e - s + step - 1) / step : (s - e + (-step) - 1) / (-step);
    if (count < 0) count = 0;
    if (count > 0) base += s * x->t.strides[ax] * 4;
    view.shape[ax] = count;
    view.strides[ax] = x->t.strides[ax] * step;
  }
  view.data = base;
  Value y;
  RTG_TRY(alloc_value(ctx, x->t.dtype, view.ndim, view.shape, false, &y));
  if (numel(view) > 0) {
    rtg_status st = copy_view(ctx, y.t, view);
    if (st != RTG_OK) {
      release_value(ctx, &y);
      return st;
    }
  }
  if (x->host_i32 && x->t.ndim == 1 && x->t.dtype == RTG_I32) {  // a slice of a shape vector stays host-readable
    auto h = std::make_shared<std::vector<int32_t>>();
    const int64_t st = view.strides[0] / (x->t.strides[0] ? x->t.strides[0] : 1);
    const int64_t first = (base - static_cast<char*>(x->t.data)) / 4 / (x->t.strides[0] ? x->t.strides[0] : 1);
    for (int64_t k = 0; k < view.shape[0]; ++k) h->push_back((*x->host_i32)[static_cast<size_t>(first + k * st)]);
    y.host_i32 = h;
  }
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- Shape  (src/ops/layout.rs:356-377)
rtg_status ShapeOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  auto h = std::make_shared<std::vector<int32_t>>();
  for (int i = 0; i < x->t.ndim; ++i) h->push_back(static_cast<int32_t>(x->t.shape[i]));
  const int64_t n = x->t.ndim;
  Value y;
  RTG_TRY(alloc_value(ctx, RTG_I32, 1, &n, false, &y));
  if (n > 0) RTG_CUDA(cudaMemcpyAsync(y.t.data, h->data(), static_cast<size_t>(n) * 4, cudaMemcpyHostToDevice, ctx->stream));
  y.host_i32 = h;  // keeps the source of the asynchronous copy alive and serves host_ints without a round trip
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- Expand  (src/ops/layout.rs:17-110)
rtg_status ExpandOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *x, *shape_v;
  RTG_TRY(in.require(0, &x));
  RTG_TRY(in.require(1, &shape_v));
  if (shape_v->t.ndim != 1) return set_error(RTG_ERR_INVALID_VALUE, "shape must have 1 dims");
  std::vector<int64_t> want;
  RTG_TRY(host_ints(ctx, *shape_v, &want));
  // broadcast_shapes(input_shape, shape): align at the trailing dimension, sizes must match or one of them be 1
  const int nd = static_cast<int>(std::max<size_t>(want.size(), static_cast<size_t>(x->t.ndim)));
  if (nd > RTG_MAX_DIMS) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "too many dimensions");
  rtg_tensor view = x->t;
  view.ndim = nd;
  for (int i = 0; i < nd; ++i) {
    const int ix = i - (nd - x->t.ndim), iw = i - (nd - static_cast<int>(want.size()));
    const int64_t a = ix >= 0 ? x->t.shape[ix] : 1, b = iw >= 0 ? want[static_cast<size_t>(iw)] : 1;
    if (a != b && a != 1 && b != 1)
      return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Cannot broadcast input with target shape");
    view.shape[i] = a == 1 ? b : a;
    view.strides[i] = (ix >= 0 && a != 1) ? x->t.strides[ix] : 0;
  }
  Value y;
  RTG_TRY(alloc_value(ctx, x->t.dtype, nd, view.shape, false, &y));
  if (numel(view) > 0) {
    rtg_status st = copy_view(ctx, y.t, view);
    if (st != RTG_OK) {
      release_value(ctx, &y);
      return st;
    }
  }
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- Where  (src/ops/binary_elementwise.rs:850-915)
namespace {
struct Strided3 {
  int nd;
  int64_t shape[RTG_MAX_DIMS], sc[RTG_MAX_DIMS], sx[RTG_MAX_DIMS], sy[RTG_MAX_DIMS];
};
__global__ void where_kernel(uint32_t* __restrict__ out, const int32_t* __restrict__ cond, const uint32_t* __restrict__ x,
                             const uint32_t* __restrict__ y, Strided3 g, int64_t total) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int64_t r = i, oc = 0, ox = 0, oy = 0;
    for (int d = g.nd - 1; d >= 0; --d) {
      const int64_t k = r % g.shape[d];
      r /= g.shape[d];
      oc += k * g.sc[d], ox += k * g.sx[d], oy += k * g.sy[d];
    }
    out[i] = cond[oc] != 0 ? x[ox] : y[oy];  // 4-byte elements of either type, copied bit for bit
  }
}

// trailing-aligned broadcast of `shape` (ndim `nd`) against `t`; false if incompatible
bool broadcast_into(int64_t* shape, int* nd, const rtg_tensor& t) {
  const int n = std::max(*nd, t.ndim);
  int64_t out[RTG_MAX_DIMS];
  for (int i = 0; i < n; ++i) {
    const int ia = i - (n - *nd), ib = i - (n - t.ndim);
    const int64_t a = ia >= 0 ? shape[ia] : 1, b = ib >= 0 ? t.shape[ib] : 1;
    if (a != b && a != 1 && b != 1) return false;
    out[i] = a == 1 ? b : a;
  }
  for (int i = 0; i < n; ++i) shape[i] = out[i];
  *nd = n;
  return true;
}
void broadcast_strides(const rtg_tensor& t, int nd, int64_t* st) {
  for (int i = 0; i < nd; ++i) {
    const int it = i - (nd - t.ndim);
    st[i] = (it >= 0 && t.shape[it] != 1) ? t.strides[it] : 0;
  }
}
}  // namespace

rtg_status WhereOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *c, *x, *y;
  RTG_TRY(in.require(0, &c));
  RTG_TRY(in.require(1, &x));
  RTG_TRY(in.require(2, &y));
  if (c->t.dtype != RTG_I32 || x->t.dtype != y->t.dtype) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  Strided3 g;
  g.nd = x->t.ndim;
  for (int i = 0; i < g.nd; ++i) g.shape[i] = x->t.shape[i];
  if (!broadcast_into(g.shape, &g.nd, y->t) || !broadcast_into(g.shape, &g.nd, c->t))
    return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Cannot broadcast inputs");
  broadcast_strides(c->t, g.nd, g.sc);
  broadcast_strides(x->t, g.nd, g.sx);
  broadcast_strides(y->t, g.nd, g.sy);
  Value r;
  RTG_TRY(alloc_value(ctx, x->t.dtype, g.nd, g.shape, false, &r));
  const int64_t total = numel(r.t);
  if (total > 0) {
    const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(total, 256), static_cast<int64_t>(ctx->num_sms) * 16));
    where_kernel<<<blocks, 256, 0, ctx->stream>>>(static_cast<uint32_t*>(r.t.data), static_cast<const int32_t*>(c->t.data),
                                                  static_cast<const uint32_t*>(x->t.data),
                                                  static_cast<const uint32_t*>(y->t.data), g, total);
    ctx->launches++;
    if (cudaGetLastError() != cudaSuccess) {
      release_value(ctx, &r);
      return set_error(RTG_ERR_CUDA, "where_kernel launch failed");
    }
  }
  out->push_back(r);
  return RTG_OK;
}

// ---------------------------------------------------------------- ReduceMean  (src/ops/reduce.rs:228-353)
namespace {
// one warp per row of a dense [rows, len] matrix: lanes stride over the row, shuffle tree, sum / len
__global__ void row_mean_kernel(float* __restrict__ out, const float* __restrict__ x, int64_t rows, int64_t len) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += nwarps) {
    const float* p = x + r * len;
    float s = 0.f;
    for (int64_t i = lane; i < len; i += 32) s += p[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[r] = __fdiv_rn(s, static_cast<float>(len));
  }
}
}  // namespace

rtg_status ReduceMeanOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  if (x->t.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  const int nd = x->t.ndim;
  bool red[RTG_MAX_DIMS] = {};
  if (!has_axes || axes.empty()) {
    for (int i = 0; i < nd; ++i) red[i] = true;
  } else {
    for (int64_t a : axes) {
      int ax;
      RTG_TRY(resolve_axis(nd, a, &ax));
      red[ax] = true;
    }
  }
  // kept dims first, reduced dims last: a dense [rows, len] matrix (gathered unless the input already has that form)
  rtg_tensor view = x->t;
  int64_t oshape[RTG_MAX_DIMS];
  int k = 0, ond = 0;
  int64_t rows = 1, len = 1;
  for (int i = 0; i < nd; ++i)
    if (!red[i]) view.shape[k] = x->t.shape[i], view.strides[k] = x->t.strides[i], rows *= x->t.shape[i], ++k;
  for (int i = 0; i < nd; ++i)
    if (red[i]) view.shape[k] = x->t.shape[i], view.strides[k] = x->t.strides[i], len *= x->t.shape[i], ++k;
  for (int i = 0; i < nd; ++i) {
    if (!red[i]) oshape[ond++] = x->t.shape[i];
    else if (keep_dims) oshape[ond++] = 1;
  }
  if (len == 0) return set_error(RTG_ERR_INVALID_VALUE, "Cannot reduce empty tensor");
  Value y;
  RTG_TRY(alloc_value(ctx, RTG_F32, ond, oshape, false, &y));
  if (rows > 0) {
    Scratch tmp;
    const float* src = static_cast<const float*>(x->t.data);
    if (!is_contiguous(view)) {
      rtg_status st = tmp.alloc(ctx, static_cast<size_t>(rows) * len * 4);
      rtg_tensor dst = view;
      dst.data = tmp.p;
      set_contiguous_strides(dst);
      if (st == RTG_OK) st = copy_view(ctx, dst, view);
      if (st != RTG_OK) {
        release_value(ctx, &y);
        return st;
      }
      src = tmp.f32();
    }
    const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(rows * 32, 256), static_cast<int64_t>(ctx->num_sms) * 16));
    row_mean_kernel<<<blocks, 256, 0, ctx->stream>>>(static_cast<float*>(y.t.data), src, rows, len);
    ctx->launches++;
  }
  out->push_back(y);
  return RTG_OK;
}

}  // namespace rtg
