// The remaining operators the target graphs need: Softmax, LayerNormalization, Transpose, Gather,
// BatchNormalization, Unsqueeze / Squeeze.  Row reductions use warp shuffles; everything is HBM-bound.
#include "operators.h"
#include "vecmath.cuh"

namespace rtg {

namespace {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// softmax over `len` elements spaced `stride` apart; one warp per lane (outer, inner).
// reference: rten-vecmath/src/softmax.rs (max-subtract, exp, normalise), src/ops/norm.rs softmax.
__global__ void softmax_kernel(float* __restrict__ out, const float* __restrict__ in, int64_t lanes, int len,
                               int64_t inner) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < lanes; r += nwarps) {
    const int64_t base = (r / inner) * len * inner + (r % inner);
    float m = __int_as_float(0xff800000);
    for (int i = lane; i < len; i += 32) m = fmaxf(m, in[base + i * inner]);
    m = warp_max(m);
    float s = 0.f;
    for (int i = lane; i < len; i += 32) {
      const float e = vm_exp(in[base + i * inner] - m);
      out[base + i * inner] = e;
      s += e;
    }
    s = warp_sum(s);
    const float inv = __fdiv_rn(1.0f, s);
    for (int i = lane; i < len; i += 32) out[base + i * inner] = out[base + i * inner] * inv;
  }
}

// One warp per row of `len` contiguous elements (src/ops/norm.rs:245-299).
__global__ void layer_norm_kernel(float* __restrict__ out, const float* __restrict__ in,
                                  const float* __restrict__ scale, const float* __restrict__ bias, int64_t rows,
                                  int len, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += nwarps) {
    const float* x = in + r * len;
    float* y = out + r * len;
    float s = 0.f;
    for (int i = lane; i < len; i += 32) s += x[i];
    const float mean = warp_sum(s) / static_cast<float>(len);
    float q = 0.f;
    for (int i = lane; i < len; i += 32) {
      const float d = x[i] - mean;
      q += d * d;
    }
    const float inv = 1.0f / sqrtf(warp_sum(q) / static_cast<float>(len) + eps);
    for (int i = lane; i < len; i += 32) {
      float v = (x[i] - mean) * inv * scale[i];
      if (bias) v += bias[i];
      y[i] = v;
    }
  }
}

// Row-in-registers variants for contiguous rows (the BERT shapes: 128-element score rows, 768-element hidden rows): one
// warp per row, the row is read once with 16-byte loads, kept in registers for the reductions and written once -- the
// kernels above re-read it two or three times with 4-byte loads and reached 1.1-1.3 TB/s.  MAXV = float4s per lane.
template <int MAXV, bool ADD>
__global__ void __launch_bounds__(256) softmax_rows_kernel(float* __restrict__ out, const float* __restrict__ in,
                                                            int64_t rows, int len, const float* __restrict__ addend,
                                                            int64_t rows_per_addend) {
  const int lane = threadIdx.x & 31;
  const int nv = len >> 2;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < rows; r += nwarps) {
    const float4* x = reinterpret_cast<const float4*>(in + r * len);
    float4* y = reinterpret_cast<float4*>(out + r * len);
    // fused `Add -> Softmax` (BERT's attention mask): row r adds row r / rows_per_addend of the addend, the same
    // round-to-nearest add the separate operator performs
    const float4* a = ADD ? reinterpret_cast<const float4*>(addend + (r / rows_per_addend) * len) : nullptr;
    float4 v[MAXV];
    float m = __int_as_float(0xff800000);
#pragma unroll
    for (int k = 0; k < MAXV; ++k) {
      const int i = lane + 32 * k;
      if (i < nv) {
        v[k] = x[i];
        if (ADD) {
          const float4 t = __ldg(a + i);
          v[k].x = __fadd_rn(v[k].x, t.x), v[k].y = __fadd_rn(v[k].y, t.y);
          v[k].z = __fadd_rn(v[k].z, t.z), v[k].w = __fadd_rn(v[k].w, t.w);
        }
        m = fmaxf(fmaxf(m, fmaxf(v[k].x, v[k].y)), fmaxf(v[k].z, v[k].w));
      }
    }
    m = warp_max(m);
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < MAXV; ++k) {
      if (lane + 32 * k < nv) {
        v[k].x = vm_exp(v[k].x - m), v[k].y = vm_exp(v[k].y - m), v[k].z = vm_exp(v[k].z - m), v[k].w = vm_exp(v[k].w - m);
        s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
      }
    }
    s = warp_sum(s);
    const float inv = __fdiv_rn(1.0f, s);
#pragma unroll
    for (int k = 0; k < MAXV; ++k) {
      const int i = lane + 32 * k;
      if (i < nv) y[i] = make_float4(v[k].x * inv, v[k].y * inv, v[k].z * inv, v[k].w * inv);
    }
  }
}

template <int MAXV>
__global__ void __launch_bounds__(256) layer_norm_rows_kernel(float* __restrict__ out, const float* __restrict__ in,
                                                               const float* __restrict__ scale,
                                                               const float* __restrict__ bias, int64_t rows, int len,
                                                               float eps) {
  const int lane = threadIdx.x & 31;
  const int nv = len >> 2;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const float4* sc = reinterpret_cast<const float4*>(scale);
  const float4* bi = reinterpret_cast<const float4*>(bias);
  for (int64_t r = warp; r < rows; r += nwarps) {
    const float4* x = reinterpret_cast<const float4*>(in + r * len);
    float4* y = reinterpret_cast<float4*>(out + r * len);
    float4 v[MAXV];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < MAXV; ++k) {
      const int i = lane + 32 * k;
      if (i < nv) {
        v[k] = x[i];
        s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
      }
    }
    const float mean = warp_sum(s) / static_cast<float>(len);
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < MAXV; ++k) {
      if (lane + 32 * k < nv) {
        v[k].x -= mean, v[k].y -= mean, v[k].z -= mean, v[k].w -= mean;
        q += (v[k].x * v[k].x + v[k].y * v[k].y) + (v[k].z * v[k].z + v[k].w * v[k].w);
      }
    }
    const float inv = 1.0f / sqrtf(warp_sum(q) / static_cast<float>(len) + eps);
#pragma unroll
    for (int k = 0; k < MAXV; ++k) {
      const int i = lane + 32 * k;
      if (i < nv) {
        const float4 g = __ldg(sc + i);
        float4 o = make_float4(v[k].x * inv * g.x, v[k].y * inv * g.y, v[k].z * inv * g.z, v[k].w * inv * g.w);
        if (bias) {
          const float4 b = __ldg(bi + i);
          o.x += b.x, o.y += b.y, o.z += b.z, o.w += b.w;
        }
        y[i] = o;
      }
    }
  }
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// out[i, :] = table[idx[i], :]   (Gather on axis 0, the embedding lookup; src/ops/gather.rs)
__global__ void gather_rows_kernel(uint32_t* __restrict__ out, const uint32_t* __restrict__ table,
                                   const int32_t* __restrict__ idx, int64_t n_idx, int64_t row, int64_t n_rows,
                                   int* __restrict__ err) {
  const int64_t total = n_idx * row;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t r = i / row, c = i % row;
    int64_t k = idx[r];
    if (k < 0) k += n_rows;
    if (k < 0 || k >= n_rows) {
      *err = 1;
      continue;
    }
    out[i] = table[k * row + c];
  }
}

// y = (x - mean[c]) * (scale[c] / sqrt(var[c] + eps)) + bias[c]  with channel stride handling (NCHW or NHWC)
__global__ void batch_norm_kernel(float* __restrict__ out, const float* __restrict__ in, const float* __restrict__ scale,
                                  const float* __restrict__ bias, const float* __restrict__ mean,
                                  const float* __restrict__ var, int64_t n, int c, int64_t inner, int chan_fastest,
                                  float eps) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int ch = chan_fastest ? static_cast<int>(i % c) : static_cast<int>((i / inner) % c);
    const float k = __fdiv_rn(scale[ch], __fsqrt_rn(__fadd_rn(var[ch], eps)));
    out[i] = __fadd_rn(__fmul_rn(__fsub_rn(in[i], mean[ch]), k), bias[ch]);
  }
}

int blocks_for(Context* ctx, int64_t items) {
  return static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(div_ceil(items, 256), static_cast<int64_t>(ctx->num_sms) * 16)));
}

rtg_status resolve_axis(int ndim, int64_t axis, int* out) {
  if (axis < -ndim || axis >= ndim) return set_error(RTG_ERR_INVALID_VALUE, "Axis is invalid");
  *out = static_cast<int>(axis < 0 ? axis + ndim : axis);
  return RTG_OK;
}

rtg_status make_contiguous(Context* ctx, const Value& v, Value* tmp, rtg_tensor* view) {
  if (is_contiguous(v.t)) {
    *view = v.t;
    return RTG_OK;
  }
  RTG_TRY(alloc_value(ctx, v.t.dtype, v.t.ndim, v.t.shape, false, tmp));
  RTG_TRY(copy_view(ctx, tmp->t, v.t));
  *view = tmp->t;
  return RTG_OK;
}

}  // namespace

// ---------------------------------------------------------------- Softmax
// `addend` (optional, fused Add -> Softmax): a contiguous f32 tensor of len-long rows, row r of x adds its row
// r / rows_per_addend.  Only the rows kernel takes it; *fused_ok tells the caller whether it was applied.
static rtg_status softmax_impl(Context* ctx, const rtg_tensor& x, int64_t axis_attr, const rtg_tensor& y,
                               const float* addend = nullptr, int64_t rows_per_addend = 1, bool* fused_ok = nullptr) {
  if (fused_ok) *fused_ok = false;
  if (x.ndim == 0) return set_error(RTG_ERR_INVALID_VALUE, "Axis is invalid");
  int axis;
  RTG_TRY(resolve_axis(x.ndim, axis_attr, &axis));
  const int64_t len = x.shape[axis];
  int64_t inner = 1, outer = 1;
  for (int i = axis + 1; i < x.ndim; ++i) inner *= x.shape[i];
  for (int i = 0; i < axis; ++i) outer *= x.shape[i];
  const int64_t lanes = outer * inner;
  if (lanes == 0 || len == 0) return RTG_OK;
  float* yp = static_cast<float*>(y.data);
  const float* xp = static_cast<const float*>(x.data);
  const int grid = blocks_for(ctx, lanes * 32);
  if (inner == 1 && len % 4 == 0 && len <= 2048 && aligned16(xp) && aligned16(yp) && (!addend || aligned16(addend))) {
    const int L = static_cast<int>(len);
    const float* ad = addend;
    const int64_t rpa = rows_per_addend;
#define RTG_SOFTMAX_ROWS(V)                                                                      \
  do {                                                                                           \
    if (ad) softmax_rows_kernel<V, true><<<grid, 256, 0, ctx->stream>>>(yp, xp, lanes, L, ad, rpa); \
    else softmax_rows_kernel<V, false><<<grid, 256, 0, ctx->stream>>>(yp, xp, lanes, L, ad, rpa);   \
  } while (0)
    if (len <= 128) RTG_SOFTMAX_ROWS(1);
    else if (len <= 256) RTG_SOFTMAX_ROWS(2);
    else if (len <= 512) RTG_SOFTMAX_ROWS(4);
    else if (len <= 1024) RTG_SOFTMAX_ROWS(8);
    else RTG_SOFTMAX_ROWS(16);
#undef RTG_SOFTMAX_ROWS
    if (fused_ok) *fused_ok = true;
  } else if (addend) {
    return RTG_OK;  // the caller adds first and comes back without the addend
  } else {
    softmax_kernel<<<grid, 256, 0, ctx->stream>>>(yp, xp, lanes, static_cast<int>(len), inner);
  }
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

// Fused `Add -> Softmax` (GraphOptimizer, model.cu): which rows of `a` row r of `x` adds.  True when x has the
// broadcast shape and `a` is contiguous with the shape [x's leading dims..., 1, ..., 1, len] (any number of leading dims,
// including none): then rows_per_addend = the product of x's dims that a's ones stand for.
static bool addend_rows(const rtg_tensor& x, const rtg_tensor& a, int axis, int64_t* rows_per_addend) {
  if (a.dtype != RTG_F32 || x.dtype != RTG_F32 || axis != x.ndim - 1 || a.ndim > x.ndim || !is_contiguous(a)) return false;
  const int off = x.ndim - a.ndim;
  if (a.ndim == 0 || a.shape[a.ndim - 1] != x.shape[x.ndim - 1]) return false;
  int k = 0;  // leading dims of x that `a` matches (dims x has and `a` lacks count as ones of `a`)
  if (off == 0)
    while (k < x.ndim - 1 && a.shape[k] == x.shape[k]) ++k;
  int64_t rpa = 1;
  for (int i = k; i < x.ndim - 1; ++i) {
    if (i >= off && a.shape[i - off] != 1) return false;
    rpa *= x.shape[i];
  }
  *rows_per_addend = rpa > 0 ? rpa : 1;
  return true;
}

rtg_status SoftmaxOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  if (x->t.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  const Value* addend = has_addend ? in.get(1) : nullptr;
  if (has_addend && !addend) return set_error(RTG_ERR_MISSING_INPUTS, "Missing inputs");
  Value sum;  // the unfused Add, when the operands do not fit the fused kernel
  if (addend) {
    int ax = 0;
    int64_t rpa = 1;
    bool fits = resolve_axis(x->t.ndim, axis, &ax) == RTG_OK && is_contiguous(x->t);
    const Value* big = x;
    if (fits && !addend_rows(x->t, addend->t, ax, &rpa)) {
      // Add is commutative: the mask may be the first operand
      fits = resolve_axis(addend->t.ndim, axis, &ax) == RTG_OK && is_contiguous(addend->t) &&
             addend_rows(addend->t, x->t, ax, &rpa);
      if (fits) big = addend, addend = x;
    }
    if (fits) {
      Value y;
      RTG_TRY(alloc_value(ctx, RTG_F32, big->t.ndim, big->t.shape, false, &y));
      bool fused_ok = false;
      rtg_status st = softmax_impl(ctx, big->t, axis, y.t, static_cast<const float*>(addend->t.data), rpa, &fused_ok);
      if (st == RTG_OK && fused_ok) {
        out->push_back(y);
        return RTG_OK;
      }
      release_value(ctx, &y);
      if (st != RTG_OK) return st;
    }
    // general broadcast, other axes, unaligned rows: the two operators one after the other
    const Value* other = in.get(1);
    int64_t shape[RTG_MAX_DIMS];
    int32_t nd;
    RTG_TRY(binary_out_shape(&x->t, &other->t, shape, &nd));
    RTG_TRY(alloc_value(ctx, RTG_F32, nd, shape, false, &sum));
    rtg_status st = binary_forward(ctx, RTG_BIN_ADD, x->t, other->t, sum.t);
    if (st != RTG_OK) {
      release_value(ctx, &sum);
      return st;
    }
    x = &sum;
  }
  Value tmp, y;
  rtg_tensor xv;
  rtg_status st = make_contiguous(ctx, *x, &tmp, &xv);
  if (st == RTG_OK) st = alloc_value(ctx, RTG_F32, xv.ndim, xv.shape, false, &y);
  if (st == RTG_OK) st = softmax_impl(ctx, xv, axis, y.t);
  release_value(ctx, &tmp);
  release_value(ctx, &sum);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

rtg_status SoftmaxOp::run_in_place(Context* ctx, Value* io, const OpInputs& rest) {
  if (io->t.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  const Value* addend = has_addend ? rest.get(0) : nullptr;
  if (has_addend && !addend) return set_error(RTG_ERR_MISSING_INPUTS, "Missing inputs");
  int ax = 0;
  int64_t rpa = 1;
  const bool fused_fits = addend && is_contiguous(io->t) && resolve_axis(io->t.ndim, axis, &ax) == RTG_OK &&
                          addend_rows(io->t, addend->t, ax, &rpa);
  if (!is_contiguous(io->t) || (addend && !fused_fits)) {
    OpInputs in;
    in.v = {io};
    if (addend) in.v.push_back(addend);
    std::vector<Value> outs;
    RTG_TRY(run(ctx, in, &outs));
    release_value(ctx, io);
    *io = outs[0];
    return RTG_OK;
  }
  io->host_f32.reset();
  if (!addend) return softmax_impl(ctx, io->t, axis, io->t);
  bool fused_ok = false;
  RTG_TRY(softmax_impl(ctx, io->t, axis, io->t, static_cast<const float*>(addend->t.data), rpa, &fused_ok));
  if (fused_ok) return RTG_OK;
  // rows the fused kernel does not take (length not a multiple of four, ...): add in place, then the plain form
  RTG_TRY(binary_forward(ctx, RTG_BIN_ADD, io->t, addend->t, io->t));
  return softmax_impl(ctx, io->t, axis, io->t);
}

// ---------------------------------------------------------------- LayerNormalization
rtg_status LayerNormOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *x, *scale;
  RTG_TRY(in.require(0, &x));
  RTG_TRY(in.require(1, &scale));
  const Value* bias = in.get(2);
  int ax;
  RTG_TRY(resolve_axis(x->t.ndim, axis, &ax));
  int64_t len = 1, rows = 1;
  for (int i = ax; i < x->t.ndim; ++i) len *= x->t.shape[i];
  for (int i = 0; i < ax; ++i) rows *= x->t.shape[i];
  // scale / bias must cover exactly the normalised dims (the common case; general broadcast is not implemented)
  if (numel(scale->t) != len) return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "`scale` cannot be broadcast to input shape");
  if (bias && numel(bias->t) != len) return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "`bias` cannot be broadcast to input shape");
  Value tmp, y;
  rtg_tensor xv;
  RTG_TRY(make_contiguous(ctx, *x, &tmp, &xv));
  rtg_status st = alloc_value(ctx, RTG_F32, xv.ndim, xv.shape, false, &y);
  if (st == RTG_OK && rows * len > 0) {
    float* yp = static_cast<float*>(y.t.data);
    const float* xp = static_cast<const float*>(xv.data);
    const float* sp = static_cast<const float*>(scale->t.data);
    const float* bp = bias ? static_cast<const float*>(bias->t.data) : nullptr;
    const int grid = blocks_for(ctx, rows * 32);
    const int L = static_cast<int>(len);
    if (len % 4 == 0 && len <= 1024 && aligned16(xp) && aligned16(yp) && aligned16(sp) && aligned16(bp) &&
        is_contiguous(scale->t) && (!bias || is_contiguous(bias->t))) {
      if (len <= 256) layer_norm_rows_kernel<2><<<grid, 256, 0, ctx->stream>>>(yp, xp, sp, bp, rows, L, epsilon);
      else if (len <= 512) layer_norm_rows_kernel<4><<<grid, 256, 0, ctx->stream>>>(yp, xp, sp, bp, rows, L, epsilon);
      else layer_norm_rows_kernel<8><<<grid, 256, 0, ctx->stream>>>(yp, xp, sp, bp, rows, L, epsilon);
    } else {
      layer_norm_kernel<<<grid, 256, 0, ctx->stream>>>(yp, xp, sp, bp, rows, L, epsilon);
    }
    ctx->launches++;
    if (cudaGetLastError() != cudaSuccess) st = set_error(RTG_ERR_CUDA, "layer_norm launch failed");
  }
  release_value(ctx, &tmp);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- Transpose
rtg_status TransposeOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  const int nd = x->t.ndim;
  std::vector<int64_t> p = perm;
  if (!has_perm) {
    p.clear();
    for (int i = nd - 1; i >= 0; --i) p.push_back(i);  // default: reverse the dims
  }
  if (static_cast<int>(p.size()) != nd) return set_error(RTG_ERR_INVALID_VALUE, "Permutation is invalid");
  std::vector<char> seen(nd, 0);
  for (int64_t a : p) {
    if (a < 0 || a >= nd || seen[a]) return set_error(RTG_ERR_INVALID_VALUE, "Permutation is invalid");
    seen[a] = 1;
  }
  rtg_tensor view = x->t;  // permuted view of the input
  for (int i = 0; i < nd; ++i) view.shape[i] = x->t.shape[p[i]], view.strides[i] = x->t.strides[p[i]];
  Value y;
  RTG_TRY(alloc_value(ctx, x->t.dtype, nd, view.shape, false, &y));
  rtg_status st = copy_view(ctx, y.t, view);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- Gather (axis 0 fast path: embedding lookup)
// ---------------------------------------------------------------- Cast
namespace {
__global__ void cast_i32_f32_kernel(float* __restrict__ y, const int32_t* __restrict__ x, int64_t n) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x)
    y[i] = static_cast<float>(x[i]);  // round to nearest even, like Rust's `as f32`
}
__global__ void cast_f32_i32_kernel(int32_t* __restrict__ y, const float* __restrict__ x, int64_t n) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float v = x[i];
    // Rust `as i32`: truncate toward zero, saturate, NaN -> 0
    y[i] = v != v ? 0 : (v >= 2147483648.f ? INT32_MAX : (v <= -2147483648.f ? INT32_MIN : static_cast<int32_t>(v)));
  }
}
}  // namespace

rtg_status CastOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  Value xt;
  rtg_tensor xv;
  RTG_TRY(make_contiguous(ctx, *x, &xt, &xv));
  Value y;
  rtg_status st = alloc_value(ctx, to, xv.ndim, xv.shape, false, &y);
  if (st == RTG_OK) {
    const int64_t n = numel(xv);
    if (n > 0) {
      if (xv.dtype == to) {
        cudaMemcpyAsync(y.t.data, xv.data, static_cast<size_t>(n) * 4, cudaMemcpyDeviceToDevice, ctx->stream);
      } else if (to == RTG_F32) {
        cast_i32_f32_kernel<<<blocks_for(ctx, n), 256, 0, ctx->stream>>>(static_cast<float*>(y.t.data),
                                                                         static_cast<const int32_t*>(xv.data), n);
        ctx->launches++;
      } else {
        cast_f32_i32_kernel<<<blocks_for(ctx, n), 256, 0, ctx->stream>>>(static_cast<int32_t*>(y.t.data),
                                                                         static_cast<const float*>(xv.data), n);
        ctx->launches++;
      }
    }
    if (to == RTG_I32 && x->host_f32) {
      auto h = std::make_shared<std::vector<int32_t>>();
      for (float v : *x->host_f32) h->push_back(static_cast<int32_t>(v));
      y.host_i32 = h;
    }
    if (to == RTG_F32 && x->host_i32) {
      auto h = std::make_shared<std::vector<float>>();
      for (int32_t v : *x->host_i32) h->push_back(static_cast<float>(v));
      y.host_f32 = h;
    }
  }
  release_value(ctx, &xt);
  if (st != RTG_OK) return st;
  out->push_back(y);
  return RTG_OK;
}

rtg_status GatherOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value *x, *idx;
  RTG_TRY(in.require(0, &x));
  RTG_TRY(in.require(1, &idx));
  if (idx->t.dtype != RTG_I32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  int ax;
  RTG_TRY(resolve_axis(x->t.ndim, axis, &ax));
  if (ax != 0) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "Gather is implemented for axis 0 only on the GPU path");
  Value xt, it;
  rtg_tensor xv, iv;
  RTG_TRY(make_contiguous(ctx, *x, &xt, &xv));
  rtg_status st = make_contiguous(ctx, *idx, &it, &iv);
  int64_t shape[RTG_MAX_DIMS];
  int nd = 0;
  for (int i = 0; i < iv.ndim; ++i) shape[nd++] = iv.shape[i];
  for (int i = 1; i < xv.ndim; ++i) {
    if (nd >= RTG_MAX_DIMS) st = set_error(RTG_ERR_UNSUPPORTED_VALUE, "too many dimensions");
    else shape[nd++] = xv.shape[i];
  }
  Value y;
  if (st == RTG_OK) st = alloc_value(ctx, xv.dtype, nd, shape, false, &y);
  if (st == RTG_OK) {
    const int64_t n_idx = numel(iv), row = xv.shape[0] ? numel(xv) / xv.shape[0] : 0;
    if (n_idx * row > 0) {
      Scratch flag;
      st = flag.alloc(ctx, 4);
      if (st == RTG_OK) {
        cudaMemsetAsync(flag.p, 0, 4, ctx->stream);
        gather_rows_kernel<<<blocks_for(ctx, n_idx * row), 256, 0, ctx->stream>>>(
            static_cast<uint32_t*>(y.t.data), static_cast<const uint32_t*>(xv.data),
            static_cast<const int32_t*>(iv.data), n_idx, row, xv.shape[0], static_cast<int*>(flag.p));
        ctx->launches++;
        int bad = 0;
        cudaMemcpyAsync(&bad, flag.p, 4, cudaMemcpyDeviceToHost, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        if (bad) st = set_error(RTG_ERR_INVALID_VALUE, "Entry in `indices` is out of range");
      }
    }
  }
  release_value(ctx, &xt);
  release_value(ctx, &it);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}

// ---------------------------------------------------------------- BatchNormalization
static rtg_status batch_norm_impl(Context* ctx, float eps, const rtg_tensor& x, const OpInputs& params, size_t first,
                                  const rtg_tensor& y) {
  const Value* p[4];
  for (int i = 0; i < 4; ++i) {
    RTG_TRY(params.require(first + i, &p[i]));
    if (p[i]->t.ndim != 1) return set_error(RTG_ERR_INVALID_VALUE, "scale must have 1 dims");
    if (!is_contiguous(p[i]->t)) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "batch norm parameters must be dense");
  }
  if (x.ndim < 3) return set_error(RTG_ERR_INVALID_VALUE, "Input must have at least 3 dims");
  const int c = static_cast<int>(x.shape[1]);
  for (int i = 0; i < 4; ++i)
    if (p[i]->t.shape[0] != c) return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "batch norm parameter length must match channels");
  const int64_t n = numel(x);
  if (n == 0) return RTG_OK;
  const bool nhwc = x.ndim == 4 && is_nhwc_dense(x) && !is_contiguous(x);
  if (!nhwc && !is_contiguous(x)) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "batch norm input must be dense");
  int64_t inner = 1;
  for (int i = 2; i < x.ndim; ++i) inner *= x.shape[i];
  batch_norm_kernel<<<blocks_for(ctx, n), 256, 0, ctx->stream>>>(
      static_cast<float*>(y.data), static_cast<const float*>(x.data), static_cast<const float*>(p[0]->t.data),
      static_cast<const float*>(p[1]->t.data), static_cast<const float*>(p[2]->t.data),
      static_cast<const float*>(p[3]->t.data), n, c, inner, nhwc ? 1 : 0, eps);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

rtg_status BatchNormOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  Value y;
  const bool nhwc = x->t.ndim == 4 && is_nhwc_dense(x->t) && !is_contiguous(x->t);
  RTG_TRY(alloc_value(ctx, RTG_F32, x->t.ndim, x->t.shape, nhwc, &y));
  rtg_status st = batch_norm_impl(ctx, epsilon, x->t, in, 1, y.t);
  if (st != RTG_OK) {
    release_value(ctx, &y);
    return st;
  }
  out->push_back(y);
  return RTG_OK;
}
rtg_status BatchNormOp::run_in_place(Context* ctx, Value* io, const OpInputs& rest) {
  return batch_norm_impl(ctx, epsilon, io->t, rest, 0, io->t);
}

// ---------------------------------------------------------------- Unsqueeze / Squeeze
rtg_status ShapeOpsOp::new_shape(Context* ctx, const rtg_tensor& x, const Value* axes_v, int64_t* shape, int* nd) const {
  std::vector<int64_t> axes;
  if (axes_v) RTG_TRY(host_ints(ctx, *axes_v, &axes));
  if (squeeze) {
    std::vector<char> drop(x.ndim, 0);
    if (!axes_v) {
      for (int i = 0; i < x.ndim; ++i) drop[i] = x.shape[i] == 1;
    } else {
      for (int64_t a : axes) {
        int r;
        RTG_TRY(resolve_axis(x.ndim, a, &r));
        if (x.shape[r] != 1) return set_error(RTG_ERR_INVALID_VALUE, "Can only remove dimensions of size 1");
        drop[r] = 1;
      }
    }
    *nd = 0;
    for (int i = 0; i < x.ndim; ++i)
      if (!drop[i]) shape[(*nd)++] = x.shape[i];
    return RTG_OK;
  }
  if (!axes_v) return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  const int out_nd = x.ndim + static_cast<int>(axes.size());
  if (out_nd > RTG_MAX_DIMS) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "too many dimensions");
  std::vector<char> is_new(out_nd, 0);
  for (int64_t a : axes) {
    int r;
    RTG_TRY(resolve_axis(out_nd, a, &r));
    if (is_new[r]) return set_error(RTG_ERR_INVALID_VALUE, "Axes must be unique");
    is_new[r] = 1;
  }
  int src = 0;
  for (int i = 0; i < out_nd; ++i) shape[i] = is_new[i] ? 1 : x.shape[src++];
  *nd = out_nd;
  return RTG_OK;
}

rtg_status ShapeOpsOp::run(Context* ctx, const OpInputs& in, std::vector<Value>* out) {
  const Value* x;
  RTG_TRY(in.require(0, &x));
  int64_t shape[RTG_MAX_DIMS];
  int nd;
  RTG_TRY(new_shape(ctx, x->t, in.get(1), shape, &nd));
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

rtg_status ShapeOpsOp::run_in_place(Context* ctx, Value* io, const OpInputs& rest) {
  int64_t shape[RTG_MAX_DIMS];
  int nd;
  RTG_TRY(new_shape(ctx, io->t, rest.get(0), shape, &nd));
  if (!is_contiguous(io->t)) {
    OpInputs in;
    in.v = {io, rest.get(0)};
    std::vector<Value> outs;
    RTG_TRY(run(ctx, in, &outs));
    release_value(ctx, io);
    *io = outs[0];
    return RTG_OK;
  }
  io->t.ndim = nd;
  for (int i = 0; i < nd; ++i) io->t.shape[i] = shape[i];
  set_contiguous_strides(io->t);
  return RTG_OK;
}

}  // namespace rtg
