// Elementwise operators and strided copies: HBM-bound, vectorised (float4), grid sized to the SM count.
//   unary   Relu / Clip / Gelu / Sigmoid / Silu / Erf   reference: src/ops/unary_elementwise.rs:22-63,331-339,423-445,571
//   binary  Add / Sub / Mul / Div with broadcasting       reference: src/ops/binary_elementwise.rs:65-157,261-322
//   copy    view -> view                                  reference: TensorBase::to_contiguous / copy_from
// Algorithmic bytes: unary 8 B/element (4 in place... read + write), binary 12 B/element, copy 8 B/element.
#include "ops.h"
#include "vecmath.cuh"

namespace rtg {

namespace {

struct Strides8 {
  int32_t ndim;
  int64_t shape[RTG_MAX_DIMS];
  int64_t sa[RTG_MAX_DIMS];
  int64_t sb[RTG_MAX_DIMS];
  int64_t so[RTG_MAX_DIMS];
};

__device__ __forceinline__ float bin_op(int op, float a, float b) {
  switch (op) {
    case RTG_BIN_ADD: return __fadd_rn(a, b);
    case RTG_BIN_SUB: return __fsub_rn(a, b);
    case RTG_BIN_MUL: return __fmul_rn(a, b);
    case RTG_BIN_DIV: return __fdiv_rn(a, b);
    default:  // Pow, with the reference's fast paths (src/ops/binary_elementwise.rs:743-751)
      if (b == 2.f) return __fmul_rn(a, a);
      if (b == 3.f) return __fmul_rn(__fmul_rn(a, a), a);
      return powf(a, b);
  }
}

template <bool VEC>
__global__ void unary_kernel(float* __restrict__ out, const float* __restrict__ in, int64_t n, int act, float lo,
                             float hi) {
  const int64_t step = static_cast<int64_t>(gridDim.x) * blockDim.x;
  int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (VEC) {
    const int64_t nv = n >> 2;
    const float4* in4 = reinterpret_cast<const float4*>(in);
    float4* out4 = reinterpret_cast<float4*>(out);
    for (; i < nv; i += step) {
      float4 v = in4[i];
      v.x = apply_act(v.x, act, lo, hi);
      v.y = apply_act(v.y, act, lo, hi);
      v.z = apply_act(v.z, act, lo, hi);
      v.w = apply_act(v.w, act, lo, hi);
      out4[i] = v;
    }
    // tail
    for (int64_t j = (nv << 2) + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; j < n; j += step)
      out[j] = apply_act(in[j], act, lo, hi);
  } else {
    for (; i < n; i += step) out[i] = apply_act(in[i], act, lo, hi);
  }
}

// out[i] = a[i] op b[(i / repeats) % cycle]  -- the reference's fast broadcast form ("cycles x repeats",
// src/ops/binary_elementwise.rs:261-322); repeats == 1 && cycle == n is the same-shape case.
template <bool VEC>
__global__ void binary_cyclic_kernel(float* __restrict__ out, const float* __restrict__ a, const float* __restrict__ b,
                                     int64_t n, int64_t cycle, int64_t repeats, int op, bool swap) {
  const int64_t step = static_cast<int64_t>(gridDim.x) * blockDim.x;
  int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (VEC) {  // same shape, 16-byte aligned
    const int64_t nv = n >> 2;
    const float4* a4 = reinterpret_cast<const float4*>(a);
    const float4* b4 = reinterpret_cast<const float4*>(b);
    float4* o4 = reinterpret_cast<float4*>(out);
    for (; i < nv; i += step) {
      const float4 x = a4[i], y = b4[i];
      o4[i] = make_float4(bin_op(op, x.x, y.x), bin_op(op, x.y, y.y), bin_op(op, x.z, y.z), bin_op(op, x.w, y.w));
    }
    for (int64_t j = (nv << 2) + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; j < n; j += step)
      out[j] = bin_op(op, a[j], b[j]);
  } else {
    for (; i < n; i += step) {
      const float x = a[i], y = b[(i / repeats) % cycle];
      out[i] = swap ? bin_op(op, y, x) : bin_op(op, x, y);
    }
  }
}

__global__ void binary_strided_kernel(float* __restrict__ out, const float* __restrict__ a, const float* __restrict__ b,
                                      int64_t n, Strides8 s, int op) {
  const int64_t step = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += step) {
    int64_t r = i, oa = 0, ob = 0, oo = 0;
    for (int d = s.ndim - 1; d >= 0; --d) {
      const int64_t idx = r % s.shape[d];
      r /= s.shape[d];
      oa += idx * s.sa[d];
      ob += idx * s.sb[d];
      oo += idx * s.so[d];
    }
    out[oo] = bin_op(op, a[oa], b[ob]);
  }
}

// Broadcast binary operator on a contiguous output seen as [rows][L]: `rows` enumerates up to four outer dimensions
// (after merging nested ones), the inner dimension is dense in the output and either dense or broadcast (stride 0) in
// each operand.  A thread owns one float4 column of the row and walks rows with an odometer, so there is no division
// per element; grid.y splits long inner dimensions (a scalar operand collapses the whole tensor into one row).
struct RowGeom {
  uint32_t size[4];  // outer dimensions, outermost first, right-aligned (unused leading slots: huge size, stride 0)
  int64_t sa[4], sb[4];
  int64_t L, rows;
  int32_t a_in, b_in;  // inner strides: 1 or 0
};

template <bool VEC>
__global__ void binary_rows_kernel(float* __restrict__ out, const float* __restrict__ a, const float* __restrict__ b,
                                   RowGeom g, int op, int64_t rows_per_block, int64_t seg) {
  const int64_t r_begin = static_cast<int64_t>(blockIdx.x) * rows_per_block;
  const int64_t r_end = min(g.rows, r_begin + rows_per_block);
  const int64_t c_begin = static_cast<int64_t>(blockIdx.y) * seg, c_end = min(g.L, c_begin + seg);
  int64_t r = r_begin + threadIdx.y;
  if (r >= r_end) return;
  uint32_t idx[4] = {0, 0, 0, 0};
  {
    int64_t t = r;
#pragma unroll
    for (int d = 3; d >= 0; --d) {
      idx[d] = static_cast<uint32_t>(t % g.size[d]);
      t /= g.size[d];
    }
  }
  for (; r < r_end; r += blockDim.y) {
    int64_t oa = 0, ob = 0;
#pragma unroll
    for (int d = 0; d < 4; ++d) oa += idx[d] * g.sa[d], ob += idx[d] * g.sb[d];
    const float* pa = a + oa;
    const float* pb = b + ob;
    float* po = out + r * g.L;
    if (VEC) {
      for (int64_t c = c_begin + 4 * threadIdx.x; c < c_end; c += 4 * blockDim.x) {
        float4 x, y;
        if (g.a_in) x = *reinterpret_cast<const float4*>(pa + c);
        else x.x = x.y = x.z = x.w = __ldg(pa);
        if (g.b_in) y = *reinterpret_cast<const float4*>(pb + c);
        else y.x = y.y = y.z = y.w = __ldg(pb);
        *reinterpret_cast<float4*>(po + c) =
            make_float4(bin_op(op, x.x, y.x), bin_op(op, x.y, y.y), bin_op(op, x.z, y.z), bin_op(op, x.w, y.w));
      }
    } else {
      for (int64_t c = c_begin + threadIdx.x; c < c_end; c += blockDim.x)
        po[c] = bin_op(op, pa[g.a_in ? c : 0], pb[g.b_in ? c : 0]);
    }
    // advance the odometer by blockDim.y rows
    idx[3] += blockDim.y;
#pragma unroll
    for (int d = 3; d > 0; --d) {
      while (idx[d] >= g.size[d]) idx[d] -= g.size[d], idx[d - 1]++;
    }
  }
}

// Permuting copy that keeps the innermost dimension: dense output [rows][L], source rows located by the odometer.
template <bool VEC>
__global__ void copy_rows_kernel(uint32_t* __restrict__ out, const uint32_t* __restrict__ in, RowGeom g,
                                 int64_t rows_per_block, int64_t seg) {
  const int64_t r_begin = static_cast<int64_t>(blockIdx.x) * rows_per_block;
  const int64_t r_end = min(g.rows, r_begin + rows_per_block);
  const int64_t c_begin = static_cast<int64_t>(blockIdx.y) * seg, c_end = min(g.L, c_begin + seg);
  int64_t r = r_begin + threadIdx.y;
  if (r >= r_end) return;
  uint32_t idx[4];
  {
    int64_t t = r;
#pragma unroll
    for (int d = 3; d >= 0; --d) {
      idx[d] = static_cast<uint32_t>(t % g.size[d]);
      t /= g.size[d];
    }
  }
  for (; r < r_end; r += blockDim.y) {
    int64_t oa = 0;
#pragma unroll
    for (int d = 0; d < 4; ++d) oa += idx[d] * g.sa[d];
    const uint32_t* pi = in + oa;
    uint32_t* po = out + r * g.L;
    if (VEC) {
      for (int64_t c = c_begin + 4 * threadIdx.x; c < c_end; c += 4 * blockDim.x)
        *reinterpret_cast<uint4*>(po + c) = *reinterpret_cast<const uint4*>(pi + c);
    } else {
      for (int64_t c = c_begin + threadIdx.x; c < c_end; c += blockDim.x) po[c] = pi[c];
    }
    idx[3] += blockDim.y;
#pragma unroll
    for (int d = 3; d > 0; --d) {
      while (idx[d] >= g.size[d]) idx[d] -= g.size[d], idx[d - 1]++;
    }
  }
}

__global__ void copy_strided_nd_kernel(uint32_t* __restrict__ out, const uint32_t* __restrict__ in, int64_t n,
                                       Strides8 s) {
  const int64_t step = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += step) {
    int64_t r = i, oa = 0, oo = 0;
    for (int d = s.ndim - 1; d >= 0; --d) {
      const int64_t idx = r % s.shape[d];
      r /= s.shape[d];
      oa += idx * s.sa[d];
      oo += idx * s.so[d];
    }
    out[oo] = in[oa];
  }
}

int grid_for(Context* ctx, int64_t work_items) {
  return static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(div_ceil(work_items, 256), static_cast<int64_t>(ctx->num_sms) * 16)));
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Do two views of the same shape address memory in the same order (identical strides on non-unit dims) and densely?
bool same_dense_order(const rtg_tensor& a, const rtg_tensor& b) {
  if (a.ndim != b.ndim) return false;
  for (int i = 0; i < a.ndim; ++i) {
    if (a.shape[i] != b.shape[i]) return false;
    if (a.shape[i] != 1 && a.strides[i] != b.strides[i]) return false;
  }
  return span_of(a) == numel(a);  // a permutation of a dense block
}

// Block / grid shape shared by the row kernels: tx threads along the inner dimension, ty rows per pass.
struct RowLaunch {
  dim3 grid, block;
  int64_t rpb, seg;
  bool ok;
};
RowLaunch row_launch(Context* ctx, const RowGeom& g, bool vec) {
  RowLaunch l = {};
  const int64_t per_thread = vec ? 4 : 1;
  int tx = 8;
  while (tx < 256 && tx * per_thread < g.L) tx *= 2;
  const int ty = 256 / tx;
  const int64_t span = static_cast<int64_t>(tx) * per_thread;
  l.seg = std::min<int64_t>(g.L, span * 4);
  if (div_ceil(g.L, l.seg) > 65535) l.seg = div_ceil(div_ceil(g.L, 65535), span) * span;
  const int64_t gy = div_ceil(g.L, l.seg);
  const int64_t want_x = std::max<int64_t>(1, static_cast<int64_t>(ctx->num_sms) * 16 / gy);
  l.rpb = std::max<int64_t>(ty, div_ceil(div_ceil(g.rows, want_x), ty) * ty);
  const int64_t gx = div_ceil(g.rows, l.rpb);
  l.ok = gy <= 65535 && gx <= 0x7fffffff;
  l.grid = dim3(static_cast<unsigned>(gx), static_cast<unsigned>(gy));
  l.block = dim3(tx, ty);
  return l;
}

}  // namespace

rtg_status copy_view(Context* ctx, const rtg_tensor& dst, const rtg_tensor& src) {
  const int64_t n = numel(dst);
  if (n == 0) return RTG_OK;
  if (same_dense_order(dst, src)) {
    RTG_CUDA(cudaMemcpyAsync(dst.data, src.data, static_cast<size_t>(n) * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    return RTG_OK;
  }
  // 2-D transposes go through the tiled kernel
  if (dst.ndim == 2 && is_contiguous(dst))
    return copy_strided_2d(ctx, static_cast<float*>(dst.data), dst.shape[1], static_cast<const float*>(src.data),
                           dst.shape[0], dst.shape[1], src.strides[0], src.strides[1]);
  // permutations that keep the innermost dimension: rows x inner, vectorised
  if (is_contiguous(dst)) {
    int64_t shp[RTG_MAX_DIMS], sa[RTG_MAX_DIMS];
    int nd = 0;
    for (int i = 0; i < dst.ndim; ++i) {
      if (dst.shape[i] == 1) continue;
      if (nd > 0 && sa[nd - 1] == dst.shape[i] * src.strides[i]) shp[nd - 1] *= dst.shape[i], sa[nd - 1] = src.strides[i];
      else shp[nd] = dst.shape[i], sa[nd] = src.strides[i], ++nd;
    }
    bool ok = nd >= 2 && nd <= 5 && sa[nd - 1] == 1 && shp[nd - 1] >= 4;
    for (int i = 0; ok && i + 1 < nd; ++i) ok = shp[i] <= 0x7fffffff && sa[i] >= 0;
    if (ok) {
      RowGeom g = {};
      g.rows = 1;
      const int lead = 4 - (nd - 1);
      for (int i = 0; i < lead; ++i) g.size[i] = 0x7fffffffu;
      for (int i = 0; i + 1 < nd; ++i) g.size[lead + i] = static_cast<uint32_t>(shp[i]), g.sa[lead + i] = sa[i], g.rows *= shp[i];
      g.L = shp[nd - 1];
      bool vec = g.L % 4 == 0 && aligned16(dst.data) && aligned16(src.data);
      for (int i = 0; vec && i + 1 < nd; ++i) vec = sa[i] % 4 == 0;
      const RowLaunch l = row_launch(ctx, g, vec);
      if (l.ok) {
        uint32_t* po = static_cast<uint32_t*>(dst.data);
        const uint32_t* pi = static_cast<const uint32_t*>(src.data);
        if (vec) copy_rows_kernel<true><<<l.grid, l.block, 0, ctx->stream>>>(po, pi, g, l.rpb, l.seg);
        else copy_rows_kernel<false><<<l.grid, l.block, 0, ctx->stream>>>(po, pi, g, l.rpb, l.seg);
        ctx->launches++;
        RTG_CUDA(cudaGetLastError());
        return RTG_OK;
      }
    }
  }
  Strides8 s = {};
  s.ndim = dst.ndim;
  for (int i = 0; i < dst.ndim; ++i) s.shape[i] = dst.shape[i], s.sa[i] = src.strides[i], s.so[i] = dst.strides[i];
  copy_strided_nd_kernel<<<grid_for(ctx, n), 256, 0, ctx->stream>>>(static_cast<uint32_t*>(dst.data),
                                                                   static_cast<const uint32_t*>(src.data), n, s);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

rtg_status unary_forward(Context* ctx, int act, const rtg_tensor& in, float lo, float hi, const rtg_tensor& out) {
  if (in.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  const int64_t n = numel(in);
  if (n == 0) return RTG_OK;
  if (!same_dense_order(out, in)) {
    // gather to the output's order first, then run in place
    RTG_TRY(copy_view(ctx, out, in));
    if (span_of(out) != n) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "unary op output must be dense");
    rtg_tensor o2 = out;
    return unary_forward(ctx, act, o2, lo, hi, out);
  }
  const bool vec = aligned16(in.data) && aligned16(out.data);
  if (vec)
    unary_kernel<true><<<grid_for(ctx, n / 4 + 1), 256, 0, ctx->stream>>>(static_cast<float*>(out.data),
                                                                         static_cast<const float*>(in.data), n, act, lo, hi);
  else
    unary_kernel<false><<<grid_for(ctx, n), 256, 0, ctx->stream>>>(static_cast<float*>(out.data),
                                                                  static_cast<const float*>(in.data), n, act, lo, hi);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

rtg_status binary_out_shape(const rtg_tensor* a, const rtg_tensor* b, int64_t* shape, int32_t* ndim) {
  if (!a || !b) return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  const int nd = std::max(a->ndim, b->ndim);
  for (int i = 0; i < nd; ++i) {
    const int ia = i - (nd - a->ndim), ib = i - (nd - b->ndim);
    const int64_t da = ia >= 0 ? a->shape[ia] : 1, db = ib >= 0 ? b->shape[ib] : 1;
    if (da == db) shape[i] = da;
    else if (da == 1) shape[i] = db;
    else if (db == 1) shape[i] = da;
    else return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Cannot broadcast inputs to same shape");
  }
  *ndim = nd;
  return RTG_OK;
}

rtg_status binary_forward(Context* ctx, int op, const rtg_tensor& a, const rtg_tensor& b, const rtg_tensor& out) {
  if (a.dtype != RTG_F32 || b.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  const int64_t n = numel(out);
  if (n == 0) return RTG_OK;
  float* po = static_cast<float*>(out.data);
  const float* pa = static_cast<const float*>(a.data);
  const float* pb = static_cast<const float*>(b.data);
  // same shape, same memory order: one flat vectorised pass
  if (same_dense_order(out, a) && same_dense_order(out, b)) {
    const bool vec = aligned16(pa) && aligned16(pb) && aligned16(po);
    if (vec)
      binary_cyclic_kernel<true><<<grid_for(ctx, n / 4 + 1), 256, 0, ctx->stream>>>(po, pa, pb, n, n, 1, op, false);
    else
      binary_cyclic_kernel<false><<<grid_for(ctx, n), 256, 0, ctx->stream>>>(po, pa, pb, n, n, 1, op, false);
    ctx->launches++;
    RTG_CUDA(cudaGetLastError());
    return RTG_OK;
  }
  // broadcast onto a dense output: merge nested dimensions, then rows x inner with an odometer (no per-element division)
  if (is_contiguous(out)) {
    int64_t shp[RTG_MAX_DIMS], sa[RTG_MAX_DIMS], sb[RTG_MAX_DIMS];
    int nd = 0;
    for (int i = 0; i < out.ndim; ++i) {
      if (out.shape[i] == 1) continue;
      const int ia = i - (out.ndim - a.ndim), ib = i - (out.ndim - b.ndim);
      const int64_t ta = (ia >= 0 && a.shape[ia] != 1) ? a.strides[ia] : 0;
      const int64_t tb = (ib >= 0 && b.shape[ib] != 1) ? b.strides[ib] : 0;
      if (nd > 0 && sa[nd - 1] == out.shape[i] * ta && sb[nd - 1] == out.shape[i] * tb) {
        shp[nd - 1] *= out.shape[i], sa[nd - 1] = ta, sb[nd - 1] = tb;
      } else {
        shp[nd] = out.shape[i], sa[nd] = ta, sb[nd] = tb, ++nd;
      }
    }
    bool ok = nd >= 1 && nd <= 5 && (sa[nd - 1] == 0 || sa[nd - 1] == 1) && (sb[nd - 1] == 0 || sb[nd - 1] == 1);
    for (int i = 0; ok && i + 1 < nd; ++i) ok = shp[i] <= 0x7fffffff;
    if (ok) {
      RowGeom g = {};
      g.rows = 1;
      const int lead = 4 - (nd - 1);
      for (int i = 0; i < lead; ++i) g.size[i] = 0x7fffffffu;
      for (int i = 0; i + 1 < nd; ++i)
        g.size[lead + i] = static_cast<uint32_t>(shp[i]), g.sa[lead + i] = sa[i], g.sb[lead + i] = sb[i], g.rows *= shp[i];
      g.L = shp[nd - 1];
      g.a_in = static_cast<int32_t>(sa[nd - 1]), g.b_in = static_cast<int32_t>(sb[nd - 1]);
      bool vec = g.L % 4 == 0 && aligned16(po) && (!g.a_in || aligned16(pa)) && (!g.b_in || aligned16(pb));
      for (int i = 0; vec && i + 1 < nd; ++i) vec = (!g.a_in || sa[i] % 4 == 0) && (!g.b_in || sb[i] % 4 == 0);
      const RowLaunch l = row_launch(ctx, g, vec);
      if (l.ok) {
        if (vec) binary_rows_kernel<true><<<l.grid, l.block, 0, ctx->stream>>>(po, pa, pb, g, op, l.rpb, l.seg);
        else binary_rows_kernel<false><<<l.grid, l.block, 0, ctx->stream>>>(po, pa, pb, g, op, l.rpb, l.seg);
        ctx->launches++;
        RTG_CUDA(cudaGetLastError());
        return RTG_OK;
      }
    }
  }
  // general broadcast: per-element index arithmetic with zero strides on broadcast dims
  Strides8 s = {};
  s.ndim = out.ndim;
  for (int i = 0; i < out.ndim; ++i) {
    const int ia = i - (out.ndim - a.ndim), ib = i - (out.ndim - b.ndim);
    s.shape[i] = out.shape[i];
    s.so[i] = out.strides[i];
    s.sa[i] = (ia >= 0 && a.shape[ia] != 1) ? a.strides[ia] : 0;
    s.sb[i] = (ib >= 0 && b.shape[ib] != 1) ? b.strides[ib] : 0;
  }
  binary_strided_kernel<<<grid_for(ctx, n), 256, 0, ctx->stream>>>(po, pa, pb, n, s, op);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

}  // namespace rtg
