// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
// Restates src/ops/matmul.rs (gemm_op :27-81, matmul_impl :123-239).
#include <algorithm>

#include "ops.h"

namespace oracle {

static size_t prod(const size_t* s, size_t n) {
  size_t p = 1;
  for (size_t i = 0; i < n; i++) p *= s[i];
  return p;
}

// to_contiguous(): row-major copy of a strided view.
static std::vector<float> to_contiguous(const TensorView& t) {
  size_t total = prod(t.shape, t.ndim);
  std::vector<float> out(total);
  for (size_t idx = 0; idx < total; idx++) {
    size_t rem = idx;
    ptrdiff_t off = 0;
    for (size_t d = t.ndim; d-- > 0;) {
      off += (ptrdiff_t)(rem % t.shape[d]) * t.strides[d];
      rem /= t.shape[d];
    }
    out[idx] = t.data[off];
  }
  return out;
}

OpError matmul(const Kernel* gk, const TensorView& a, const TensorView& b, std::vector<float>* out,
               std::vector<size_t>* out_shape) {
  if (a.ndim < 2 || b.ndim < 2) return {OP_INVALID_VALUE, "Inputs must have >= 2 dimensions"};
  const size_t a_rows = a.shape[a.ndim - 2], a_cols = a.shape[a.ndim - 1];
  const size_t b_rows = b.shape[b.ndim - 2], b_cols = b.shape[b.ndim - 1];
  if (a_cols != b_rows)
    return {OP_INCOMPATIBLE_INPUT_SHAPES, "Columns of first matrix does not match rows of second matrix"};
  const size_t a_pn = a.ndim - 2, b_pn = b.ndim - 2;
  const size_t num_a = prod(a.shape, a_pn), num_b = prod(b.shape, b_pn);

  // broadcast_shapes(a_prefix, b_prefix)
  size_t pn = std::max(a_pn, b_pn);
  std::vector<size_t> out_prefix(pn);
  for (size_t i = 0; i < pn; i++) {
    size_t da = i + a_pn >= pn ? a.shape[i + a_pn - pn] : 1;
    size_t db = i + b_pn >= pn ? b.shape[i + b_pn - pn] : 1;
    if (da == db) out_prefix[i] = da;
    else if (da == 1) out_prefix[i] = db;
    else if (db == 1) out_prefix[i] = da;
    else return {OP_INCOMPATIBLE_INPUT_SHAPES, "Cannot broadcast shapes"};
  }
  *out_shape = out_prefix;
  out_shape->push_back(a_rows);
  out_shape->push_back(b_cols);

  // [A..., M, K] x [K, N]  ->  one [A*M, K] x [K, N] GEMM   (matmul.rs:162-169)
  if (num_a > 1 && num_b == 1) {
    std::vector<float> a_contig = to_contiguous(a);
    TensorView a2{a_contig.data(), 2, {num_a * a_rows, a_cols}, {(ptrdiff_t)a_cols, 1}};
    std::vector<size_t> shape2;
    return matmul(gk, a2, b, out, &shape2);
  }

  size_t total = prod(out_shape->data(), out_shape->size());
  out->assign(total, 0.f);
  if (total == 0) return {};

  const size_t n_out = prod(out_prefix.data(), pn);
  const size_t out_row_stride = b_cols;

  // Matrix index -> data offset of a broadcast operand.
  auto batch_offset = [&](const TensorView& t, size_t t_pn, size_t idx) {
    ptrdiff_t off = 0;
    size_t rem = idx;
    for (size_t d = pn; d-- > 0;) {
      size_t coord = rem % out_prefix[d];
      rem /= out_prefix[d];
      if (d + t_pn >= pn) {
        size_t td = d + t_pn - pn;
        if (t.shape[td] != 1) off += (ptrdiff_t)coord * t.strides[td];
      }
    }
    return off;
  };
  auto mat_at = [&](const TensorView& t, ptrdiff_t off) {
    return Matrix{t.data + off, t.shape[t.ndim - 2], t.shape[t.ndim - 1], t.strides[t.ndim - 2],
                  t.strides[t.ndim - 1]};
  };

  PackedAMatrix packed_a;
  PackedBMatrix packed_b;
  const bool use_pa = num_a == 1 && num_b > 1 && a_rows > 1;
  const bool use_pb = num_a > 1 && num_b == 1 && a_rows > 1;
  if (use_pa) packed_a = prepack_a(gk, mat_at(a, 0));
  if (use_pb) packed_b = prepack_b(gk, mat_at(b, 0));

  ThreadPool::instance().parallel_for(n_out, [&](size_t i) {
    Matrix am = mat_at(a, batch_offset(a, a_pn, i)), bm = mat_at(b, batch_offset(b, b_pn, i));
    GemmInputA ia;
    if (use_pa) ia.packed = &packed_a;
    else ia.unpacked = &am;
    GemmInputB ib;
    if (use_pb) ib.packed = &packed_b;
    else ib.unpacked = &bm;
    gemm_impl(gk, out->data() + i * a_rows * out_row_stride, a_rows * out_row_stride, out_row_stride, ia, ib, 1.f, 0.f,
              nullptr, 0);
  });
  return {};
}

OpError gemm_op(const Kernel* gk, const TensorView& a0, const TensorView& b0, const TensorView* c, float alpha,
                float beta, bool transpose_a, bool transpose_b, std::vector<float>* out, size_t out_shape[2]) {
  if (a0.ndim != 2 || b0.ndim != 2) return {OP_INVALID_VALUE, "a must have 2 dims"};
  Matrix a{a0.data, a0.shape[0], a0.shape[1], a0.strides[0], a0.strides[1]};
  Matrix b{b0.data, b0.shape[0], b0.shape[1], b0.strides[0], b0.strides[1]};
  if (transpose_a) {
    std::swap(a.rows, a.cols);
    std::swap(a.row_stride, a.col_stride);
  }
  if (transpose_b) {
    std::swap(b.rows, b.cols);
    std::swap(b.row_stride, b.col_stride);
  }
  out_shape[0] = a.rows;
  out_shape[1] = b.cols;
  const size_t M = a.rows, N = b.cols;
  GemmInputA ia;
  ia.unpacked = &a;
  GemmInputB ib;
  ib.unpacked = &b;
  if (c && beta != 0.f) {
    // can_broadcast_to(out_shape) + expand_to
    if (c->ndim > 2) return {OP_INCOMPATIBLE_INPUT_SHAPES, "Cannot broadcast c to output shape"};
    size_t cs[2] = {1, 1};
    ptrdiff_t cst[2] = {0, 0};
    for (size_t i = 0; i < c->ndim; i++) {
      cs[2 - c->ndim + i] = c->shape[i];
      cst[2 - c->ndim + i] = c->strides[i];
    }
    if ((cs[0] != M && cs[0] != 1) || (cs[1] != N && cs[1] != 1))
      return {OP_INCOMPATIBLE_INPUT_SHAPES, "Cannot broadcast c to output shape"};
    out->assign(M * N, 0.f);
    for (size_t i = 0; i < M; i++)
      for (size_t j = 0; j < N; j++)
        (*out)[i * N + j] = c->data[(cs[0] == 1 ? 0 : (ptrdiff_t)i * cst[0]) + (cs[1] == 1 ? 0 : (ptrdiff_t)j * cst[1])];
    if (gemm_impl(gk, out->data(), M * N, N, ia, ib, alpha, beta, nullptr, 0) != 0)
      return {OP_PANIC, "Columns of matrix `a` must match rows of matrix `b`"};
  } else {
    out->assign(M * N, 0.f);
    if (gemm_impl(gk, out->data(), M * N, N, ia, ib, alpha, 0.f, nullptr, 0) != 0)
      return {OP_PANIC, "Columns of matrix `a` must match rows of matrix `b`"};
  }
  return {};
}

}  // namespace oracle
