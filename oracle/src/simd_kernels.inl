// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
//
// Generic micro-kernels, included once per ISA translation unit with a vector
// policy `V` (scalar / AVX2 / AVX-512).  Everything is in an anonymous namespace so
// each TU keeps its own ISA-specific copy.
//
// Restates src/gemm/kernels.rs:26-316 (simd_gemv, simd_gemv_transposed,
// simd_gemv_fallback, simd_gemm).  Operation ORDER follows the reference exactly:
// sequential fused multiply-adds over k, alpha/beta store variants by case.
#include <cmath>

#include "gemm.h"

namespace {

using oracle::Matrix;

// src/gemm/kernels.rs:177-194
inline void simd_gemv_fallback(float* out, size_t out_len, const float* a, size_t a_len, Matrix b, float alpha,
                               float beta) {
  for (size_t col = 0; col < out_len; col++) {
    float acc = 0.f;
    for (size_t k = 0; k < a_len; k++) acc = std::fmaf(a[k], b.at(k, col), acc);
    acc *= alpha;
    if (beta == 0.f) out[col] = acc;
    else out[col] = acc + beta * out[col];
  }
}

// src/gemm/kernels.rs:107-169
template <class V>
inline void simd_gemv_transposed(float* out, size_t out_len, const float* a, size_t a_len, Matrix b, float alpha,
                                 float beta) {
  const float* b_ptr = b.data;
  const ptrdiff_t b_col_stride = b.col_stride;
  constexpr size_t COL_TILE = 8;
  size_t n_full = out_len / COL_TILE * COL_TILE;
  for (size_t col0 = 0; col0 < n_full; col0 += COL_TILE) {
    typename V::T vacc[COL_TILE];
    for (size_t i = 0; i < COL_TILE; i++) vacc[i] = V::zero();
    size_t k_full = a_len / V::LEN * V::LEN;
    for (size_t k0 = 0; k0 < k_full; k0 += V::LEN) {
      typename V::T a_tile = V::load(a + k0);
      for (size_t i = 0; i < COL_TILE; i++) {
        const float* b_col = b_ptr + (ptrdiff_t)(col0 + i) * b_col_stride;
        vacc[i] = V::mul_add(a_tile, V::load(b_col + k0), vacc[i]);
      }
    }
    float acc[COL_TILE];
    for (size_t i = 0; i < COL_TILE; i++) acc[i] = V::sum(vacc[i]);
    for (size_t k = k_full; k < a_len; k++) {
      float ak = a[k];
      for (size_t i = 0; i < COL_TILE; i++) {
        const float* b_col = b_ptr + (ptrdiff_t)(col0 + i) * b_col_stride;
        acc[i] = std::fmaf(ak, b_col[k], acc[i]);
      }
    }
    if (beta == 0.f) {
      for (size_t i = 0; i < COL_TILE; i++) out[col0 + i] = alpha * acc[i];
    } else {
      for (size_t i = 0; i < COL_TILE; i++) out[col0 + i] = alpha * acc[i] + beta * out[col0 + i];
    }
  }
  if (n_full < out_len) {
    Matrix tail = b;
    tail.data = b.data + (ptrdiff_t)n_full * b.col_stride;
    tail.cols = out_len - n_full;
    simd_gemv_fallback(out + n_full, out_len - n_full, a, a_len, tail, alpha, beta);
  }
}

// src/gemm/kernels.rs:26-103
template <class V, size_t NR_REGS>
inline void simd_gemv(float* out, size_t out_len, const float* a, size_t a_len, Matrix b, float alpha, float beta) {
  if (b.row_stride == 1) return simd_gemv_transposed<V>(out, out_len, a, a_len, b, alpha, beta);
  if (b.col_stride != 1) return simd_gemv_fallback(out, out_len, a, a_len, b, alpha, beta);

  const float* b_ptr = b.data;
  const ptrdiff_t brs = b.row_stride;
  constexpr size_t TILE = NR_REGS * V::LEN;
  size_t n_full = out_len / TILE * TILE;
  for (size_t c0 = 0; c0 < n_full; c0 += TILE) {
    typename V::T acc[NR_REGS];
    for (size_t i = 0; i < NR_REGS; i++) acc[i] = V::zero();
    for (size_t k = 0; k < a_len; k++) {
      typename V::T a_elts = V::splat(a[k]);
      for (size_t i = 0; i < NR_REGS; i++) {
        typename V::T b_elts = V::load(b_ptr + (ptrdiff_t)k * brs + c0 + i * V::LEN);
        acc[i] = V::mul_add(a_elts, b_elts, acc[i]);
      }
    }
    if (alpha != 1.f) {
      typename V::T av = V::splat(alpha);
      for (size_t i = 0; i < NR_REGS; i++) acc[i] = V::mul(acc[i], av);
    }
    if (beta == 0.f) {
      for (size_t i = 0; i < NR_REGS; i++) V::store(out + c0 + i * V::LEN, acc[i]);
    } else if (beta == 1.f) {
      for (size_t i = 0; i < NR_REGS; i++) {
        float* p = out + c0 + i * V::LEN;
        V::store(p, V::add(V::load(p), acc[i]));
      }
    } else {
      typename V::T bv = V::splat(beta);
      for (size_t i = 0; i < NR_REGS; i++) {
        float* p = out + c0 + i * V::LEN;
        V::store(p, V::mul_add(V::load(p), bv, acc[i]));
      }
    }
  }
  for (size_t c = n_full; c < out_len; c++) {
    float acc = 0.f;
    // Plain `acc += a*b` in the reference (kernels.rs:97): mul then add, not fused.
    for (size_t k = 0; k < a_len; k++) {
      volatile float prod = a[k] * b_ptr[(ptrdiff_t)k * brs + c];
      acc += prod;
    }
    float tmp = beta == 0.f ? 0.f : out[c];
    volatile float t1 = beta * tmp;
    volatile float t2 = acc * alpha;
    out[c] = t1 + t2;
  }
}

// src/gemm/kernels.rs:206-316
template <class V, size_t MR, size_t NR_REGS>
inline void simd_gemm(float* tile_ptr, size_t tile_row_stride, const float* a, const float* b, size_t depth,
                      float alpha, float beta) {
  typename V::T tmp[MR][NR_REGS];
  typename V::T b_rows[NR_REGS];
  for (size_t i = 0; i < MR; i++)
    for (size_t j = 0; j < NR_REGS; j++) tmp[i][j] = V::zero();

  for (size_t k = 0; k < depth; k++) {
    const float* a_k = a + k * MR;
    const float* b_k = b + k * NR_REGS * V::LEN;
#pragma GCC unroll 8
    for (size_t j = 0; j < NR_REGS; j++) b_rows[j] = V::load(b_k + j * V::LEN);
#pragma GCC unroll 8
    for (size_t i = 0; i < MR; i++) {
      typename V::T a_b = V::splat(a_k[i]);
#pragma GCC unroll 8
      for (size_t j = 0; j < NR_REGS; j++) tmp[i][j] = V::mul_add(a_b, b_rows[j], tmp[i][j]);
    }
  }

  auto out_ptr = [&](size_t i, size_t j) { return tile_ptr + tile_row_stride * i + j * V::LEN; };
  if (beta == 0.f && alpha == 1.f) {
    for (size_t i = 0; i < MR; i++)
      for (size_t j = 0; j < NR_REGS; j++) V::store(out_ptr(i, j), tmp[i][j]);
  } else if (beta == 1.f && alpha == 1.f) {
    for (size_t i = 0; i < MR; i++)
      for (size_t j = 0; j < NR_REGS; j++) V::store(out_ptr(i, j), V::add(V::load(out_ptr(i, j)), tmp[i][j]));
  } else if (beta == 0.f) {
    typename V::T av = V::splat(alpha);
    for (size_t i = 0; i < MR; i++)
      for (size_t j = 0; j < NR_REGS; j++) V::store(out_ptr(i, j), V::mul(tmp[i][j], av));
  } else {
    typename V::T av = V::splat(alpha), bv = V::splat(beta);
    for (size_t i = 0; i < MR; i++)
      for (size_t j = 0; j < NR_REGS; j++) {
        typename V::T o = V::mul(V::load(out_ptr(i, j)), bv);
        V::store(out_ptr(i, j), V::mul_add(tmp[i][j], av, o));
      }
  }
}

}  // namespace
