// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
// BaseKernel: src/gemm/kernels.rs:413-479 (MR=8, NR=4, f32 as a 1-lane "SIMD" type).
#include "simd_kernels.inl"

namespace {
struct VScalar {
  using T = float;
  static constexpr size_t LEN = 1;
  static T zero() { return 0.f; }
  static T splat(float v) { return v; }
  static T load(const float* p) { return *p; }
  static void store(float* p, T v) { *p = v; }
  static T mul_add(T a, T b, T c) { return std::fmaf(a, b, c); }  // f32::mul_add is fused
  static T mul(T a, T b) { return a * b; }
  static T add(T a, T b) { return a + b; }
  static float sum(T v) { return v; }
};
void base_kernel(float* t, size_t rs, const float* a, const float* b, size_t depth, float alpha, float beta) {
  simd_gemm<VScalar, 8, 4>(t, rs, a, b, depth, alpha, beta);
}
void base_gemv(float* out, size_t out_len, const float* a, size_t a_len, Matrix b, float alpha, float beta) {
  simd_gemv<VScalar, 4>(out, out_len, a, a_len, b, alpha, beta);
}
}  // namespace

namespace oracle {
extern const Kernel kBaseKernel = {KERNEL_BASE, 8, 4, "base", base_kernel, base_gemv};
}
