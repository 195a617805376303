// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
// Avx512Kernel: src/gemm/kernels/x86_64.rs:132-227 (MR=6, NR=32, __m512; the
// reference's opt-in `avx512` cargo feature).  Built with -mavx512f -mavx512vl -mfma.
#include <immintrin.h>

#include "simd_kernels.inl"

namespace {
struct VAvx512 {
  using T = __m512;
  static constexpr size_t LEN = 16;
  static T zero() { return _mm512_setzero_ps(); }
  static T splat(float v) { return _mm512_set1_ps(v); }
  static T load(const float* p) { return _mm512_loadu_ps(p); }
  static void store(float* p, T v) { _mm512_storeu_ps(p, v); }
  static T mul_add(T a, T b, T c) { return _mm512_fmadd_ps(a, b, c); }
  static T mul(T a, T b) { return _mm512_mul_ps(a, b); }
  static T add(T a, T b) { return _mm512_add_ps(a, b); }
  static float sum(T v) { return _mm512_reduce_add_ps(v); }
};
void avx512_kernel(float* t, size_t rs, const float* a, const float* b, size_t depth, float alpha, float beta) {
  simd_gemm<VAvx512, 6, 2>(t, rs, a, b, depth, alpha, beta);
}
void avx512_gemv(float* out, size_t out_len, const float* a, size_t a_len, Matrix b, float alpha, float beta) {
  simd_gemv<VAvx512, 2>(out, out_len, a, a_len, b, alpha, beta);
}
}  // namespace

namespace oracle {
extern const Kernel kAvx512Kernel = {KERNEL_AVX512, 6, 32, "avx512", avx512_kernel, avx512_gemv};
}
