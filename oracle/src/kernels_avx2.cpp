// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
// FmaKernel: src/gemm/kernels/x86_64.rs:19-130 (MR=6, NR=16, __m256).  Built with -mavx2 -mfma.
#include <immintrin.h>

#include "simd_kernels.inl"

namespace {
struct VAvx2 {
  using T = __m256;
  static constexpr size_t LEN = 8;
  static T zero() { return _mm256_setzero_ps(); }
  static T splat(float v) { return _mm256_set1_ps(v); }
  static T load(const float* p) { return _mm256_loadu_ps(p); }
  static void store(float* p, T v) { _mm256_storeu_ps(p, v); }
  static T mul_add(T a, T b, T c) { return _mm256_fmadd_ps(a, b, c); }
  static T mul(T a, T b) { return _mm256_mul_ps(a, b); }
  static T add(T a, T b) { return _mm256_add_ps(a, b); }
  // rten-simd/src/arch/x86_64.rs:236-248 horizontal sum order
  static float sum(T v) {
    __m128 hi4 = _mm256_extractf128_ps(v, 1);
    __m128 lo4 = _mm256_castps256_ps128(v);
    __m128 s4 = _mm_add_ps(lo4, hi4);
    __m128 s2 = _mm_add_ps(s4, _mm_movehl_ps(s4, s4));
    __m128 s1 = _mm_add_ps(s2, _mm_shuffle_ps(s2, s2, 0x1));
    return _mm_cvtss_f32(s1);
  }
};
void fma_kernel(float* t, size_t rs, const float* a, const float* b, size_t depth, float alpha, float beta) {
  simd_gemm<VAvx2, 6, 2>(t, rs, a, b, depth, alpha, beta);
}
void fma_gemv(float* out, size_t out_len, const float* a, size_t a_len, Matrix b, float alpha, float beta) {
  simd_gemv<VAvx2, 4>(out, out_len, a, a_len, b, alpha, beta);
}
}  // namespace

namespace oracle {
extern const Kernel kFmaKernel = {KERNEL_FMA, 6, 16, "fma", fma_kernel, fma_gemv};
}
