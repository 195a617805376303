// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
// Elementwise epilogue ops, MatMul/Gemm operators and the few extra operators the
// benchmark graphs need (pooling), restated from the reference.
#include "ops.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

namespace oracle {

// ---------------------------------------------------------------------------
// rten-vecmath/src/exp.rs:10-137 (Cody-Waite reduction + degree-6 polynomial).
// ---------------------------------------------------------------------------
static inline float bits_to_f(int32_t i) {
  float f;
  std::memcpy(&f, &i, 4);
  return f;
}

float vm_exp(float x) {
  const float INV_LOG2 = 1.44269504088896340736f;  // std::f32::consts::LOG2_E
  const float ROUNDING_MAGIC = 12582912.f;
  const float LOG2_HI = -6.93145752e-1f, LOG2_LO = -1.42860677e-6f;
  const float P0 = 1.0f, P1 = 1.0f, P2 = 4.99999851e-1f, P3 = 1.66664720e-1f, P4 = 4.16695364e-2f,
              P5 = 8.37312452e-3f, P6 = 1.37805939e-3f;
  float j = std::fmaf(x, INV_LOG2, ROUNDING_MAGIC);
  j = j - ROUNDING_MAGIC;
  float r = std::fmaf(j, LOG2_HI, x);
  r = std::fmaf(j, LOG2_LO, r);
  int32_t k = (int32_t)j;  // to_int_trunc
  float tmp = P6;
  tmp = std::fmaf(tmp, r, P5);
  tmp = std::fmaf(tmp, r, P4);
  tmp = std::fmaf(tmp, r, P3);
  tmp = std::fmaf(tmp, r, P2);
  tmp = std::fmaf(tmp, r, P1);
  r = std::fmaf(tmp, r, P0);
  int32_t ia = k > 0 ? 0 : (int32_t)0x83000000u;
  int32_t is = (int32_t)((uint32_t)ia + 0x7f000000u);
  int32_t it = (int32_t)(((uint32_t)k << 23) - (uint32_t)ia);
  r = r * bits_to_f(is);
  r = r * bits_to_f(it);
  if (x >= 104.0f) r = std::numeric_limits<float>::infinity();
  if (x <= -104.0f) r = 0.f;
  return r;
}

// rten-vecmath/src/exp.rs:144-148
float vm_sigmoid(float x) { return 1.0f / (1.0f + vm_exp(-x)); }

// rten-vecmath/src/erf.rs:28-58 (Abramowitz & Stegun 7.1.26)
float vm_erf(float x0) {
  bool neg = x0 < 0.f;
  float x = neg ? -x0 : x0;
  const float p = 0.3275911f;
  const float a0 = 0.254829592f, a1 = -0.284496736f, a2 = 1.421413741f, a3 = -1.453152027f, a4 = 1.061405429f;
  float t = 1.0f / std::fmaf(x, p, 1.0f);
  // poly_eval (rten-simd/src/vec.rs:221-227): Horner then one final multiply by t.
  float y = a4;
  y = std::fmaf(y, t, a3);
  y = std::fmaf(y, t, a2);
  y = std::fmaf(y, t, a1);
  y = std::fmaf(y, t, a0);
  float at = y * t;
  float exp_mx2 = vm_exp(-(x * x));
  float r = 1.0f - at * exp_mx2;
  return neg ? -r : r;
}

// rten-vecmath/src/erf.rs:82-91
float vm_gelu(float x) {
  const float SQRT_2_RCP = 1.0f / 1.41421356237309504880f;
  float half_x = x * 0.5f;
  float y = x * SQRT_2_RCP;
  y = vm_erf(y) + 1.0f;
  return half_x * y;
}

// src/ops/unary_elementwise.rs (Relu :571, Sigmoid/Gelu/Erf/Silu via vecmath, Clip :325-329).
void unary(UnaryOp op, const float* in, float* out, size_t n, float lo, float hi) {
  const size_t CHUNK = 32 * 1024;  // src/ops/unary_elementwise.rs:151-185
  ThreadPool::instance().parallel_for(div_ceil(n, CHUNK), [&](size_t c) {
    size_t i0 = c * CHUNK, i1 = std::min(n, i0 + CHUNK);
    for (size_t i = i0; i < i1; i++) {
      float x = in[i], y;
      switch (op) {
        case UNARY_RELU: y = std::fmax(x, 0.f); break;  // f32::max
        case UNARY_GELU: y = vm_gelu(x); break;
        case UNARY_SIGMOID: y = vm_sigmoid(x); break;
        case UNARY_ERF: y = vm_erf(x); break;
        case UNARY_SILU: y = x * vm_sigmoid(x); break;
        case UNARY_CLIP: y = x < lo ? lo : (x > hi ? hi : x); break;  // f32::clamp
        default: y = x;
      }
      out[i] = y;
    }
  });
}

// src/ops/binary_elementwise.rs:65-157 (general broadcast) — restated with index
// arithmetic; Add/Sub/Mul/Div are exact per element so iteration order is irrelevant.
int binary(BinaryOp op, const float* a, const size_t* a_shape, size_t a_nd, const float* b, const size_t* b_shape,
           size_t b_nd, float* out, size_t* out_shape, size_t* out_nd) {
  size_t nd = std::max(a_nd, b_nd);
  size_t as[8], bs[8], os[8];
  if (nd > 8) return -1;
  for (size_t i = 0; i < nd; i++) {
    as[i] = i + a_nd >= nd ? a_shape[i + a_nd - nd] : 1;
    bs[i] = i + b_nd >= nd ? b_shape[i + b_nd - nd] : 1;
    if (as[i] == bs[i]) os[i] = as[i];
    else if (as[i] == 1) os[i] = bs[i];
    else if (bs[i] == 1) os[i] = as[i];
    else return -1;  // "Cannot broadcast inputs to same shape"
  }
  size_t a_str[8], b_str[8], total = 1;
  size_t sa = 1, sb = 1;
  for (size_t i = nd; i-- > 0;) {
    a_str[i] = as[i] == 1 ? 0 : sa;
    b_str[i] = bs[i] == 1 ? 0 : sb;
    sa *= as[i];
    sb *= bs[i];
    total *= os[i];
  }
  *out_nd = nd;
  for (size_t i = 0; i < nd; i++) out_shape[i] = os[i];
  if (!out) return 0;
  const size_t CHUNK = 32 * 1024;
  ThreadPool::instance().parallel_for(div_ceil(total, CHUNK), [&](size_t c) {
    size_t i0 = c * CHUNK, i1 = std::min(total, i0 + CHUNK);
    for (size_t idx = i0; idx < i1; idx++) {
      size_t rem = idx, ao = 0, bo = 0;
      for (size_t d = nd; d-- > 0;) {
        size_t coord = rem % os[d];
        rem /= os[d];
        ao += coord * a_str[d];
        bo += coord * b_str[d];
      }
      float x = a[ao], y = b[bo], r;
      switch (op) {
        case BINARY_ADD: r = x + y; break;
        case BINARY_SUB: r = x - y; break;
        case BINARY_MUL: r = x * y; break;
        case BINARY_DIV: r = x / y; break;
        default: r = x;
      }
      out[idx] = r;
    }
  });
  return 0;
}

// src/ops/pooling.rs:95-255 pool_impl, :358-375 max_pool, :257-282 average_pool.
OpError pool2d(bool is_max, const float* input, const size_t in_shape[4], const size_t kernel_size[2],
               const size_t strides[2], const Padding& padding, bool count_include_pad, float* out,
               size_t out_shape[4]) {
  const size_t batch = in_shape[0], in_c = in_shape[1], in_h = in_shape[2], in_w = in_shape[3];
  size_t out_h, out_w, fixed[4];
  OpError err = calc_output_size_and_padding(in_h, in_w, kernel_size[0], kernel_size[1], strides[0], strides[1],
                                             padding, 1, 1, &out_h, &out_w, fixed);
  if (!err.ok()) return err;
  const size_t pad_top = fixed[0], pad_left = fixed[1];
  out_shape[0] = batch;
  out_shape[1] = in_c;
  out_shape[2] = out_h;
  out_shape[3] = out_w;
  if (!out) return {};
  const float kernel_len = (float)(kernel_size[0] * kernel_size[1]);
  ThreadPool::instance().parallel_for(batch * in_c, [&](size_t nc) {
    const float* in_view = input + nc * in_h * in_w;
    float* o = out + nc * out_h * out_w;
    for (size_t out_y = 0; out_y < out_h; out_y++)
      for (size_t out_x = 0; out_x < out_w; out_x++) {
        float acc = is_max ? -std::numeric_limits<float>::infinity() : 0.f;
        size_t non_pad = 0;
        for (size_t k_y = 0; k_y < kernel_size[0]; k_y++)
          for (size_t k_x = 0; k_x < kernel_size[1]; k_x++) {
            size_t in_y = out_y * strides[0] + k_y, in_x = out_x * strides[1] + k_x;
            if (in_y >= pad_top && in_y < in_h + pad_top && in_x >= pad_left && in_x < in_w + pad_left) {
              float v = in_view[(in_y - pad_top) * in_w + (in_x - pad_left)];
              acc = is_max ? std::fmax(acc, v) : acc + v;
              non_pad++;
            }
          }
        o[out_y * out_w + out_x] = is_max ? acc : (count_include_pad ? acc / kernel_len : acc / (float)non_pad);
      }
  });
  return {};
}

// src/ops/pooling.rs:284-345: sequential (y, x) sum per channel, divided by h*w.
void global_average_pool(const float* input, size_t batch, size_t chans, size_t in_h, size_t in_w, float* out) {
  // Single-threaded in the reference as well (no Rayon in this function).
  for (size_t nc = 0; nc < batch * chans; nc++) {
    const float* p = input + nc * in_h * in_w;
    float sum = 0.f;
    for (size_t i = 0; i < in_h * in_w; i++) sum += p[i];
    out[nc] = sum / (float)(in_h * in_w);
  }
}

}  // namespace oracle
