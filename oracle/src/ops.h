// ORACLE — TEST INFRASTRUCTURE ONLY.  See common.h.
#pragma once
#include "conv.h"

namespace oracle {

enum UnaryOp { UNARY_RELU = 0, UNARY_GELU = 1, UNARY_SIGMOID = 2, UNARY_ERF = 3, UNARY_SILU = 4, UNARY_CLIP = 5 };
enum BinaryOp { BINARY_ADD = 0, BINARY_SUB = 1, BINARY_MUL = 2, BINARY_DIV = 3 };

float vm_exp(float x);
float vm_sigmoid(float x);
float vm_erf(float x);
float vm_gelu(float x);

void unary(UnaryOp op, const float* in, float* out, size_t n, float lo, float hi);
int binary(BinaryOp op, const float* a, const size_t* a_shape, size_t a_nd, const float* b, const size_t* b_shape,
           size_t b_nd, float* out, size_t* out_shape, size_t* out_nd);

OpError pool2d(bool is_max, const float* input, const size_t in_shape[4], const size_t kernel_size[2],
               const size_t strides[2], const Padding& padding, bool count_include_pad, float* out,
               size_t out_shape[4]);
void global_average_pool(const float* input, size_t batch, size_t chans, size_t in_h, size_t in_w, float* out);

// Strided N-d tensor view used by the MatMul / Gemm operator restatements.
struct TensorView {
  const float* data;
  size_t ndim;
  size_t shape[8];
  ptrdiff_t strides[8];
};

// src/ops/matmul.rs:119-239 matmul / matmul_impl
OpError matmul(const Kernel* gk, const TensorView& a, const TensorView& b, std::vector<float>* out,
               std::vector<size_t>* out_shape);

// src/ops/matmul.rs:27-81 gemm_op
OpError gemm_op(const Kernel* gk, const TensorView& a, const TensorView& b, const TensorView* c, float alpha,
                float beta, bool transpose_a, bool transpose_b, std::vector<float>* out, size_t out_shape[2]);

}  // namespace oracle
