// Operator classes used by the graph executor (declarations; see operators.cu / operators_nn.cu).
#pragma once
#include "graph.h"

namespace rtg {

// struct Conv {groups, dilations, padding, strides}  src/ops/conv.rs:282-311
class ConvOp : public Operator {
 public:
  rtg_conv_attrs attrs = {};
  // fused epilogue (set by optimize_graph): out = act(conv + bias + residual); input 3 is the residual
  bool has_residual = false;
  int act = ACT_NONE;
  float act_lo = 0.f, act_hi = 0.f;
  const char* name() const override { return act != ACT_NONE || has_residual ? "FusedConv" : "Conv"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
  rtg_status prepare(Context* ctx, const OpInputs& consts) override;
  void release(Context* ctx) override;

 private:
  rtg_status partial_plan(const rtg_tensor& w, ConvPlan* p) const;
  ConvWeights weights;
  bool prepared = false, prepared_pointwise = false, prepared_depthwise = false;
  const void* prepared_from = nullptr;
};

// struct ConvTranspose {padding, strides}  src/ops/conv.rs:547-577
class ConvTransposeOp : public Operator {
 public:
  rtg_conv_transpose_attrs attrs = {};
  const char* name() const override { return "ConvTranspose"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};

// UnaryFloatOp blanket impl  src/ops/unary_elementwise.rs:22-63 ; Clip :335-383
class UnaryOp : public Operator {
 public:
  int act;
  float act_lo = 0.f, act_hi = 0.f;
  const char* op_name;
  UnaryOp(int a, const char* n) : act(a), op_name(n) {}
  const char* name() const override { return op_name; }
  bool can_run_in_place() const override { return true; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
  rtg_status run_in_place(Context* ctx, Value* io, const OpInputs& rest) override;

 private:
  rtg_status bounds(Context* ctx, const Value* lo_v, const Value* hi_v, float* lo, float* hi) const;
};

// Add / Sub / Mul / Div  src/ops/binary_elementwise.rs:395-560
class BinaryOp : public Operator {
 public:
  int op;
  const char* op_name;
  BinaryOp(int o, const char* n) : op(o), op_name(n) {}
  const char* name() const override { return op_name; }
  bool can_run_in_place() const override { return true; }
  bool is_commutative() const override { return op == RTG_BIN_ADD || op == RTG_BIN_MUL; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
  rtg_status run_in_place(Context* ctx, Value* io, const OpInputs& rest) override;
};

// MaxPool / AveragePool / GlobalAveragePool  src/ops/pooling.rs
class PoolOp : public Operator {
 public:
  int kind = RTG_POOL_MAX;
  bool global = false;
  rtg_pool_attrs attrs = {};
  const char* name() const override { return global ? "GlobalAveragePool" : kind == RTG_POOL_MAX ? "MaxPool" : "AveragePool"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};

// Flatten {axis} / Reshape {allow_zero}  src/ops/layout.rs:131-340
class ReshapeOp : public Operator {
 public:
  bool flatten = false;
  int64_t axis = 1;
  bool allow_zero = false;
  const char* name() const override { return flatten ? "Flatten" : "Reshape"; }
  bool can_run_in_place() const override { return true; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
  rtg_status run_in_place(Context* ctx, Value* io, const OpInputs& rest) override;

 private:
  rtg_status target_shape(Context* ctx, const rtg_tensor& x, const Value* shape_v, int64_t* shape, int* nd) const;
};

class IdentityOp : public Operator {
 public:
  const char* name() const override { return "Identity"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};

// struct Gemm {alpha, beta, transpose_a, transpose_b}  src/ops/matmul.rs:13-19,83-117
class GemmOp : public Operator {
 public:
  float alpha = 1.f, beta = 1.f;
  bool trans_a = false, trans_b = false;
  const char* name() const override { return "Gemm"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
  // constant weights (the classifier head): packed once at load, like MatMulOp
  rtg_status prepare(Context* ctx, const OpInputs& consts) override;
  void release(Context* ctx) override;

 private:
  BPack b_pack;
  const void* b_pack_from = nullptr;
};

// MatMul  src/ops/matmul.rs:241-254 ; FusedTranspose(MatMul) is expressed by a TransposeOp whose result is a view
class MatMulOp : public Operator {
 public:
  bool has_bias = false;  // fused `+ bias[N]` (input 2) and activation, set by optimize_graph
  // fused `(MatMul [+ bias]) -> Add(residual)` (input 3: an activation of the output's shape; the residual connections of
  // a transformer block).  Same order as the unfused chain: (acc + bias) + residual, then the activation
  bool has_residual = false;
  int act = ACT_NONE;
  // fused `MatMul -> Mul(scalar)` / `MatMul -> Div(scalar)` (attention score scaling)
  float scale = 1.f;
  bool scale_div = false;
  // FusedTranspose(MatMul) (src/ops/fused.rs:46-78, GraphOptimizer::fuse_transpose src/optimize.rs:333-377): the
  // operand is consumed through a permuted view instead of a materialised Transpose
  bool has_perm[2] = {false, false};
  bool perm_reverse[2] = {false, false};  // Transpose without `perm`: reverse the dims
  std::vector<int64_t> perm[2];
  bool fused() const { return has_bias || has_residual || act != ACT_NONE || scale != 1.f || has_perm[0] || has_perm[1]; }
  const char* name() const override { return fused() ? "FusedMatMul" : "MatMul"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
  // a constant 2-D right operand (linear-layer weights) gets its packed copy (BPack) once at load (make_b_pack),
  // the counterpart of prepack_b for RTG_MATH_TF32_BF16X2
  rtg_status prepare(Context* ctx, const OpInputs& consts) override;
  void release(Context* ctx) override;

 private:
  BPack b_pack;
  const void* b_pack_from = nullptr;
  int64_t b_pack_dims[4] = {0, 0, 0, 0};  // K, N, row stride, column stride of the operand it was made from
};

// operators_nn.cu: the remaining operators BERT-class graphs need
class SoftmaxOp : public Operator {
 public:
  int64_t axis = -1;
  bool has_addend = false;  // GraphOptimizer: `Add -> Softmax` fused, the other Add operand is input 1
  const char* name() const override { return "Softmax"; }
  bool can_run_in_place() const override { return true; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
  rtg_status run_in_place(Context* ctx, Value* io, const OpInputs& rest) override;
};
class LayerNormOp : public Operator {
 public:
  int64_t axis = -1;
  float epsilon = 1e-5f;
  const char* name() const override { return "LayerNormalization"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class TransposeOp : public Operator {
 public:
  std::vector<int64_t> perm;
  bool has_perm = false;
  const char* name() const override { return "Transpose"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class GatherOp : public Operator {
 public:
  int64_t axis = 0;
  const char* name() const override { return "Gather"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
// Cast  src/ops/convert.rs: f32 <-> i32 (`as` conversion: truncation toward zero, saturating, NaN -> 0)
class CastOp : public Operator {
 public:
  int to = RTG_F32;
  const char* name() const override { return "Cast"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class BatchNormOp : public Operator {
 public:
  float epsilon = 1e-5f;
  const char* name() const override { return "BatchNormalization"; }
  bool can_run_in_place() const override { return true; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
  rtg_status run_in_place(Context* ctx, Value* io, const OpInputs& rest) override;
};
// operators_layout.cu: index-driven copies (Concat, Slice, Shape, Expand)
class ConcatOp : public Operator {  // src/ops/concat.rs
 public:
  int64_t axis = 0;
  const char* name() const override { return "Concat"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class SliceOp : public Operator {  // src/ops/slice.rs
 public:
  const char* name() const override { return "Slice"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class ShapeOp : public Operator {  // src/ops/layout.rs Shape
 public:
  const char* name() const override { return "Shape"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class ExpandOp : public Operator {  // src/ops/layout.rs Expand
 public:
  const char* name() const override { return "Expand"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class WhereOp : public Operator {  // src/ops/binary_elementwise.rs where_op
 public:
  const char* name() const override { return "Where"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class ReduceMeanOp : public Operator {  // src/ops/reduce.rs reduce_mean
 public:
  std::vector<int64_t> axes;
  bool has_axes = false, keep_dims = true;
  const char* name() const override { return "ReduceMean"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class EinsumOp : public Operator {  // src/ops/einsum.rs; einsum.cu
 public:
  std::string equation;
  const char* name() const override { return "Einsum"; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
};
class ShapeOpsOp : public Operator {  // Unsqueeze / Squeeze: metadata-only changes of a contiguous tensor
 public:
  bool squeeze = false;
  const char* name() const override { return squeeze ? "Squeeze" : "Unsqueeze"; }
  bool can_run_in_place() const override { return true; }
  rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) override;
  rtg_status run_in_place(Context* ctx, Value* io, const OpInputs& rest) override;

 private:
  rtg_status new_shape(Context* ctx, const rtg_tensor& x, const Value* axes, int64_t* shape, int* nd) const;
};

}  // namespace rtg
