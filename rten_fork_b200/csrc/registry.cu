// OpRegistry: `.rten` OperatorNode -> Operator (reference: src/op_registry.rs:31-62 registry, :239-245
// padding_from_attrs, :385-817 per-operator readers).  Attribute defaults follow the reference readers exactly:
// strides / dilations default to [1, 1], pads to [0; 4], auto_pad Same wins over explicit pads.
#include "operators.h"
#include "schema_ids.h"

namespace rtg {

using namespace schema;

const char* OpRegistry::type_name(uint8_t type) { return type < Op_COUNT ? kOperatorTypeNames[type] : "Unknown"; }

rtg_status OpRegistry::read_op(const OpNodeView& node, std::unique_ptr<Operator>* out) const {
  auto it = ops_.find(node.type);
  if (it == ops_.end())
    // ReadOpError::OperatorUnavailable (src/op_registry.rs:205-215)
    return set_errorf(RTG_ERR_FORMAT, "operator %s is not available in this registry", type_name(node.type));
  return it->second(node, out);
}

namespace {

rtg_status attr_error(const char* what) {
  // ReadOpError::AttrError
  return set_errorf(RTG_ERR_FORMAT, "operator attributes missing or invalid (%s)", what);
}

// padding_from_attrs: field ids per attrs table are passed in because they differ between tables.
void read_padding(const fb::Table& a, int auto_pad_field, int pads_field, int32_t* pad_mode, int64_t* pads,
                  int32_t* n_pads) {
  const uint8_t auto_pad = a.scalar<uint8_t>(auto_pad_field, AutoPad_Same);
  fb::Vec<uint32_t> p = a.vec<uint32_t>(pads_field);
  if (auto_pad == AutoPad_Same) {
    *pad_mode = RTG_PAD_SAME;
    *n_pads = 0;
  } else if (p.present) {
    *pad_mode = RTG_PAD_FIXED;
    *n_pads = static_cast<int32_t>(std::min<uint32_t>(p.size(), 4));
    for (int i = 0; i < *n_pads; ++i) pads[i] = p[i];
    if (p.size() > 4) *n_pads = 5;  // reported as "Expected 4 padding values" at run time
  } else {
    *pad_mode = RTG_PAD_FIXED;
    *n_pads = 4;
    for (int i = 0; i < 4; ++i) pads[i] = 0;
  }
}

template <class Op>
ReadOpFn simple(std::function<void(Op*)> init = nullptr) {
  return [init](const OpNodeView&, std::unique_ptr<Operator>* out) -> rtg_status {
    Op* op = new Op();
    if (init) init(op);
    out->reset(op);
    return RTG_OK;
  };
}

ReadOpFn unary(int act, const char* name) {
  return [act, name](const OpNodeView&, std::unique_ptr<Operator>* out) -> rtg_status {
    out->reset(new UnaryOp(act, name));
    return RTG_OK;
  };
}
ReadOpFn binary(int op, const char* name) {
  return [op, name](const OpNodeView&, std::unique_ptr<Operator>* out) -> rtg_status {
    out->reset(new BinaryOp(op, name));
    return RTG_OK;
  };
}

rtg_status read_conv(const OpNodeView& n, std::unique_ptr<Operator>* out) {
  if (n.attrs_type != Attrs_ConvAttrs || !n.attrs.valid) return attr_error("ConvAttrs");
  // ConvAttrs: auto_pad:0 pads:1 groups:2 strides:3 dilations:4   (src/schema.fbs:253-262, reader :432-442)
  ConvOp* op = new ConvOp();
  rtg_conv_attrs& a = op->attrs;
  a.groups = static_cast<int32_t>(n.attrs.scalar<uint32_t>(2, 0));
  read_padding(n.attrs, 0, 1, &a.pad_mode, a.pads, &a.n_pads);
  fb::Vec<uint32_t> s = n.attrs.vec<uint32_t>(3), d = n.attrs.vec<uint32_t>(4);
  a.n_strides = s.present ? static_cast<int32_t>(s.size()) : 2;
  a.n_dilations = d.present ? static_cast<int32_t>(d.size()) : 2;
  for (int i = 0; i < 2; ++i) {
    a.strides[i] = s.present ? (static_cast<uint32_t>(i) < s.size() ? s[i] : 1) : 1;
    a.dilations[i] = d.present ? (static_cast<uint32_t>(i) < d.size() ? d[i] : 1) : 1;
  }
  out->reset(op);
  return RTG_OK;
}

rtg_status read_pool(const OpNodeView& n, std::unique_ptr<Operator>* out, int kind) {
  const uint8_t want = kind == RTG_POOL_MAX ? Attrs_MaxPoolAttrs : Attrs_AveragePoolAttrs;
  if (n.attrs_type != want || !n.attrs.valid) return attr_error("pool attrs");
  // {Max,Average}PoolAttrs: kernel_size:0 auto_pad:1 pads:2 strides:3 (count_include_pad:4)
  fb::Vec<uint32_t> k = n.attrs.vec<uint32_t>(0), s = n.attrs.vec<uint32_t>(3);
  if (!k.present || k.size() < 2) return attr_error("kernel_size");  // array_from_iter panics in the reference
  if (s.present && s.size() < 2) return attr_error("strides");
  PoolOp* op = new PoolOp();
  op->kind = kind;
  op->attrs.kernel[0] = k[0], op->attrs.kernel[1] = k[1];
  op->attrs.strides[0] = s.present ? s[0] : 1, op->attrs.strides[1] = s.present ? s[1] : 1;
  int32_t n_pads = 0;
  read_padding(n.attrs, 1, 2, &op->attrs.pad_mode, op->attrs.pads, &n_pads);
  if (op->attrs.pad_mode == RTG_PAD_FIXED && n_pads != 4) {
    delete op;
    return attr_error("pads");
  }
  op->attrs.count_include_pad = kind == RTG_POOL_AVG ? n.attrs.scalar<uint8_t>(4, 0) : 0;
  out->reset(op);
  return RTG_OK;
}

}  // namespace

const OpRegistry& OpRegistry::with_all_ops() {
  static OpRegistry* reg = [] {
    OpRegistry* r = new OpRegistry();
    r->register_op(Op_Conv, read_conv);
    r->register_op(Op_ConvTranspose, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_ConvTransposeAttrs || !n.attrs.valid) return attr_error("ConvTransposeAttrs");
      // ConvTransposeAttrs: strides:0 auto_pad:1 (default NotSet) pads:2   (src/schema.fbs:264-272, reader :459-466)
      ConvTransposeOp* op = new ConvTransposeOp();
      rtg_conv_transpose_attrs& a = op->attrs;
      const uint8_t auto_pad = n.attrs.scalar<uint8_t>(1, AutoPad_NotSet);
      fb::Vec<uint32_t> p = n.attrs.vec<uint32_t>(2), s = n.attrs.vec<uint32_t>(0);
      if (auto_pad == AutoPad_Same) {
        a.pad_mode = RTG_PAD_SAME;
      } else {
        a.pad_mode = RTG_PAD_FIXED;
        a.n_pads = p.present ? static_cast<int32_t>(std::min<uint32_t>(p.size(), 5)) : 4;
        for (int i = 0; i < 4; ++i) a.pads[i] = (p.present && static_cast<uint32_t>(i) < p.size()) ? p[i] : 0;
      }
      a.n_strides = s.present ? static_cast<int32_t>(std::min<uint32_t>(s.size(), 3)) : 2;  // default [1, 1]
      for (int i = 0; i < 2; ++i) a.strides[i] = s.present ? (static_cast<uint32_t>(i) < s.size() ? s[i] : 1) : 1;
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_MaxPool, [](const OpNodeView& n, std::unique_ptr<Operator>* o) -> rtg_status { return read_pool(n, o, RTG_POOL_MAX); });
    r->register_op(Op_AveragePool, [](const OpNodeView& n, std::unique_ptr<Operator>* o) -> rtg_status { return read_pool(n, o, RTG_POOL_AVG); });
    r->register_op(Op_GlobalAveragePool, simple<PoolOp>([](PoolOp* p) { p->global = true, p->kind = RTG_POOL_AVG; }));
    r->register_op(Op_Relu, unary(ACT_RELU, "Relu"));
    r->register_op(Op_Clip, unary(ACT_CLIP, "Clip"));
    r->register_op(Op_Gelu, unary(ACT_GELU, "Gelu"));
    r->register_op(Op_Sigmoid, unary(ACT_SIGMOID, "Sigmoid"));
    r->register_op(Op_Erf, unary(ACT_ERF, "Erf"));
    r->register_op(Op_Tanh, unary(ACT_TANH, "Tanh"));
    r->register_op(Op_Sqrt, unary(ACT_SQRT, "Sqrt"));
    r->register_op(Op_Add, binary(RTG_BIN_ADD, "Add"));
    r->register_op(Op_Sub, binary(RTG_BIN_SUB, "Sub"));
    r->register_op(Op_Mul, binary(RTG_BIN_MUL, "Mul"));
    r->register_op(Op_Div, binary(RTG_BIN_DIV, "Div"));
    r->register_op(Op_Pow, binary(RTG_BIN_POW, "Pow"));
    r->register_op(Op_Identity, simple<IdentityOp>());
    r->register_op(Op_MatMul, simple<MatMulOp>());
    r->register_op(Op_Gemm, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_GemmAttrs || !n.attrs.valid) return attr_error("GemmAttrs");
      GemmOp* op = new GemmOp();  // GemmAttrs: alpha:0 beta:1 transpose_a:2 transpose_b:3 (reader :500-507)
      op->alpha = n.attrs.scalar<float>(0, 0.f);
      op->beta = n.attrs.scalar<float>(1, 0.f);
      op->trans_a = n.attrs.scalar<uint8_t>(2, 0) != 0;
      op->trans_b = n.attrs.scalar<uint8_t>(3, 0) != 0;
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Flatten, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_FlattenAttrs || !n.attrs.valid) return attr_error("FlattenAttrs");
      ReshapeOp* op = new ReshapeOp();
      op->flatten = true;
      op->axis = n.attrs.scalar<int32_t>(0, 0);
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Reshape, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_ReshapeAttrs || !n.attrs.valid) return attr_error("ReshapeAttrs");
      ReshapeOp* op = new ReshapeOp();
      op->allow_zero = n.attrs.scalar<uint8_t>(0, 0) != 0;
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Softmax, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_SoftmaxAttrs || !n.attrs.valid) return attr_error("SoftmaxAttrs");
      SoftmaxOp* op = new SoftmaxOp();
      op->axis = n.attrs.scalar<int32_t>(0, 0);
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_LayerNormalization, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_LayerNormalizationAttrs || !n.attrs.valid) return attr_error("LayerNormalizationAttrs");
      LayerNormOp* op = new LayerNormOp();
      op->axis = n.attrs.scalar<int32_t>(0, 0);
      op->epsilon = n.attrs.scalar<float>(1, 0.f);
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Transpose, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_TransposeAttrs || !n.attrs.valid) return attr_error("TransposeAttrs");
      TransposeOp* op = new TransposeOp();
      fb::Vec<uint32_t> p = n.attrs.vec<uint32_t>(0);
      op->has_perm = p.present;
      for (uint32_t i = 0; i < p.size(); ++i) op->perm.push_back(p[i]);
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Gather, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_GatherAttrs || !n.attrs.valid) return attr_error("GatherAttrs");
      GatherOp* op = new GatherOp();
      op->axis = n.attrs.scalar<int32_t>(0, 0);
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_BatchNormalization, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_BatchNormalizationAttrs || !n.attrs.valid) return attr_error("BatchNormalizationAttrs");
      BatchNormOp* op = new BatchNormOp();
      op->epsilon = n.attrs.scalar<float>(0, 0.f);
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Cast, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_CastAttrs || !n.attrs.valid) return attr_error("CastAttrs");
      CastOp* op = new CastOp();  // CastAttrs: to:0 (DataType: Int32 = 0, Float = 1; src/schema.fbs)
      const uint8_t to = n.attrs.scalar<uint8_t>(0, DataType_Int32);
      op->to = to == DataType_Float ? RTG_F32 : RTG_I32;
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Concat, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_ConcatAttrs || !n.attrs.valid) return attr_error("ConcatAttrs");
      ConcatOp* op = new ConcatOp();
      op->axis = n.attrs.scalar<int32_t>(0, 0);  // ConcatAttrs: axis:0  (src/schema.fbs:232-234)
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Where, simple<WhereOp>());
    r->register_op(Op_ReduceMean, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_ReduceMeanAttrs || !n.attrs.valid) return attr_error("ReduceMeanAttrs");
      // ReduceMeanAttrs: axes:0 keep_dims:1   (src/schema.fbs:387-390, reduce_axes reader src/op_registry.rs:698-712)
      ReduceMeanOp* op = new ReduceMeanOp();
      fb::Vec<int32_t> a = n.attrs.vec<int32_t>(0);
      op->has_axes = a.present;
      for (uint32_t i = 0; a.present && i < a.size(); ++i) op->axes.push_back(a[i]);
      op->keep_dims = n.attrs.scalar<uint8_t>(1, 0) != 0;
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Einsum, [](const OpNodeView& n, std::unique_ptr<Operator>* out) -> rtg_status {
      if (n.attrs_type != Attrs_EinsumAttrs || !n.attrs.valid) return attr_error("EinsumAttrs");
      EinsumOp* op = new EinsumOp();  // EinsumAttrs: equation:0   (src/schema.fbs:274-276)
      op->equation = n.attrs.str(0);
      out->reset(op);
      return RTG_OK;
    });
    r->register_op(Op_Slice, simple<SliceOp>());
    r->register_op(Op_Shape, simple<ShapeOp>());
    r->register_op(Op_Expand, simple<ExpandOp>());
    r->register_op(Op_Unsqueeze, simple<ShapeOpsOp>());
    r->register_op(Op_Squeeze, simple<ShapeOpsOp>([](ShapeOpsOp* p) { p->squeeze = true; }));
    return r;
  }();
  return *reg;
}

}  // namespace rtg
