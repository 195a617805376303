// `.rten` loading (reference: src/header.rs:60-110 V2 container, src/model.rs:265-522 load_impl / load_graph /
// add_graph_* , docs/rten-file-format.md) and the load-time graph optimiser.
#include <stdio.h>

#include <fstream>
#include <set>

#include "operators.h"
#include "schema_ids.h"

namespace rtg {

using namespace schema;

namespace {

struct Header {
  bool present = false;
  uint64_t model_offset = 0, model_len = 0, tensor_data_offset = 0;
};

// Header::from_buf.  Returns RTG_OK with present=false for headerless (V1) files, i.e. when the magic is absent.
rtg_status read_header(const uint8_t* p, size_t n, Header* h) {
  if (n < 4 || memcmp(p, "RTEN", 4) != 0) return RTG_OK;  // HeaderError::InvalidMagic => V1 format
  if (n < 32) return set_error(RTG_ERR_FORMAT, "invalid header: TooShort");
  uint32_t version;
  memcpy(&version, p + 4, 4);
  if (version != 2) return set_error(RTG_ERR_FORMAT, "invalid header: UnsupportedVersion");
  memcpy(&h->model_offset, p + 8, 8);
  memcpy(&h->model_len, p + 16, 8);
  memcpy(&h->tensor_data_offset, p + 24, 8);
  if (h->model_offset < 32 || h->model_offset > n) return set_error(RTG_ERR_FORMAT, "invalid header: InvalidOffset");
  if (h->model_len > n - h->model_offset) return set_error(RTG_ERR_FORMAT, "invalid header: InvalidLength");
  if (h->tensor_data_offset < 32 || h->tensor_data_offset > n)
    return set_error(RTG_ERR_FORMAT, "invalid header: InvalidOffset");
  h->present = true;
  return RTG_OK;
}

// Upload a constant; small tensors also keep a host mirror for shape-consuming operators.
rtg_status make_constant(Context* ctx, int dtype, const std::vector<int64_t>& shape, const void* host, Node* node) {
  Value v;
  int64_t n = 1;
  for (int64_t d : shape)
    if (d < 0 || __builtin_mul_overflow(n, d, &n) || n > (INT64_MAX >> 2))
      return set_error(RTG_ERR_FORMAT, "graph error: constant data length does not match shape");
  if (shape.size() > RTG_MAX_DIMS) return set_error(RTG_ERR_FORMAT, "constant has too many dimensions");
  RTG_TRY(alloc_value(ctx, dtype, static_cast<int>(shape.size()), shape.data(), false, &v));
  if (n > 0) RTG_CUDA(cudaMemcpyAsync(v.t.data, host, static_cast<size_t>(n) * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (n <= 4096) {
    if (dtype == RTG_I32)
      v.host_i32 = std::make_shared<std::vector<int32_t>>(static_cast<const int32_t*>(host), static_cast<const int32_t*>(host) + n);
    else
      v.host_f32 = std::make_shared<std::vector<float>>(static_cast<const float*>(host), static_cast<const float*>(host) + n);
  }
  node->constant = v;
  return RTG_OK;
}

}  // namespace

rtg_status load_model(Context* ctx, const uint8_t* data, size_t len, bool optimize, Model** out) {
  Header hdr;
  RTG_TRY(read_header(data, len, &hdr));
  const uint8_t* model_data = hdr.present ? data + hdr.model_offset : data;
  const size_t model_len = hdr.present ? static_cast<size_t>(hdr.model_len) : len;
  std::unique_ptr<Model> model(new Model());
  model->ctx = ctx;
  const OpRegistry& registry = OpRegistry::with_all_ops();
  // every error return below hands the constants already uploaded (and whatever the operators prepared) back to the pool
  struct GraphGuard {
    Context* ctx;
    Graph* g;
    bool armed = true;
    ~GraphGuard() {
      if (armed) g->release(ctx);
    }
  } guard{ctx, &model->graph};
  try {
    if (model_len < 8) throw fb::Error("file too short");
    fb::Table root = fb::root(model_data, model_len);
    // Model: schema_version:0 graph:1 metadata:2
    if (root.scalar<int32_t>(0, 0) != 1) return set_error(RTG_ERR_FORMAT, "unsupported schema version");
    fb::Table graph = root.table(1);
    if (!graph.valid) return set_error(RTG_ERR_FORMAT, "parse error: model has no graph");
    // Graph: nodes:0 inputs:1 outputs:2 captures:3
    fb::Vec<uint32_t> nodes = graph.vec<uint32_t>(0);
    Graph& g = model->graph;
    g.nodes.resize(nodes.size());
    for (uint32_t i = 0; i < nodes.size(); ++i) {
      fb::Table nt = graph.vec_table(nodes, i);
      Node& node = g.nodes[i];
      node.name = nt.str(0);  // Node: name:0 data_type:1 data:2
      const uint8_t kind = nt.scalar<uint8_t>(1, Kind_NONE);
      fb::Table d = nt.table(2);
      if (kind == Kind_OperatorNode) {
        // OperatorNode: type:0 attrs_type:1 attrs:2 inputs:3 outputs:4   (add_graph_operator, src/model.rs:384-445)
        node.kind = NodeKind::Operator;
        OpNodeView view;
        view.type = d.scalar<uint8_t>(0, 0);
        view.attrs_type = d.scalar<uint8_t>(1, 0);
        view.attrs = d.table(2);
        node.op_type = view.type;
        rtg_status st = registry.read_op(view, &node.op);
        if (st != RTG_OK) {
          std::string msg = rtg_last_error();
          return set_errorf(st, "operator error: in node \"%s\": %s", node.name.c_str(), msg.c_str());
        }
        fb::Vec<int32_t> ins = d.vec<int32_t>(3), outs = d.vec<int32_t>(4);
        for (uint32_t k = 0; k < ins.size(); ++k) node.inputs.push_back(ins[k]);
        for (uint32_t k = 0; k < outs.size(); ++k) node.outputs.push_back(outs[k]);
        for (int32_t id : node.inputs)
          if (id >= static_cast<int32_t>(nodes.size())) return set_error(RTG_ERR_FORMAT, "graph error: operator input id out of range");
        for (int32_t id : node.outputs)
          if (id >= static_cast<int32_t>(nodes.size())) return set_error(RTG_ERR_FORMAT, "graph error: operator output id out of range");
      } else if (kind == Kind_ConstantNode) {
        // ConstantNode: shape:0 data_type:1 data:2 dtype:3 data_offset:4   (add_graph_constant, src/model.rs:468-522)
        node.kind = NodeKind::Constant;
        fb::Vec<uint32_t> shp = d.vec<uint32_t>(0);
        std::vector<int64_t> shape;
        int64_t n = 1;
        for (uint32_t k = 0; k < shp.size(); ++k) {
          shape.push_back(shp[k]);
          // a product that leaves int64 (or whose byte count does) cannot match any data length
          if (__builtin_mul_overflow(n, static_cast<int64_t>(shp[k]), &n) || n > (INT64_MAX >> 2))
            return set_error(RTG_ERR_FORMAT, "graph error: constant data length does not match shape");
        }
        const uint8_t data_kind = d.scalar<uint8_t>(1, Data_NONE);
        if (d.has(4)) {
          // tensor data stored after the model buffer (V2 only)
          if (!hdr.present) return set_error(RTG_ERR_FORMAT, "graph error: tensor data section missing");
          uint64_t off = 0;
          if (__builtin_add_overflow(static_cast<uint64_t>(hdr.tensor_data_offset), d.scalar<uint64_t>(4, 0), &off))
            return set_error(RTG_ERR_FORMAT, "graph error: invalid tensor data offset");
          if (!d.has(3)) return set_error(RTG_ERR_FORMAT, "graph error: constant with data offset has no dtype");
          const uint16_t dt = d.scalar<uint16_t>(3, 0);
          if (dt > 1) return set_error(RTG_ERR_FORMAT, "graph error: unsupported data type for external constant");
          if (off > len || static_cast<uint64_t>(n) * 4 > len - off)
            return set_error(RTG_ERR_FORMAT, "graph error: invalid tensor data offset");
          RTG_TRY(make_constant(ctx, dt == 0 ? RTG_I32 : RTG_F32, shape, data + off, &node));
        } else if (data_kind == Data_FloatData || data_kind == Data_IntData) {
          fb::Table dt = d.table(2);
          fb::Vec<uint32_t> raw = dt.vec<uint32_t>(0);  // 4-byte elements either way
          if (!raw.present || static_cast<int64_t>(raw.size()) != n)
            return set_error(RTG_ERR_FORMAT, "graph error: constant data length does not match shape");
          RTG_TRY(make_constant(ctx, data_kind == Data_IntData ? RTG_I32 : RTG_F32, shape, raw.data(), &node));
        } else {
          return set_error(RTG_ERR_FORMAT, "graph error: unsupported constant data type");
        }
      } else if (kind == Kind_ValueNode) {
        // ValueNode: shape:0 ([Dim{value:0, name:1}])   (add_graph_value, src/model.rs:447-466)
        node.kind = NodeKind::Value;
        fb::Vec<uint32_t> dims = d.vec<uint32_t>(0);
        node.has_shape = dims.present;
        for (uint32_t k = 0; k < dims.size(); ++k) {
          fb::Table dim = d.vec_table(dims, k);
          DimSpec ds;
          ds.name = dim.str(1);
          ds.value = ds.name.empty() ? static_cast<int64_t>(dim.scalar<uint32_t>(0, 0)) : -1;
          node.shape.push_back(ds);
        }
      } else {
        return set_error(RTG_ERR_FORMAT, "graph error: unknown node type");
      }
    }
    fb::Vec<uint32_t> ins = graph.vec<uint32_t>(1), outs = graph.vec<uint32_t>(2);
    for (uint32_t i = 0; i < ins.size(); ++i) g.input_ids.push_back(static_cast<int32_t>(ins[i]));
    for (uint32_t i = 0; i < outs.size(); ++i) g.output_ids.push_back(static_cast<int32_t>(outs[i]));
    for (int32_t id : g.input_ids)
      if (id < 0 || id >= static_cast<int32_t>(g.nodes.size())) return set_error(RTG_ERR_FORMAT, "graph error: input id out of range");
    for (int32_t id : g.output_ids)
      if (id < 0 || id >= static_cast<int32_t>(g.nodes.size())) return set_error(RTG_ERR_FORMAT, "graph error: output id out of range");
  } catch (const fb::Error& e) {
    return set_errorf(RTG_ERR_FORMAT, "parse error: %s", e.what());
  }
  RTG_CUDA(cudaStreamSynchronize(ctx->stream));  // constants are uploaded from the caller's buffer

  if (optimize) {
    // GraphOptimizer::optimize (src/optimize.rs:286-297): constant propagation first, then the fusions
    int replaced = 0;
    rtg_status st = propagate_constants(ctx, &model->graph, &replaced);
    if (st != RTG_OK) {
      std::string msg = rtg_last_error();
      return set_errorf(st, "optimization failed: %s", msg.c_str());  // ModelLoadError::OptimizeError
    }
    optimize_graph(&model->graph);
  }

  // Let operators pre-arrange constant operands (weights) once.
  Graph& g = model->graph;
  for (Node& node : g.nodes) {
    if (node.kind != NodeKind::Operator) continue;
    OpInputs consts;
    for (int32_t id : node.inputs)
      consts.v.push_back(id >= 0 && g.nodes[id].kind == NodeKind::Constant ? &g.nodes[id].constant : nullptr);
    RTG_TRY(node.op->prepare(ctx, consts));
  }
  guard.armed = false;
  *out = model.release();
  return RTG_OK;
}

// ---------------------------------------------------------------- optimiser
// The reference's GraphOptimizer fuses Transpose->MatMul, Silu, Gelu and LayerNormalization patterns
// (src/optimize.rs:286-297).  On the GPU the profitable fusions are the epilogues, because every separate
// elementwise operator costs a full read + write of the activation tensor:
//     Conv -> [Add(residual)] -> [Relu | Clip]          => one FusedConv launch
//     MatMul -> Add(bias vector) -> [Gelu | Relu | ...]  => one FusedMatMul launch
// Each fused node computes exactly the same expression in the same order as the unfused chain, so results stay
// within the parity tolerance (the intermediate values are simply never stored).
int optimize_graph(Graph* g) {
  const int32_t n = static_cast<int32_t>(g->nodes.size());
  std::vector<int32_t> uses(n, 0), producer(n, -1);
  std::vector<std::vector<int32_t>> consumers(n);
  std::vector<char> dead(n, 0);
  auto rebuild = [&]() {
    std::fill(uses.begin(), uses.end(), 0);
    std::fill(producer.begin(), producer.end(), -1);
    for (auto& c : consumers) c.clear();
    for (int32_t i = 0; i < n; ++i) {
      Node& nd = g->nodes[i];
      if (nd.kind != NodeKind::Operator || dead[i]) continue;
      for (int32_t in : nd.inputs)
        if (in >= 0) {
          uses[in]++;
          consumers[in].push_back(i);
        }
      for (int32_t o : nd.outputs)
        if (o >= 0) producer[o] = i;
    }
    for (int32_t o : g->output_ids) uses[o] += 1000;  // graph outputs must stay materialised
  };
  rebuild();
  int fused = 0;

  // ---- Gelu: x * (Erf(x / sqrt(2)) + 1) * 0.5  =>  Gelu(x)      (GraphOptimizer::fuse_gelu, src/optimize.rs:401-424)
  auto scalar_const = [&](int32_t id, float want) {
    if (id < 0) return false;
    const Node& c = g->nodes[id];
    if (c.kind != NodeKind::Constant || !c.constant.host_f32 || c.constant.host_f32->size() != 1) return false;
    const float v = (*c.constant.host_f32)[0];
    return fabsf(v - want) <= 1e-6f * fabsf(want);
  };
  auto scalar_value = [&](int32_t id, float* v) {
    if (id < 0) return false;
    const Node& c = g->nodes[id];
    if (c.kind != NodeKind::Constant || !c.constant.host_f32 || c.constant.host_f32->size() != 1) return false;
    *v = (*c.constant.host_f32)[0];
    return true;
  };
  auto bin_node = [&](int32_t value, int op) -> Node* {  // the binary operator of kind `op` producing `value`
    if (value < 0 || producer[value] < 0 || dead[producer[value]]) return nullptr;
    Node& nd = g->nodes[producer[value]];
    BinaryOp* b = dynamic_cast<BinaryOp*>(nd.op.get());
    return (b && b->op == op && nd.inputs.size() == 2 && nd.outputs.size() == 1) ? &nd : nullptr;
  };
  for (int32_t i = 0; i < n; ++i) {
    Node& half = g->nodes[i];
    if (half.kind != NodeKind::Operator || dead[i]) continue;
    BinaryOp* hb = dynamic_cast<BinaryOp*>(half.op.get());
    if (!hb || hb->op != RTG_BIN_MUL || half.inputs.size() != 2 || half.outputs.size() != 1) continue;
    const int hc = scalar_const(half.inputs[1], 0.5f) ? 1 : (scalar_const(half.inputs[0], 0.5f) ? 0 : -1);
    if (hc < 0) continue;
    const int32_t t_val = half.inputs[1 - hc];
    Node* mul = bin_node(t_val, RTG_BIN_MUL);
    if (!mul || uses[t_val] != 1) continue;
    for (int side = 0; side < 2; ++side) {
      const int32_t x = mul->inputs[side], a_val = mul->inputs[1 - side];
      Node* add = bin_node(a_val, RTG_BIN_ADD);
      if (!add || uses[a_val] != 1) continue;
      const int ac = scalar_const(add->inputs[1], 1.f) ? 1 : (scalar_const(add->inputs[0], 1.f) ? 0 : -1);
      if (ac < 0) continue;
      const int32_t e_val = add->inputs[1 - ac];
      if (e_val < 0 || producer[e_val] < 0 || uses[e_val] != 1) continue;
      Node& erf = g->nodes[producer[e_val]];
      UnaryOp* eu = dynamic_cast<UnaryOp*>(erf.op.get());
      if (!eu || eu->act != ACT_ERF || erf.inputs.size() != 1) continue;
      const int32_t d_val = erf.inputs[0];
      Node* div = bin_node(d_val, RTG_BIN_DIV);
      if (!div || uses[d_val] != 1 || div->inputs[0] != x || !scalar_const(div->inputs[1], 1.41421356237f)) continue;
      // rewrite: the final Mul becomes Gelu(x); the four inner operators die
      dead[producer[t_val]] = dead[producer[a_val]] = dead[producer[e_val]] = dead[producer[d_val]] = 1;
      half.op.reset(new UnaryOp(ACT_GELU, "Gelu"));
      half.inputs.assign(1, x);
      ++fused;
      break;
    }
  }
  if (fused) rebuild();

  // ---- Silu: x * Sigmoid(x)  =>  Silu(x)      (GraphOptimizer::fuse_silu, src/optimize.rs:379-399).  The reference has
  //      no serialised Silu operator: this fusion is the only way one comes to exist.  Same arithmetic as the pair
  //      (x * sigmoid(x), rten-vecmath/src/exp.rs:194-197), so fused and unfused results are bit-identical.
  {
    int sf = 0;
    for (int32_t i = 0; i < n; ++i) {
      Node& mul = g->nodes[i];
      if (mul.kind != NodeKind::Operator || dead[i]) continue;
      BinaryOp* mb = dynamic_cast<BinaryOp*>(mul.op.get());
      if (!mb || mb->op != RTG_BIN_MUL || mul.inputs.size() != 2 || mul.outputs.size() != 1) continue;
      for (int side = 0; side < 2; ++side) {
        const int32_t x = mul.inputs[side], s_val = mul.inputs[1 - side];
        if (x < 0 || s_val < 0 || producer[s_val] < 0 || dead[producer[s_val]]) continue;
        Node& sig = g->nodes[producer[s_val]];
        UnaryOp* su = dynamic_cast<UnaryOp*>(sig.op.get());
        if (!su || su->act != ACT_SIGMOID || sig.inputs.size() != 1 || sig.inputs[0] != x) continue;
        if (uses[s_val] == 1) dead[producer[s_val]] = 1;  // otherwise the Sigmoid stays for its other consumers
        mul.op.reset(new UnaryOp(ACT_SILU, "Silu"));
        mul.inputs.assign(1, x);
        ++sf;
        break;
      }
    }
    if (sf) rebuild();
    fused += sf;
  }

  // ---- LayerNormalization spelled out as primitives      (GraphOptimizer::fuse_layer_norm, src/optimize.rs:427-513)
  //        c = x - ReduceMean(x, axes=[-1])
  //        n = c / Sqrt(epsilon + ReduceMean(Pow(c, 2), axes=[-1]))
  //        y = n * scale + bias                 (scale, bias, epsilon constants)
  //      => LayerNormalization{axis = -1, epsilon}(x, scale, bias), matched from the final Add backwards like the
  //      reference's three patterns.  One pass over x instead of nine.
  {
    auto is_const = [&](int32_t id) { return id >= 0 && g->nodes[id].kind == NodeKind::Constant; };
    auto unary_node = [&](int32_t value, int act) -> Node* {
      if (value < 0 || producer[value] < 0 || dead[producer[value]]) return nullptr;
      Node& nd = g->nodes[producer[value]];
      UnaryOp* u = dynamic_cast<UnaryOp*>(nd.op.get());
      return (u && u->act == act && nd.inputs.size() >= 1 && nd.outputs.size() == 1) ? &nd : nullptr;
    };
    // a ReduceMean over the last axis given as -1 (attribute or constant input), the only form the reference fuses
    auto mean_last_axis = [&](int32_t value) -> Node* {
      if (value < 0 || producer[value] < 0 || dead[producer[value]]) return nullptr;
      Node& nd = g->nodes[producer[value]];
      ReduceMeanOp* r = dynamic_cast<ReduceMeanOp*>(nd.op.get());
      if (!r || nd.inputs.empty() || nd.outputs.size() != 1) return nullptr;
      if (r->has_axes) return (r->axes.size() == 1 && r->axes[0] == -1) ? &nd : nullptr;
      if (nd.inputs.size() > 1 && is_const(nd.inputs[1])) {
        const Value& ax = g->nodes[nd.inputs[1]].constant;
        if (ax.host_i32 && ax.host_i32->size() == 1 && (*ax.host_i32)[0] == -1) return &nd;
      }
      return nullptr;
    };
    int lf = 0;
    for (int32_t i = 0; i < n; ++i) {
      Node& fin = g->nodes[i];
      if (fin.kind != NodeKind::Operator || dead[i]) continue;
      BinaryOp* fb_ = dynamic_cast<BinaryOp*>(fin.op.get());
      if (!fb_ || fb_->op != RTG_BIN_ADD || fin.inputs.size() != 2 || fin.outputs.size() != 1) continue;
      bool done = false;
      for (int a_side = 0; a_side < 2 && !done; ++a_side) {           // (n * scale) + bias, either operand order
        const int32_t bias = fin.inputs[a_side], m_val = fin.inputs[1 - a_side];
        if (!is_const(bias)) continue;
        Node* mul = bin_node(m_val, RTG_BIN_MUL);
        if (!mul) continue;
        for (int m_side = 0; m_side < 2 && !done; ++m_side) {
          const int32_t scale = mul->inputs[m_side], n_val = mul->inputs[1 - m_side];
          if (!is_const(scale)) continue;
          Node* div = bin_node(n_val, RTG_BIN_DIV);                    // c / Sqrt(...)
          if (!div) continue;
          const int32_t c_val = div->inputs[0];
          Node* sq = unary_node(div->inputs[1], ACT_SQRT);
          if (!sq) continue;
          Node* add_eps = bin_node(sq->inputs[0], RTG_BIN_ADD);        // epsilon + ReduceMean(...), either order
          if (!add_eps) continue;
          for (int e_side = 0; e_side < 2 && !done; ++e_side) {
            const int32_t eps_id = add_eps->inputs[e_side], var_val = add_eps->inputs[1 - e_side];
            float eps = 0.f;
            if (!scalar_value(eps_id, &eps)) continue;
            Node* var_mean = mean_last_axis(var_val);
            if (!var_mean) continue;
            Node* pw = bin_node(var_mean->inputs[0], RTG_BIN_POW);     // Pow(c, 2)
            if (!pw || pw->inputs[0] != c_val || !scalar_const(pw->inputs[1], 2.f)) continue;
            Node* sub = bin_node(c_val, RTG_BIN_SUB);                  // x - ReduceMean(x)
            if (!sub) continue;
            const int32_t x = sub->inputs[0];
            Node* ctr_mean = mean_last_axis(sub->inputs[1]);
            if (!ctr_mean || ctr_mean->inputs[0] != x) continue;
            LayerNormOp* ln = new LayerNormOp();
            ln->axis = -1;
            ln->epsilon = eps;
            fin.op.reset(ln);
            fin.inputs = {x, scale, bias};
            ++lf;
            done = true;
          }
        }
      }
    }
    if (lf) {
      // operators that only fed the fused expression die (the planner would never schedule them anyway)
      rebuild();
      bool changed = true;
      while (changed) {
        changed = false;
        for (int32_t i = 0; i < n; ++i) {
          Node& nd = g->nodes[i];
          if (nd.kind != NodeKind::Operator || dead[i] || nd.outputs.empty()) continue;
          bool used = false;
          for (int32_t o : nd.outputs)
            if (o >= 0 && uses[o] > 0) used = true;
          if (!used) {
            dead[i] = 1;
            changed = true;
          }
        }
        if (changed) rebuild();
      }
    }
    fused += lf;
  }

  // ---- Transpose -> MatMul  =>  FusedTranspose(MatMul): the permuted view feeds the TMA maps directly
  //      (GraphOptimizer::fuse_transpose, src/optimize.rs:333-377)
  {
    int tf = 0;
    for (int32_t i = 0; i < n; ++i) {
      Node& nd = g->nodes[i];
      if (nd.kind != NodeKind::Operator || dead[i]) continue;
      MatMulOp* mm = dynamic_cast<MatMulOp*>(nd.op.get());
      if (!mm || nd.inputs.size() < 2) continue;
      for (int side = 0; side < 2; ++side) {
        const int32_t v = nd.inputs[side];
        if (v < 0 || producer[v] < 0 || dead[producer[v]] || mm->has_perm[side]) continue;
        Node& tn = g->nodes[producer[v]];
        TransposeOp* tp = dynamic_cast<TransposeOp*>(tn.op.get());
        if (!tp || tn.inputs.size() != 1 || tn.inputs[0] < 0) continue;
        nd.inputs[side] = tn.inputs[0];
        mm->has_perm[side] = true;
        mm->perm_reverse[side] = !tp->has_perm;
        mm->perm[side] = tp->perm;
        if (--uses[v] == 0) dead[producer[v]] = 1;
        ++tf;
      }
    }
    if (tf) rebuild();
    fused += tf;
  }
  auto single_consumer = [&](int32_t value) -> int32_t {
    if (value < 0 || uses[value] != 1 || consumers[value].size() != 1) return -1;
    return consumers[value][0];
  };
  auto act_of = [&](const Node& nd, int* act, float* lo, float* hi) -> bool {
    UnaryOp* u = dynamic_cast<UnaryOp*>(nd.op.get());
    if (!u) return false;
    *lo = 0.f, *hi = 0.f;
    if (u->act == ACT_CLIP) {
      // min / max must be constants to be folded into the epilogue
      *lo = -3.40282347e+38f, *hi = 3.40282347e+38f;
      for (size_t k = 1; k < nd.inputs.size() && k < 3; ++k) {
        const int32_t id = nd.inputs[k];
        if (id < 0) continue;
        const Node& c = g->nodes[id];
        if (c.kind != NodeKind::Constant || !c.constant.host_f32 || c.constant.host_f32->size() != 1) return false;
        (k == 1 ? *lo : *hi) = (*c.constant.host_f32)[0];
      }
    }
    *act = u->act;
    return true;
  };
  for (int32_t i = 0; i < n; ++i) {
    Node& nd = g->nodes[i];
    if (nd.kind != NodeKind::Operator || dead[i] || nd.outputs.size() != 1 || nd.outputs[0] < 0) continue;
    ConvOp* conv = dynamic_cast<ConvOp*>(nd.op.get());
    MatMulOp* mm = dynamic_cast<MatMulOp*>(nd.op.get());
    if (!conv && !mm) continue;
    if (conv && (conv->has_residual || conv->act != ACT_NONE)) continue;
    int32_t cur_out = nd.outputs[0];
    int32_t next = single_consumer(cur_out);
    // --- MatMul -> Mul / Div by a scalar constant: the epilogue's alpha (the unfused operators would cost a read and
    //     a write of the whole score tensor)
    if (mm && next >= 0 && !dead[next] && mm->scale == 1.f) {
      Node& sn = g->nodes[next];
      BinaryOp* b = dynamic_cast<BinaryOp*>(sn.op.get());
      float c = 1.f;
      bool ok = false, div = false;
      if (b && sn.inputs.size() == 2 && sn.outputs.size() == 1) {
        if (b->op == RTG_BIN_DIV && sn.inputs[0] == cur_out && scalar_value(sn.inputs[1], &c)) ok = div = true;
        else if (b->op == RTG_BIN_MUL && sn.inputs[0] == cur_out && scalar_value(sn.inputs[1], &c)) ok = true;
        else if (b->op == RTG_BIN_MUL && sn.inputs[1] == cur_out && scalar_value(sn.inputs[0], &c)) ok = true;
      }
      if (ok && c != 1.f) {
        mm->scale = c;
        mm->scale_div = div;
        dead[next] = 1;
        nd.outputs[0] = sn.outputs[0];
        cur_out = sn.outputs[0];
        ++fused;
      }
    }
    // --- optional Add
    next = single_consumer(cur_out);
    if (next >= 0 && !dead[next]) {
      Node& add = g->nodes[next];
      BinaryOp* b = dynamic_cast<BinaryOp*>(add.op.get());
      if (b && b->op == RTG_BIN_ADD && add.inputs.size() == 2 && add.outputs.size() == 1) {
        const int32_t other = add.inputs[0] == cur_out ? add.inputs[1] : add.inputs[0];
        bool ok = other >= 0 && other != cur_out;
        if (ok && conv) {
          // the residual must not depend on this convolution (it is a sibling branch or the block input)
          ok = g->nodes[other].kind != NodeKind::Constant && nd.inputs.size() <= 3;
          // cheap cycle check: the residual's producer must come from an earlier value than our output
          std::vector<int32_t> stack = {other};
          std::set<int32_t> seen;
          while (ok && !stack.empty()) {
            const int32_t v = stack.back();
            stack.pop_back();
            if (v == cur_out) ok = false;
            if (v < 0 || !seen.insert(v).second || producer[v] < 0) continue;
            for (int32_t in : g->nodes[producer[v]].inputs) stack.push_back(in);
          }
          if (ok) {
            while (nd.inputs.size() < 3) nd.inputs.push_back(-1);
            nd.inputs.push_back(other);
            conv->has_residual = true;
          }
        } else if (ok && mm) {
          // bias must be a constant vector [N]
          const Node& c = g->nodes[other];
          ok = c.kind == NodeKind::Constant && c.constant.t.ndim == 1 && c.constant.t.dtype == RTG_F32 && nd.inputs.size() == 2;
          if (ok) {
            nd.inputs.push_back(other);
            mm->has_bias = true;
          }
        }
        if (ok) {
          dead[next] = 1;
          nd.outputs[0] = add.outputs[0];
          cur_out = add.outputs[0];
          ++fused;
        }
      }
    }
    // --- MatMul [+ bias] -> Add(activation): the residual connection of a transformer block rides in the epilogue
    //     ((acc + bias) + residual, the order of the unfused chain).  The operand must not depend on this product.
    next = single_consumer(cur_out);
    if (mm && !mm->has_residual && mm->act == ACT_NONE && next >= 0 && !dead[next]) {
      Node& add = g->nodes[next];
      BinaryOp* b = dynamic_cast<BinaryOp*>(add.op.get());
      if (b && b->op == RTG_BIN_ADD && add.inputs.size() == 2 && add.outputs.size() == 1) {
        const int32_t other = add.inputs[0] == cur_out ? add.inputs[1] : add.inputs[0];
        bool ok = other >= 0 && other != cur_out && g->nodes[other].kind != NodeKind::Constant &&
                  nd.inputs.size() <= 3;
        // an Add that feeds a Softmax (attention mask) is left to the Add -> Softmax fusion below: the mask only
        // broadcasts, which the epilogue cannot take, and the Softmax kernel adds it for free
        const int32_t after = add.outputs[0] >= 0 ? single_consumer(add.outputs[0]) : -1;
        if (after >= 0 && !dead[after] && dynamic_cast<SoftmaxOp*>(g->nodes[after].op.get())) ok = false;
        std::vector<int32_t> stack = {other};
        std::set<int32_t> seen;
        while (ok && !stack.empty()) {
          const int32_t v = stack.back();
          stack.pop_back();
          if (v == cur_out) ok = false;
          if (v < 0 || !seen.insert(v).second || producer[v] < 0) continue;
          for (int32_t in : g->nodes[producer[v]].inputs) stack.push_back(in);
        }
        if (ok) {
          while (nd.inputs.size() < 3) nd.inputs.push_back(-1);
          nd.inputs.push_back(other);
          mm->has_residual = true;
          dead[next] = 1;
          nd.outputs[0] = add.outputs[0];
          cur_out = add.outputs[0];
          ++fused;
        }
      }
    }
    // --- optional activation
    next = single_consumer(cur_out);
    if (next >= 0 && !dead[next] && (conv || mm->has_bias || mm->has_residual)) {
      Node& an = g->nodes[next];
      int act;
      float lo, hi;
      if (an.outputs.size() == 1 && act_of(an, &act, &lo, &hi) &&
          (act == ACT_RELU || act == ACT_CLIP || act == ACT_GELU || act == ACT_SIGMOID || act == ACT_TANH || act == ACT_SILU)) {
        if (conv) conv->act = act, conv->act_lo = lo, conv->act_hi = hi;
        else if (act != ACT_CLIP) mm->act = act;
        if (conv || act != ACT_CLIP) {
          dead[next] = 1;
          nd.outputs[0] = an.outputs[0];
          ++fused;
        }
      }
    }
  }
  // --- Add -> Softmax (the attention mask of a transformer block): the add happens in the Softmax kernel's registers,
  //     one pass over the scores instead of two.  Same bits: the kernel performs the same round-to-nearest add.
  for (int32_t i = 0; i < n; ++i) {
    Node& nd = g->nodes[i];
    if (nd.kind != NodeKind::Operator || dead[i] || nd.outputs.size() != 1 || nd.outputs[0] < 0) continue;
    BinaryOp* b = dynamic_cast<BinaryOp*>(nd.op.get());
    if (!b || b->op != RTG_BIN_ADD || nd.inputs.size() != 2 || nd.inputs[0] < 0 || nd.inputs[1] < 0) continue;
    const int32_t next = single_consumer(nd.outputs[0]);
    if (next < 0 || dead[next]) continue;
    Node& sn = g->nodes[next];
    SoftmaxOp* sm = dynamic_cast<SoftmaxOp*>(sn.op.get());
    if (!sm || sm->has_addend || sn.inputs.size() != 1 || sn.inputs[0] != nd.outputs[0]) continue;
    sn.inputs = {nd.inputs[0], nd.inputs[1]};
    sm->has_addend = true;
    dead[i] = 1;
    ++fused;
  }
  // dead operators become value-less placeholders (ids of other nodes must not shift)
  for (int32_t i = 0; i < n; ++i)
    if (dead[i]) {
      g->nodes[i].kind = NodeKind::Value;
      g->nodes[i].op.reset();
      g->nodes[i].inputs.clear();
      g->nodes[i].outputs.clear();
    }
  return fused;
}

}  // namespace rtg
