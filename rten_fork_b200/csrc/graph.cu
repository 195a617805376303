// Graph planner and executor (reference: src/graph.rs:733-1073 run/run_plan, :1256-1346 create_plan).
#include "graph.h"

#include <algorithm>
#include <set>

namespace rtg {

rtg_status OpInputs::require(size_t i, const Value** out) const {
  if (i >= v.size() || v[i] == nullptr) return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  *out = v[i];
  return RTG_OK;
}

rtg_status alloc_value(Context* ctx, int dtype, int ndim, const int64_t* shape, bool nhwc, Value* out) {
  Value v;
  v.t.dtype = dtype;
  v.t.ndim = ndim;
  v.t.device = ctx->device;
  int64_t n = 1;
  for (int i = 0; i < ndim; ++i) {
    v.t.shape[i] = shape[i];
    n *= shape[i];
  }
  set_contiguous_strides(v.t);
  if (nhwc && ndim == 4) {
    const int64_t C = shape[1], H = shape[2], W = shape[3];
    v.t.strides[0] = H * W * C, v.t.strides[1] = 1, v.t.strides[2] = W * C, v.t.strides[3] = C;
  }
  void* p = nullptr;
  RTG_TRY(ctx->pool.alloc(static_cast<size_t>(std::max<int64_t>(n, 1)) * 4, &p));
  v.t.data = p;
  v.owned = true;
  *out = v;
  return RTG_OK;
}

void release_value(Context* ctx, Value* v) {
  if (v->owned && v->t.data) ctx->pool.free(v->t.data);
  v->t.data = nullptr;
  v->owned = false;
}

rtg_status host_ints(Context* ctx, const Value& v, std::vector<int64_t>* out) {
  out->clear();
  if (v.t.dtype != RTG_I32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  const int64_t n = numel(v.t);
  if (v.host_i32 && static_cast<int64_t>(v.host_i32->size()) == n) {
    out->assign(v.host_i32->begin(), v.host_i32->end());
    return RTG_OK;
  }
  if (!is_contiguous(v.t)) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "shape-like tensors must be contiguous");
  std::vector<int32_t> tmp(static_cast<size_t>(n));
  if (n > 0) {
    RTG_CUDA(cudaMemcpyAsync(tmp.data(), v.t.data, static_cast<size_t>(n) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    RTG_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  out->assign(tmp.begin(), tmp.end());
  return RTG_OK;
}

rtg_status host_floats(Context* ctx, const Value& v, std::vector<float>* out) {
  out->clear();
  if (v.t.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  const int64_t n = numel(v.t);
  if (v.host_f32 && static_cast<int64_t>(v.host_f32->size()) == n) {
    *out = *v.host_f32;
    return RTG_OK;
  }
  if (!is_contiguous(v.t)) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "scalar-like tensors must be contiguous");
  out->resize(static_cast<size_t>(n));
  if (n > 0) {
    RTG_CUDA(cudaMemcpyAsync(out->data(), v.t.data, static_cast<size_t>(n) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    RTG_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  return RTG_OK;
}

int Graph::num_operators() const {
  int n = 0;
  for (const Node& nd : nodes) n += nd.kind == NodeKind::Operator;
  return n;
}

void Graph::release(Context* ctx) {
  for (Node& nd : nodes) {
    if (nd.kind == NodeKind::Operator && nd.op) nd.op->release(ctx);
    if (nd.kind == NodeKind::Constant) release_value(ctx, &nd.constant);
  }
}

// Depth-first from the requested outputs; an operator is scheduled once all of its inputs are resolved.
int32_t Graph::source_node(int32_t id) const {
  const int32_t n = static_cast<int32_t>(nodes.size());
  for (int32_t i = 0; i < n; ++i)
    if (nodes[i].kind == NodeKind::Operator)
      for (int32_t o : nodes[i].outputs)
        if (o == id) return i;
  return -1;
}

rtg_status Graph::create_plan(const std::vector<int32_t>& inputs, const std::vector<int32_t>& outputs,
                              std::vector<int32_t>* plan, bool allow_missing) const {
  plan->clear();
  const int32_t n = static_cast<int32_t>(nodes.size());
  std::vector<int32_t> producer(nodes.size(), -1);  // value node id -> operator node id
  for (int32_t i = 0; i < n; ++i)
    if (nodes[i].kind == NodeKind::Operator)
      for (int32_t o : nodes[i].outputs)
        if (o >= 0 && o < n) producer[o] = i;
  std::vector<char> resolved(nodes.size(), 0), scheduled(nodes.size(), 0);
  for (int32_t i = 0; i < n; ++i)
    if (nodes[i].kind == NodeKind::Constant) resolved[i] = 1;
  for (int32_t i : inputs) {
    if (i < 0 || i >= n) return set_error(RTG_ERR_INVALID_VALUE, "Input node ID is invalid");
    resolved[i] = 1;
  }
  // iterative post-order DFS over operators
  struct Frame {
    int32_t op;
    size_t next;
  };
  for (int32_t out : outputs) {
    if (out < 0 || out >= n) return set_error(RTG_ERR_INVALID_VALUE, "Output node ID is invalid");
    if (resolved[out]) continue;
    if (producer[out] < 0) {
      if (allow_missing) continue;
      return set_errorf(RTG_ERR_INVALID_VALUE, "Missing input \"%s\" for op \"<output>\"", nodes[out].name.c_str());
    }
    std::vector<Frame> stack;
    stack.push_back(Frame{producer[out], 0});
    while (!stack.empty()) {
      Frame& f = stack.back();
      const Node& op = nodes[f.op];
      if (scheduled[f.op]) {
        stack.pop_back();
        continue;
      }
      bool descended = false;
      while (f.next < op.inputs.size()) {
        const int32_t in = op.inputs[f.next++];
        if (in < 0 || resolved[in]) continue;
        if (in >= n) return set_error(RTG_ERR_INVALID_VALUE, "operator input id out of range");
        if (producer[in] < 0) {
          if (allow_missing) continue;  // a value nobody can provide: prune_plan drops what depends on it
          return set_errorf(RTG_ERR_INVALID_VALUE, "Missing input \"%s\" for op \"%s\"", nodes[in].name.c_str(),
                            op.name.c_str());
        }
        if (!scheduled[producer[in]]) {
          stack.push_back(Frame{producer[in], 0});
          descended = true;
          break;
        }
      }
      if (descended) continue;
      scheduled[f.op] = 1;
      for (int32_t o : op.outputs)
        if (o >= 0) resolved[o] = 1;
      plan->push_back(f.op);
      stack.pop_back();
    }
  }
  return RTG_OK;
}

static std::string shape_str(const rtg_tensor& t) {
  std::string s = "[";
  for (int i = 0; i < t.ndim; ++i) s += (i ? "," : "") + std::to_string(t.shape[i]);
  return s + "]";
}

rtg_status Graph::run(Context* ctx, const std::vector<std::pair<int32_t, Value>>& inputs,
                      const std::vector<int32_t>& outputs, std::vector<Value>* results,
                      std::vector<TimingRecord>* timing) {
  std::vector<int32_t> in_ids;
  for (auto& kv : inputs) in_ids.push_back(kv.first);
  const int32_t nn = static_cast<int32_t>(nodes.size());
  for (int32_t o : outputs)
    if (o < 0 || o >= nn) {
      for (auto& kv : inputs) {
        Value v = kv.second;
        release_value(ctx, &v);
      }
      return set_error(RTG_ERR_INVALID_VALUE, "Output node ID is invalid");
    }
  // get_cached_plan (src/graph.rs:768-795): re-plan only when the input / output ids change.  The cache is shared by
  // concurrent runs: it is read and replaced under the mutex, the plan itself is immutable once published.
  std::shared_ptr<const std::vector<int32_t>> plan_ptr;
  {
    std::lock_guard<std::mutex> lock(plan_mu_);
    if (!cached_plan_ || in_ids != cached_inputs_ || outputs != cached_outputs_) {
      auto fresh = std::make_shared<std::vector<int32_t>>();
      rtg_status st = create_plan(in_ids, outputs, fresh.get());
      if (st != RTG_OK) {
        for (auto& kv : inputs) {
          Value v = kv.second;
          release_value(ctx, &v);
        }
        return st;
      }
      cached_plan_ = fresh;
      cached_inputs_ = in_ids;
      cached_outputs_ = outputs;
    }
    plan_ptr = cached_plan_;
  }
  return run_plan(ctx, inputs, *plan_ptr, outputs, results, timing);
}

void Graph::prune_plan(const std::vector<int32_t>& plan, const std::vector<int32_t>& inputs,
                       const std::vector<int32_t>& outputs, std::vector<int32_t>* pruned,
                       std::vector<int32_t>* leaves) const {
  std::vector<char> resolved(nodes.size(), 0), feeds_pruned(nodes.size(), 0);
  for (size_t i = 0; i < nodes.size(); ++i)
    if (nodes[i].kind == NodeKind::Constant) resolved[i] = 1;
  for (int32_t i : inputs) resolved[i] = 1;
  std::vector<int32_t> candidates;
  for (int32_t op_id : plan) {
    const Node& op = nodes[op_id];
    bool all = true;
    for (int32_t in : op.inputs)
      if (in >= 0 && !resolved[in]) all = false;
    if (!all) {  // every operator here is deterministic (Operator::is_deterministic, src/ops/mod.rs:850-861)
      for (int32_t in : op.inputs)
        if (in >= 0 && resolved[in]) feeds_pruned[in] = 1;
      continue;
    }
    for (int32_t o : op.outputs)
      if (o >= 0) {
        resolved[o] = 1;
        candidates.push_back(o);
      }
    pruned->push_back(op_id);
  }
  for (int32_t c : candidates)
    if (feeds_pruned[c] || std::find(outputs.begin(), outputs.end(), c) != outputs.end()) leaves->push_back(c);
}

rtg_status Graph::partial_run(Context* ctx, const std::vector<std::pair<int32_t, Value>>& inputs,
                              const std::vector<int32_t>& outputs, std::vector<std::pair<int32_t, Value>>* leaves) {
  leaves->clear();
  auto drop_inputs = [&]() {
    for (auto& kv : inputs) {
      Value v = kv.second;
      release_value(ctx, &v);
    }
  };
  std::vector<int32_t> in_ids, plan, pruned, leaf_ids;
  for (auto& kv : inputs) in_ids.push_back(kv.first);
  rtg_status st = create_plan(in_ids, outputs, &plan, /*allow_missing=*/true);
  if (st != RTG_OK) {
    drop_inputs();
    return st;
  }
  prune_plan(plan, in_ids, outputs, &pruned, &leaf_ids);
  std::vector<Value> vals;
  RTG_TRY(run_plan(ctx, inputs, pruned, leaf_ids, &vals, nullptr));
  for (size_t i = 0; i < leaf_ids.size(); ++i) leaves->emplace_back(leaf_ids[i], vals[i]);
  return RTG_OK;
}

rtg_status Graph::run_plan(Context* ctx, const std::vector<std::pair<int32_t, Value>>& inputs,
                           const std::vector<int32_t>& plan, const std::vector<int32_t>& outputs,
                           std::vector<Value>* results, std::vector<TimingRecord>* timing) {
  const size_t n = nodes.size();

  // values produced during this run (operator outputs and owned inputs)
  std::vector<Value> temp(n);
  std::vector<char> live(n, 0);
  for (auto& kv : inputs) {
    temp[kv.first] = kv.second;
    live[kv.first] = 1;
  }
  // usage counts (src/graph.rs:842-862)
  std::vector<int32_t> refs(n, 0);
  for (int32_t op : plan)
    for (int32_t in : nodes[op].inputs)
      if (in >= 0) refs[in]++;
  for (int32_t o : outputs) refs[o]++;

  auto fail = [&](rtg_status st) {
    for (size_t i = 0; i < n; ++i)
      if (live[i]) release_value(ctx, &temp[i]);
    return st;
  };
  auto lookup = [&](int32_t id) -> const Value* {
    if (id < 0) return nullptr;
    if (nodes[id].kind == NodeKind::Constant) return &nodes[id].constant;
    return live[id] ? &temp[id] : nullptr;
  };

  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  if (timing) {
    cudaEventCreate(&ev0);
    cudaEventCreate(&ev1);
  }

  for (int32_t op_id : plan) {
    Node& node = nodes[op_id];
    Operator* op = node.op.get();
    // ---- in-place selection (src/graph.rs:897-931): input 0, or for commutative binary ops the largest eligible
    // operand; eligible = an owned temporary nobody else will read.
    int inplace_idx = -1;
    if (op->can_run_in_place() && !node.inputs.empty()) {
      auto eligible = [&](size_t i) {
        const int32_t id = node.inputs[i];
        return id >= 0 && nodes[id].kind != NodeKind::Constant && live[id] && temp[id].owned && refs[id] == 1;
      };
      if (op->is_commutative() && node.inputs.size() == 2) {
        int64_t best = -1;
        for (size_t i = 0; i < 2; ++i)
          if (eligible(i) && numel(temp[node.inputs[i]].t) > best) {
            best = numel(temp[node.inputs[i]].t);
            inplace_idx = static_cast<int>(i);
          }
      } else if (eligible(0)) {
        inplace_idx = 0;
      }
    }
    OpInputs in;
    for (size_t i = 0; i < node.inputs.size(); ++i) {
      const int32_t id = node.inputs[i];
      const Value* v = lookup(id);
      if (id >= 0 && v == nullptr)
        return fail(set_errorf(RTG_ERR_INVALID_VALUE, "Missing input \"%s\" for op \"%s\"", nodes[id].name.c_str(),
                               node.name.c_str()));
      in.v.push_back(v);
    }
    if (timing) cudaEventRecord(ev0, ctx->stream);
    std::vector<Value> outs;
    rtg_status st;
    std::string in_shapes;
    if (timing)
      for (const Value* v : in.v) in_shapes += v ? shape_str(v->t) + " " : "- ";
    if (inplace_idx >= 0) {
      const int32_t io_id = node.inputs[inplace_idx];
      Value io = temp[io_id];
      OpInputs rest;
      for (size_t i = 0; i < in.v.size(); ++i)
        if (static_cast<int>(i) != inplace_idx) rest.v.push_back(in.v[i]);
      st = op->run_in_place(ctx, &io, rest);
      if (st == RTG_OK) {
        live[io_id] = 0;  // ownership moved to the output
        refs[io_id]--;
        outs.push_back(io);
      }
    } else {
      st = op->run(ctx, in, &outs);
    }
    if (st != RTG_OK) {
      // RunError::OperatorError{name, error} (src/graph.rs:975-978): keep the status, prefix the node name
      std::string msg = rtg_last_error();
      for (Value& v : outs) release_value(ctx, &v);
      return fail(set_errorf(st, "operator \"%s\" (%s) failed: %s", node.name.c_str(), op->name(), msg.c_str()));
    }
    if (timing) {
      cudaEventRecord(ev1, ctx->stream);
      cudaEventSynchronize(ev1);
      float ms = 0;
      cudaEventElapsedTime(&ms, ev0, ev1);
      timing->push_back(TimingRecord{op->name(), node.name, ms, in_shapes});
    }
    if (outs.size() < node.outputs.size())
      return fail(set_errorf(RTG_ERR_INTERNAL, "operator \"%s\" returned too few outputs", node.name.c_str()));
    for (size_t i = 0; i < outs.size(); ++i) {
      const int32_t oid = i < node.outputs.size() ? node.outputs[i] : -1;
      if (oid < 0 || refs[oid] == 0) {
        release_value(ctx, &outs[i]);
        continue;
      }
      temp[oid] = outs[i];
      live[oid] = 1;
    }
    // drop inputs that nobody needs any more (src/graph.rs:1013-1037)
    for (size_t i = 0; i < node.inputs.size(); ++i) {
      const int32_t id = node.inputs[i];
      if (id < 0 || static_cast<int>(i) == inplace_idx) continue;
      if (--refs[id] == 0 && live[id]) {
        release_value(ctx, &temp[id]);
        live[id] = 0;
      }
    }
  }
  if (timing) {
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
  }

  results->clear();
  for (int32_t o : outputs) {
    const Value* v = lookup(o);
    if (!v) return fail(set_errorf(RTG_ERR_INVALID_VALUE, "output \"%s\" was not produced", nodes[o].name.c_str()));
    Value r = *v;
    if (nodes[o].kind == NodeKind::Constant || --refs[o] > 0 || !live[o]) {
      // constants, or values requested more than once: hand out an owned copy
      Value c;
      rtg_status st = alloc_value(ctx, r.t.dtype, r.t.ndim, r.t.shape, false, &c);
      if (st != RTG_OK) return fail(st);
      st = copy_view(ctx, c.t, r.t);
      if (st != RTG_OK) return fail(st);
      c.host_i32 = r.host_i32, c.host_f32 = r.host_f32;
      results->push_back(c);
    } else {
      live[o] = 0;
      results->push_back(r);
    }
  }
  for (size_t i = 0; i < n; ++i)
    if (live[i]) release_value(ctx, &temp[i]);
  return RTG_OK;
}

rtg_status propagate_constants(Context* ctx, Graph* g, int* replaced) {
  *replaced = 0;
  std::vector<std::pair<int32_t, Value>> leaves;
  RTG_TRY(g->partial_run(ctx, {}, g->output_ids, &leaves));
  rtg_status st = RTG_OK;
  for (auto& kv : leaves) {
    Value v = kv.second;
    if (st == RTG_OK && !is_contiguous(v.t)) {  // constants are stored in the reference layout
      Value c;
      st = alloc_value(ctx, v.t.dtype, v.t.ndim, v.t.shape, false, &c);
      if (st == RTG_OK) st = copy_view(ctx, c.t, v.t);
      if (st == RTG_OK) {
        c.host_i32 = v.host_i32, c.host_f32 = v.host_f32;
        release_value(ctx, &v);
        v = c;
      } else {
        release_value(ctx, &c);
      }
    }
    const int64_t cnt = numel(v.t);
    if (st == RTG_OK && cnt > 0 && cnt <= 4096) {  // host mirror, like the constants of the file (make_constant)
      if (v.t.dtype == RTG_I32 && !v.host_i32) {
        std::vector<int64_t> tmp;
        st = host_ints(ctx, v, &tmp);
        if (st == RTG_OK) v.host_i32 = std::make_shared<std::vector<int32_t>>(tmp.begin(), tmp.end());
      } else if (v.t.dtype == RTG_F32 && !v.host_f32) {
        std::vector<float> tmp;
        st = host_floats(ctx, v, &tmp);
        if (st == RTG_OK) v.host_f32 = std::make_shared<std::vector<float>>(tmp);
      }
    }
    if (st != RTG_OK) {
      release_value(ctx, &v);
      continue;
    }
    // The value node becomes the constant (node ids stay stable; the reference adds a node and rewires,
    // GraphMutator::replace_value).  Its producer is never planned again: constants are resolved from the start.
    Node& nd = g->nodes[kv.first];
    nd.kind = NodeKind::Constant;
    nd.constant = v;
    ++*replaced;
  }
  return st;
}

}  // namespace rtg
