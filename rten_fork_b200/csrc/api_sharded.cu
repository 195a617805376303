// C ABI: batch-sharded Model::run inside ONE process (the reference's own scaling advice is the
// same shape: "use Rayon to execute the model on separate chunks in parallel", docs/performance.md:137-139).
// One batch is split along dim 0 of every graph input into contiguous chunks, one per GPU; every GPU has its own
// context (stream + device pool), a replica of the weights made at load, and a host worker thread; the chunks run
// concurrently and the outputs are laid side by side in the caller's host buffers.  No collective: the graph must be
// separable by batch item, which is checked operator by operator at load.
#include <condition_variable>
#include <mutex>
#include <thread>

#include "operators.h"
#include "schema_ids.h"

using namespace rtg;

namespace {

struct Shard {
  rtg_ctx* ctx = nullptr;
  rtg_model* model = nullptr;
  std::thread worker;
  std::mutex mu;
  std::condition_variable cv;
  bool has_job = false, done = false, quit = false;
  // job
  std::vector<rtg_tensor> inputs;
  std::vector<rtg_tensor> outputs;  // host destinations (already offset to this shard's rows)
  int64_t rows = 0;
  rtg_status status = RTG_OK;
  std::string error;
};

struct Sharded {
  std::vector<std::unique_ptr<Shard>> shards;
  int32_t n_inputs = 0, n_outputs = 0;
};

// Is every operator of the graph separable by batch item (dim 0 of the graph inputs)?  Conservative: an operator is
// accepted only in a form that provably keeps dim 0 as the batch; anything else refuses the model.
rtg_status check_batch_separable(const Graph& g) {
  const int32_t n = static_cast<int32_t>(g.nodes.size());
  std::vector<char> batch(n, 0);  // value carries the batch in dim 0
  for (int32_t id : g.input_ids) batch[id] = 1;
  auto is_const = [&](int32_t id) { return id >= 0 && g.nodes[id].kind == NodeKind::Constant; };
  // plan order = node order for operators whose inputs precede them; iterate to a fixed point to be safe
  std::vector<int32_t> plan;
  RTG_TRY(g.create_plan(g.input_ids, g.output_ids, &plan));
  for (int32_t op_id : plan) {
    const Node& nd = g.nodes[op_id];
    bool any_batch = false;
    for (int32_t in : nd.inputs)
      if (in >= 0 && batch[in]) any_batch = true;
    if (!any_batch) continue;  // constants only (should have been propagated): does not touch the batch
    auto refuse = [&](const char* why) {
      return set_errorf(RTG_ERR_UNSUPPORTED_VALUE, "model is not separable by batch item: operator \"%s\" (%s) %s",
                        nd.name.c_str(), nd.op->name(), why);
    };
    const Operator* op = nd.op.get();
    bool ok = false;
    if (dynamic_cast<const ConvOp*>(op) || dynamic_cast<const ConvTransposeOp*>(op) || dynamic_cast<const PoolOp*>(op) ||
        dynamic_cast<const BatchNormOp*>(op)) {
      ok = nd.inputs[0] >= 0 && batch[nd.inputs[0]];
      for (size_t i = 1; i < nd.inputs.size() && ok; ++i) {
        const int32_t in = nd.inputs[i];
        if (in < 0 || is_const(in)) continue;
        // a fused residual (input 3 of FusedConv) is an activation of the same batch
        ok = batch[in] != 0;
      }
      if (!ok) return refuse("mixes a per-batch value with a non-constant operand");
    } else if (dynamic_cast<const UnaryOp*>(op) || dynamic_cast<const IdentityOp*>(op) || dynamic_cast<const CastOp*>(op)) {
      ok = true;
    } else if (dynamic_cast<const BinaryOp*>(op) || dynamic_cast<const WhereOp*>(op)) {
      // every non-constant operand must carry the batch; a constant must not have a leading extent that indexes it
      for (int32_t in : nd.inputs) {
        if (in < 0) continue;
        if (is_const(in)) {
          const rtg_tensor& t = g.nodes[in].constant.t;
          if (t.ndim >= 1 && t.shape[0] != 1 && t.ndim >= 4) return refuse("has a constant operand with a batch extent");
        } else if (!batch[in]) {
          return refuse("mixes a per-batch value with a batch-free one");
        }
      }
      ok = true;
    } else if (const ReshapeOp* r = dynamic_cast<const ReshapeOp*>(op)) {
      if (!r->flatten || r->axis < 1) return refuse("folds the batch dimension");
      ok = true;
    } else if (const GemmOp* ge = dynamic_cast<const GemmOp*>(op)) {
      if (ge->trans_a || !batch[nd.inputs[0]] || !is_const(nd.inputs[1]) ||
          (nd.inputs.size() > 2 && nd.inputs[2] >= 0 && !is_const(nd.inputs[2])))
        return refuse("is not activation x constant weights");
      ok = true;
    } else if (const MatMulOp* mm = dynamic_cast<const MatMulOp*>(op)) {
      if (mm->has_perm[0] || !batch[nd.inputs[0]] || !is_const(nd.inputs[1])) return refuse("is not activation x constant weights");
      // a fused bias is a constant; a fused residual (input 3) is an activation of the same batch
      for (size_t i = 2; i < nd.inputs.size(); ++i)
        if (nd.inputs[i] >= 0 && !is_const(nd.inputs[i]) && !batch[nd.inputs[i]])
          return refuse("mixes a per-batch value with a batch-free one");
      ok = true;
    } else if (const SoftmaxOp* sm = dynamic_cast<const SoftmaxOp*>(op)) {
      if (sm->axis == 0) return refuse("normalises over the batch dimension");
      // a fused `Add -> Softmax` keeps the Add's rule for both operands
      for (int32_t in : nd.inputs) {
        if (in < 0) continue;
        if (is_const(in)) {
          const rtg_tensor& t = g.nodes[in].constant.t;
          if (t.ndim >= 1 && t.shape[0] != 1 && t.ndim >= 4) return refuse("has a constant operand with a batch extent");
        } else if (!batch[in]) {
          return refuse("mixes a per-batch value with a batch-free one");
        }
      }
      ok = true;
    } else if (const LayerNormOp* ln = dynamic_cast<const LayerNormOp*>(op)) {
      if (ln->axis == 0) return refuse("normalises over the batch dimension");
      ok = true;
    } else if (const TransposeOp* tp = dynamic_cast<const TransposeOp*>(op)) {
      if (!tp->has_perm || tp->perm.empty() || tp->perm[0] != 0) return refuse("moves the batch dimension");
      ok = true;
    }
    if (!ok) return refuse("is not known to keep batch items independent");
    for (int32_t o : nd.outputs)
      if (o >= 0) batch[o] = 1;
  }
  for (int32_t o : g.output_ids)
    if (!batch[o] && !is_const(o))
      return set_errorf(RTG_ERR_UNSUPPORTED_VALUE, "model is not separable by batch item: output \"%s\" does not carry the batch",
                        g.nodes[o].name.c_str());
  return RTG_OK;
}

void worker_loop(Shard* s) {
  for (;;) {
    std::unique_lock<std::mutex> lock(s->mu);
    s->cv.wait(lock, [&] { return s->has_job || s->quit; });
    if (s->quit) return;
    s->has_job = false;
    lock.unlock();
    rtg_status st = RTG_OK;
    std::string err;
    if (s->rows > 0) {
      rtg_run* run = nullptr;
      st = rtg_model_run(s->model, s->inputs.data(), static_cast<int32_t>(s->inputs.size()), &run);
      if (st == RTG_OK) {
        int32_t n_out = 0;
        rtg_run_num_outputs(run, &n_out);
        if (n_out != static_cast<int32_t>(s->outputs.size())) st = set_error(RTG_ERR_INTERNAL, "output count changed");
        for (int32_t i = 0; i < n_out && st == RTG_OK; ++i) {
          int64_t shape[RTG_MAX_DIMS];
          int32_t nd = 0, dt = 0;
          st = rtg_run_output_info(run, i, shape, &nd, &dt);
          if (st != RTG_OK) break;
          if (nd < 1 || shape[0] != s->rows) {
            st = set_errorf(RTG_ERR_UNSUPPORTED_VALUE, "output %d does not carry the batch in dim 0 (%lld rows for a chunk of %lld)",
                            i, nd < 1 ? 0LL : (long long)shape[0], (long long)s->rows);
            break;
          }
          int64_t per_row = 1;
          for (int k = 1; k < nd; ++k) per_row *= shape[k];
          const rtg_tensor& dst = s->outputs[i];
          int64_t cap_per_row = 1;
          for (int k = 1; k < dst.ndim; ++k) cap_per_row *= dst.shape[k];
          if (dst.ndim != nd || cap_per_row != per_row) {
            st = set_errorf(RTG_ERR_INCORRECT_OUTPUT_TYPE, "output buffer %d has the wrong shape", i);
            break;
          }
          st = rtg_run_output_copy(run, i, dst.data, static_cast<size_t>(s->rows * per_row) * 4, RTG_HOST);
        }
        if (st != RTG_OK) err = rtg_last_error();
        rtg_run_free(run);
      } else {
        err = rtg_last_error();
      }
    }
    lock.lock();
    s->status = st;
    s->error = err;
    s->done = true;
    lock.unlock();
    s->cv.notify_all();
  }
}

}  // namespace

extern "C" {

rtg_status rtg_sharded_model_load(const int32_t* devices, int32_t n_devices, int32_t math, const void* data, size_t len,
                                  int32_t optimize, rtg_sharded_model** out) {
  if (!devices || n_devices < 1 || n_devices > 64 || !data || !out) return set_error(RTG_ERR_INVALID_VALUE, "bad argument");
  std::unique_ptr<Sharded> sh(new Sharded());
  auto cleanup = [&]() {
    for (auto& s : sh->shards) {
      if (s->model) rtg_model_free(s->model);
      if (s->ctx) rtg_ctx_destroy(s->ctx);
    }
  };
  for (int32_t i = 0; i < n_devices; ++i) {
    std::unique_ptr<Shard> s(new Shard());
    rtg_status st = rtg_ctx_create(devices[i], &s->ctx);
    if (st == RTG_OK) st = rtg_ctx_set_math(s->ctx, math);
    if (st == RTG_OK) st = rtg_model_load(s->ctx, data, len, optimize, &s->model);  // weights replicated per GPU
    sh->shards.push_back(std::move(s));
    if (st != RTG_OK) {
      std::string msg = rtg_last_error();
      cleanup();
      return set_errorf(st, "%s", msg.c_str());
    }
  }
  const Model* m0 = reinterpret_cast<const Model*>(sh->shards[0]->model);
  rtg_status st = check_batch_separable(m0->graph);
  if (st != RTG_OK) {
    std::string msg = rtg_last_error();
    cleanup();
    return set_errorf(st, "%s", msg.c_str());
  }
  sh->n_inputs = static_cast<int32_t>(m0->graph.input_ids.size());
  sh->n_outputs = static_cast<int32_t>(m0->graph.output_ids.size());
  for (auto& s : sh->shards) s->worker = std::thread(worker_loop, s.get());
  *out = reinterpret_cast<rtg_sharded_model*>(sh.release());
  return RTG_OK;
}

rtg_status rtg_sharded_model_free(rtg_sharded_model* model) {
  Sharded* sh = reinterpret_cast<Sharded*>(model);
  if (!sh) return RTG_OK;
  for (auto& s : sh->shards) {
    {
      std::lock_guard<std::mutex> lock(s->mu);
      s->quit = true;
    }
    s->cv.notify_all();
    if (s->worker.joinable()) s->worker.join();
    if (s->model) rtg_model_free(s->model);
    if (s->ctx) rtg_ctx_destroy(s->ctx);
  }
  delete sh;
  return RTG_OK;
}

rtg_status rtg_sharded_model_num_shards(const rtg_sharded_model* model, int32_t* n) {
  const Sharded* sh = reinterpret_cast<const Sharded*>(model);
  if (!sh || !n) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  *n = static_cast<int32_t>(sh->shards.size());
  return RTG_OK;
}

rtg_status rtg_sharded_model_run(rtg_sharded_model* model, const rtg_tensor* inputs, int32_t n_inputs, rtg_tensor* outputs,
                                 int32_t n_outputs) {
  Sharded* sh = reinterpret_cast<Sharded*>(model);
  if (!sh || (n_inputs > 0 && !inputs) || (n_outputs > 0 && !outputs)) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  if (n_inputs != sh->n_inputs)
    return set_errorf(RTG_ERR_MISSING_INPUTS, "model expects %d inputs, got %d", sh->n_inputs, n_inputs);
  if (n_outputs != sh->n_outputs)
    return set_errorf(RTG_ERR_INVALID_VALUE, "model has %d outputs, got %d buffers", sh->n_outputs, n_outputs);
  if (n_inputs == 0) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "a model without inputs has no batch to shard");
  int64_t batch = -1;
  for (int32_t i = 0; i < n_inputs; ++i) {
    const rtg_tensor& t = inputs[i];
    if (t.device != RTG_HOST || !is_contiguous(t) || t.ndim < 1)
      return set_error(RTG_ERR_UNSUPPORTED_VALUE, "rtg_sharded_model_run takes contiguous host inputs with a batch dimension");
    if (batch < 0) batch = t.shape[0];
    if (t.shape[0] != batch) return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "inputs disagree on the batch size");
  }
  for (int32_t i = 0; i < n_outputs; ++i)
    if (outputs[i].device != RTG_HOST || !outputs[i].data || !is_contiguous(outputs[i]) || outputs[i].ndim < 1 ||
        outputs[i].shape[0] != batch)
      return set_error(RTG_ERR_INCORRECT_OUTPUT_TYPE, "outputs must be contiguous host buffers with the batch in dim 0");
  const int64_t g = static_cast<int64_t>(sh->shards.size());
  const int64_t base = batch / g, extra = batch % g;
  int64_t lo = 0;
  for (int64_t r = 0; r < g; ++r) {  // the first `extra` shards take one more item (parallel.shard_range)
    Shard* s = sh->shards[r].get();
    const int64_t rows = base + (r < extra ? 1 : 0);
    std::lock_guard<std::mutex> lock(s->mu);
    s->rows = rows;
    s->inputs.assign(inputs, inputs + n_inputs);
    for (rtg_tensor& t : s->inputs) {
      int64_t per_row = 1;
      for (int k = 1; k < t.ndim; ++k) per_row *= t.shape[k];
      t.data = static_cast<char*>(t.data) + lo * per_row * 4;
      t.shape[0] = rows;
    }
    s->outputs.assign(outputs, outputs + n_outputs);
    for (rtg_tensor& t : s->outputs) {
      int64_t per_row = 1;
      for (int k = 1; k < t.ndim; ++k) per_row *= t.shape[k];
      t.data = static_cast<char*>(t.data) + lo * per_row * 4;
      t.shape[0] = rows;
    }
    s->done = false;
    s->has_job = true;
    lo += rows;
  }
  for (auto& s : sh->shards) s->cv.notify_all();
  rtg_status st = RTG_OK;
  std::string err;
  for (auto& s : sh->shards) {  // a failed shard fails the whole run
    std::unique_lock<std::mutex> lock(s->mu);
    s->cv.wait(lock, [&] { return s->done; });
    if (s->status != RTG_OK && st == RTG_OK) {
      st = s->status;
      err = s->error;
    }
  }
  if (st != RTG_OK) return set_errorf(st, "%s", err.c_str());
  return RTG_OK;
}

}  // extern "C"
