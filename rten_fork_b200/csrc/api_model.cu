// C ABI: Model::load / load_file / run and RunOptions timing (reference: src/model.rs:214-263,580-646,
// src/graph.rs:467-483,1053-1135, src/timing.rs).
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <map>
#include <memory>

#include "graph.h"

using namespace rtg;

void rtg::SubmitPipeline::release() {
  for (Slot& sl : slots) {
    for (void* p : sl.staging)
      if (p) cudaFree(p);
    if (sl.uploaded) cudaEventDestroy(sl.uploaded);
    if (sl.consumed) cudaEventDestroy(sl.consumed);
    sl = Slot();
  }
  if (upload) cudaStreamDestroy(upload);
  upload = nullptr;
}

void rtg::Model::release_captured() {
  for (auto& c : captured) {
    if (c->exec) {
      cudaGraphExecDestroy(c->exec);
      ctx->pool.pinned_addresses--;
    }
    for (Value& v : c->in_bufs) release_value(ctx, &v);
    for (Value& v : c->outs) release_value(ctx, &v);
  }
  captured.clear();
}

namespace {

// RTEN_TIMING (src/model.rs:124-147, :587-590): when set, every Model::run records per-operator times and prints the
// table; the value is a space-separated list of key=value options.  sort=name | time is honoured; by-shape is accepted
// but shapes are not recorded here.
struct TimingEnv {
  bool on = false, by_name = false;
};
const TimingEnv& timing_env() {
  static const TimingEnv env = []() {
    TimingEnv t;
    const char* e = getenv("RTEN_TIMING");
    if (!e) return t;
    t.on = true;
    std::string cfg(e);
    size_t pos = 0;
    while (pos < cfg.size()) {
      size_t end = cfg.find_first_of(" \t\n", pos);
      if (end == std::string::npos) end = cfg.size();
      const std::string tok = cfg.substr(pos, end - pos);
      pos = end + 1;
      const size_t eq = tok.find('=');
      if (tok.empty() || eq == std::string::npos) continue;
      const std::string key = tok.substr(0, eq), val = tok.substr(eq + 1);
      if (key == "sort") {
        if (val == "name") t.by_name = true;
        else if (val == "time") t.by_name = false;
        else fprintf(stderr, "Unrecognized sort order \"%s\"\n", val.c_str());
      } else if (key == "by-shape") {
        // accepted: the table is per operator type either way
      } else {
        fprintf(stderr, "Unrecognized timing option \"%s\"\n", key.c_str());
      }
    }
    return t;
  }();
  return env;
}

bool graph_capture_default() {
  static const bool on = []() {
    const char* e = getenv("RTEN_GPU_GRAPH");
    return !e || atoi(e) != 0;
  }();
  return on;
}

// Copies one caller input (host or device, any strides for device tensors) into a device value of the same shape.
rtg_status copy_input(Context* ctx, const rtg_tensor& t, const Value& dst) {
  const int64_t n = numel(t);
  if (n == 0) return RTG_OK;
  if (t.device == RTG_HOST) {
    rtg_tensor tc = t;
    if (!is_contiguous(tc)) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "host model inputs must be contiguous");
    RTG_CUDA(cudaMemcpyAsync(dst.t.data, t.data, static_cast<size_t>(n) * 4, cudaMemcpyHostToDevice, ctx->stream));
    return RTG_OK;
  }
  return copy_view(ctx, dst.t, t);
}

// The outputs of a replay live in buffers the next replay overwrites: the run handle gets its own copies (same
// strides, so an NHWC-strided result stays one).
rtg_status copy_outputs(Context* ctx, const CapturedPlan& cp, Run* run) {
  for (const Value& src : cp.outs) {
    Value dst;
    RTG_TRY(alloc_value(ctx, src.t.dtype, src.t.ndim, src.t.shape, false, &dst));
    for (int k = 0; k < src.t.ndim; ++k) dst.t.strides[k] = src.t.strides[k];
    const int64_t n = numel(src.t);
    if (n > 0) {
      cudaError_t e = cudaMemcpyAsync(dst.t.data, src.t.data, static_cast<size_t>(n) * 4, cudaMemcpyDeviceToDevice, ctx->stream);
      if (e != cudaSuccess) {
        release_value(ctx, &dst);
        return set_errorf(RTG_ERR_CUDA, "output copy failed: %s", cudaGetErrorString(e));
      }
    }
    dst.host_i32 = src.host_i32, dst.host_f32 = src.host_f32;
    run->outputs.push_back(dst);
  }
  return RTG_OK;
}

// Replays (or first captures) the plan for this key.  Returns RTG_OK with *handled = false when the caller should run
// eagerly (first run of a key, capture not possible, timing on).
rtg_status run_captured(Model* m, const rtg_tensor* inputs, int32_t n_inputs, const std::vector<int32_t>& in_ids,
                        const std::vector<int32_t>& outs_req, Run* run, bool* handled) {
  *handled = false;
  Context* ctx = m->ctx;
  if (!m->graph_capture || !graph_capture_default() || m->timing || timing_env().on || ctx->ktiming || !ctx->pool.use_pool)
    return RTG_OK;
  std::lock_guard<std::mutex> lock(m->cap_mu);
  CapturedPlan* cp = nullptr;
  for (auto& c : m->captured) {
    if (c->math != ctx->math || c->in_ids != in_ids || c->out_ids != outs_req) continue;
    bool same = true;
    for (int32_t i = 0; i < n_inputs && same; ++i) {
      same = c->in_desc[i].dtype == inputs[i].dtype && c->in_desc[i].ndim == inputs[i].ndim;
      for (int k = 0; k < inputs[i].ndim && same; ++k) same = c->in_desc[i].shape[k] == inputs[i].shape[k];
    }
    if (same) {
      cp = c.get();
      break;
    }
  }
  if (!cp) {
    if (m->captured.size() >= 8) return RTG_OK;  // a model run at many shapes: keep the eager path for the rest
    m->captured.emplace_back(new CapturedPlan());
    cp = m->captured.back().get();
    cp->in_ids = in_ids, cp->out_ids = outs_req, cp->math = ctx->math;
    cp->in_desc.assign(inputs, inputs + n_inputs);
  }
  if (cp->failed) return RTG_OK;
  if (!cp->exec) {
    if (cp->seen++ < 1) return RTG_OK;  // first run: eager
    // ---- capture
    rtg_status st = RTG_OK;
    for (int32_t i = 0; i < n_inputs && st == RTG_OK; ++i) {
      Value v;
      st = alloc_value(ctx, inputs[i].dtype, inputs[i].ndim, inputs[i].shape, false, &v);
      if (st == RTG_OK) cp->in_bufs.push_back(v);
    }
    for (int32_t i = 0; i < n_inputs && st == RTG_OK; ++i) st = copy_input(ctx, inputs[i], cp->in_bufs[i]);
    auto give_up = [&](rtg_status why) {
      for (Value& v : cp->in_bufs) release_value(ctx, &v);
      for (Value& v : cp->outs) release_value(ctx, &v);
      cp->in_bufs.clear();
      cp->outs.clear();
      cp->failed = true;
      cudaGetLastError();
      (void)why;
      return RTG_OK;  // the caller runs eagerly
    };
    if (st != RTG_OK) return give_up(st);
    if (cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed) != cudaSuccess) return give_up(RTG_ERR_CUDA);
    std::vector<std::pair<int32_t, Value>> in;
    for (int32_t i = 0; i < n_inputs; ++i) {
      Value v = cp->in_bufs[i];
      v.owned = false;  // the captured kernels read the persistent buffers; nothing may recycle or overwrite them
      in.emplace_back(in_ids[i], v);
    }
    const uint64_t l0 = ctx->launches;
    std::vector<Value> outs;
    st = m->graph.run(ctx, in, outs_req, &outs, nullptr);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(ctx->stream, &graph);
    if (st != RTG_OK || ce != cudaSuccess || !graph) {
      // an operator needed a host round trip (or failed): not capturable; nothing was executed
      if (graph) cudaGraphDestroy(graph);
      for (Value& v : outs) release_value(ctx, &v);
      ctx->launches = l0;
      return give_up(st);
    }
    cp->launches = ctx->launches - l0;
    ctx->launches = l0;
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ie = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ie != cudaSuccess) {
      for (Value& v : outs) release_value(ctx, &v);
      return give_up(RTG_ERR_CUDA);
    }
    // outputs that are views of something else (constants, requested twice) were copied by Graph::run: all owned
    cp->outs = outs;
    cp->exec = exec;
    ctx->pool.pinned_addresses++;
  } else {
    for (int32_t i = 0; i < n_inputs; ++i) RTG_TRY(copy_input(ctx, inputs[i], cp->in_bufs[i]));
  }
  RTG_CUDA(cudaGraphLaunch(cp->exec, ctx->stream));
  ctx->launches += cp->launches;
  m->replays++;
  RTG_TRY(copy_outputs(ctx, *cp, run));
  run->node_ids = outs_req;
  *handled = true;
  return RTG_OK;
}

rtg_status node_info(const Model* m, int32_t id, const char** name, int64_t* shape, int32_t* ndim, int32_t* dtype) {
  const Node& nd = m->graph.nodes[id];
  if (name) *name = nd.name.c_str();
  if (ndim) *ndim = nd.has_shape ? static_cast<int32_t>(nd.shape.size()) : -1;
  if (shape && nd.has_shape)
    for (size_t i = 0; i < nd.shape.size() && i < RTG_MAX_DIMS; ++i) shape[i] = nd.shape[i].value;
  if (dtype) *dtype = -1;  // ValueNodes carry no dtype in schema version 1
  return RTG_OK;
}

// Per-operator timing table, sorted by time, in the layout of the reference's report (src/timing.rs:146-385).
std::string format_timing(const std::vector<TimingRecord>& recs, const DevPool& pool, bool by_name = false) {
  std::map<std::string, std::pair<double, int>> by_op;
  double total = 0;
  for (const TimingRecord& r : recs) {
    by_op[r.op].first += r.ms;
    by_op[r.op].second++;
    total += r.ms;
  }
  std::vector<std::pair<std::string, std::pair<double, int>>> rows(by_op.begin(), by_op.end());
  if (by_name) std::sort(rows.begin(), rows.end(), [](auto& a, auto& b) { return a.first < b.first; });
  else std::sort(rows.begin(), rows.end(), [](auto& a, auto& b) { return a.second.first > b.second.first; });
  char line[256];
  snprintf(line, sizeof(line), "Graph run of %zu ops finished in %.3fms (device time)\n\n", recs.size(), total);
  std::string s = line;
  for (auto& r : rows) {
    snprintf(line, sizeof(line), "%-22s %9.3fms (%5.2f%%)  x%d\n", r.first.c_str(), r.second.first,
             total > 0 ? 100.0 * r.second.first / total : 0.0, r.second.second);
    s += line;
  }
  snprintf(line, sizeof(line), "\nPool allocs %llu hits %llu\n", (unsigned long long)pool.alloc_count,
           (unsigned long long)pool.hit_count);
  s += line;
  return s;
}

// Model::run proper.  `host_mirror` (rtg_model_submit): the caller's host copies of `inputs`, which then point at
// device staging buffers; small i32 inputs keep a host mirror for shape arithmetic either way.
// `input_ids` / `output_ids`: node ids (Model::run addresses values by id, src/model.rs:580-592); NULL = the graph's
// own input / output lists in order.  `partial`: Model::partial_run (src/model.rs:639-646).
rtg_status run_inputs(Model* m, const rtg_tensor* inputs, int32_t n_inputs, const rtg_tensor* host_mirror, Run** out,
                      const int32_t* input_ids = nullptr, const int32_t* output_ids = nullptr, int32_t n_outputs = 0,
                      bool partial = false) {
  Context* ctx = m->ctx;
  Graph& g = m->graph;
  if (!input_ids && n_inputs != static_cast<int32_t>(g.input_ids.size()))
    return set_errorf(RTG_ERR_MISSING_INPUTS, "model expects %zu inputs, got %d", g.input_ids.size(), n_inputs);
  const int32_t n_nodes = static_cast<int32_t>(g.nodes.size());
  for (int32_t i = 0; input_ids && i < n_inputs; ++i) {
    if (input_ids[i] < 0 || input_ids[i] >= n_nodes) return set_error(RTG_ERR_INVALID_VALUE, "Input node ID is invalid");
    for (int32_t k = 0; k < i; ++k)
      if (input_ids[k] == input_ids[i]) return set_error(RTG_ERR_INVALID_VALUE, "Inputs are not unique");  // RunError::PlanningError
  }
  std::vector<int32_t> outs_req = output_ids ? std::vector<int32_t>(output_ids, output_ids + n_outputs) : g.output_ids;
  for (size_t i = 0; i < outs_req.size(); ++i) {
    if (outs_req[i] < 0 || outs_req[i] >= n_nodes) return set_error(RTG_ERR_INVALID_VALUE, "Output node ID is invalid");
    for (size_t k = 0; k < i; ++k)
      if (outs_req[k] == outs_req[i]) return set_error(RTG_ERR_INVALID_VALUE, "Outputs are not unique");
  }
  for (int32_t i = 0; i < n_inputs; ++i)
    if (inputs[i].ndim < 0 || inputs[i].ndim > RTG_MAX_DIMS || (inputs[i].dtype != RTG_F32 && inputs[i].dtype != RTG_I32))
      return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  if (!partial) {
    std::vector<int32_t> in_ids_v(n_inputs);
    for (int32_t i = 0; i < n_inputs; ++i) in_ids_v[i] = input_ids ? input_ids[i] : g.input_ids[i];
    std::unique_ptr<Run> crun(new Run());
    crun->ctx = ctx;
    bool handled = false;
    rtg_status cst = run_captured(m, inputs, n_inputs, in_ids_v, outs_req, crun.get(), &handled);
    if (cst != RTG_OK) {
      for (Value& v : crun->outputs) release_value(ctx, &v);
      return cst;
    }
    if (handled) {
      *out = crun.release();
      return RTG_OK;
    }
  }
  std::vector<std::pair<int32_t, Value>> in;
  auto cleanup = [&]() {
    for (auto& kv : in) release_value(ctx, &kv.second);
  };
  for (int32_t i = 0; i < n_inputs; ++i) {
    const rtg_tensor& t = inputs[i];
    if (t.ndim < 0 || t.ndim > RTG_MAX_DIMS || (t.dtype != RTG_F32 && t.dtype != RTG_I32)) {
      cleanup();
      return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
    }
    // Inputs become owned temporaries (moved into the run like the reference's owned inputs, src/graph.rs:808-824)
    Value v;
    rtg_status st = alloc_value(ctx, t.dtype, t.ndim, t.shape, false, &v);
    if (st != RTG_OK) {
      cleanup();
      return st;
    }
    const int64_t n = numel(t);
    if (n > 0) {
      if (t.device == RTG_HOST) {
        rtg_tensor tc = t;
        if (!is_contiguous(tc)) {
          release_value(ctx, &v);
          cleanup();
          return set_error(RTG_ERR_UNSUPPORTED_VALUE, "host model inputs must be contiguous");
        }
        cudaError_t e = cudaMemcpyAsync(v.t.data, t.data, static_cast<size_t>(n) * 4, cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) {
          release_value(ctx, &v);
          cleanup();
          return set_errorf(RTG_ERR_CUDA, "input upload failed: %s", cudaGetErrorString(e));
        }
        if (t.dtype == RTG_I32 && n <= 4096)
          v.host_i32 = std::make_shared<std::vector<int32_t>>(static_cast<const int32_t*>(t.data), static_cast<const int32_t*>(t.data) + n);
      } else {
        if (host_mirror && t.dtype == RTG_I32 && n <= 4096) {
          const int32_t* h = static_cast<const int32_t*>(host_mirror[i].data);
          v.host_i32 = std::make_shared<std::vector<int32_t>>(h, h + n);
        }
        st = copy_view(ctx, v.t, t);
        if (st != RTG_OK) {
          release_value(ctx, &v);
          cleanup();
          return st;
        }
      }
    }
    in.emplace_back(input_ids ? input_ids[i] : g.input_ids[i], v);
  }
  std::unique_ptr<Run> run(new Run());
  run->ctx = ctx;
  std::vector<TimingRecord> recs;
  if (partial) {
    std::vector<std::pair<int32_t, Value>> leaves;
    rtg_status st = g.partial_run(ctx, in, outs_req, &leaves);
    if (st != RTG_OK) return st;
    for (auto& kv : leaves) {
      run->node_ids.push_back(kv.first);
      run->outputs.push_back(kv.second);
    }
    *out = run.release();
    return RTG_OK;
  }
  const bool timing = m->timing || timing_env().on;
  rtg_status st = g.run(ctx, in, outs_req, &run->outputs, timing ? &recs : nullptr);
  if (st != RTG_OK) return st;  // Graph::run released the inputs
  run->node_ids = outs_req;
  if (timing) m->timing_report = format_timing(recs, ctx->pool, timing_env().by_name);
  if (timing_env().on) fputs(m->timing_report.c_str(), stdout);  // the reference prints the table after the run
  *out = run.release();
  return RTG_OK;
}

}  // namespace

extern "C" {

rtg_status rtg_model_load(rtg_ctx* ctx_, const void* data, size_t len, int32_t optimize, rtg_model** out) {
  Context* ctx = reinterpret_cast<Context*>(ctx_);
  if (!ctx || !data || !out) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_CUDA(cudaSetDevice(ctx->device));
  Model* m = nullptr;
  RTG_TRY(load_model(ctx, static_cast<const uint8_t*>(data), len, optimize != 0, &m));
  *out = reinterpret_cast<rtg_model*>(m);
  return RTG_OK;
}

rtg_status rtg_model_load_file(rtg_ctx* ctx_, const char* path, int32_t optimize, rtg_model** out) {
  if (!path) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  FILE* f = fopen(path, "rb");
  if (!f) return set_errorf(RTG_ERR_FORMAT, "read error: cannot open %s", path);  // ModelLoadError::ReadFailed
  fseek(f, 0, SEEK_END);
  const long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  std::vector<uint8_t> buf(static_cast<size_t>(std::max<long>(n, 0)));
  const size_t got = n > 0 ? fread(buf.data(), 1, buf.size(), f) : 0;
  fclose(f);
  if (got != buf.size()) return set_errorf(RTG_ERR_FORMAT, "read error: short read on %s", path);
  return rtg_model_load(ctx_, buf.data(), buf.size(), optimize, out);
}

rtg_status rtg_model_free(rtg_model* model) {
  Model* m = reinterpret_cast<Model*>(model);
  if (!m) return RTG_OK;
  cudaSetDevice(m->ctx->device);
  cudaStreamSynchronize(m->ctx->stream);
  m->release_captured();
  m->graph.release(m->ctx);
  m->pipe.release();
  delete m;
  return RTG_OK;
}

rtg_status rtg_model_num_inputs(const rtg_model* model, int32_t* n) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || !n) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  *n = static_cast<int32_t>(m->graph.input_ids.size());
  return RTG_OK;
}
rtg_status rtg_model_num_outputs(const rtg_model* model, int32_t* n) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || !n) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  *n = static_cast<int32_t>(m->graph.output_ids.size());
  return RTG_OK;
}
rtg_status rtg_model_num_ops(const rtg_model* model, int32_t* n) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || !n) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  *n = m->graph.num_operators();
  return RTG_OK;
}
rtg_status rtg_model_input_info(const rtg_model* model, int32_t i, const char** name, int64_t* shape, int32_t* ndim,
                                int32_t* dtype) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || i < 0 || i >= static_cast<int32_t>(m->graph.input_ids.size()))
    return set_error(RTG_ERR_INVALID_VALUE, "input index out of range");
  return node_info(m, m->graph.input_ids[i], name, shape, ndim, dtype);
}
rtg_status rtg_model_output_info(const rtg_model* model, int32_t i, const char** name, int64_t* shape, int32_t* ndim,
                                 int32_t* dtype) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || i < 0 || i >= static_cast<int32_t>(m->graph.output_ids.size()))
    return set_error(RTG_ERR_INVALID_VALUE, "output index out of range");
  return node_info(m, m->graph.output_ids[i], name, shape, ndim, dtype);
}

rtg_status rtg_model_set_graph_capture(rtg_model* model, int32_t enabled) {
  Model* m = reinterpret_cast<Model*>(model);
  if (!m) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  std::lock_guard<std::mutex> lock(m->cap_mu);
  m->graph_capture = enabled != 0;
  return RTG_OK;
}
rtg_status rtg_model_graph_replays(rtg_model* model, uint64_t* replays) {
  Model* m = reinterpret_cast<Model*>(model);
  if (!m || !replays) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  *replays = m->replays;
  return RTG_OK;
}

rtg_status rtg_model_set_timing(rtg_model* model, int32_t enabled) {
  Model* m = reinterpret_cast<Model*>(model);
  if (!m) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  m->timing = enabled != 0;
  return RTG_OK;
}
const char* rtg_model_timing_report(rtg_model* model) {
  Model* m = reinterpret_cast<Model*>(model);
  return m ? m->timing_report.c_str() : "";
}

rtg_status rtg_model_run(rtg_model* model, const rtg_tensor* inputs, int32_t n_inputs, rtg_run** out) {
  Model* m = reinterpret_cast<Model*>(model);
  if (!m || !out || (n_inputs > 0 && !inputs)) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_CUDA(cudaSetDevice(m->ctx->device));
  Run* run = nullptr;
  RTG_TRY(run_inputs(m, inputs, n_inputs, nullptr, &run));
  *out = reinterpret_cast<rtg_run*>(run);
  return RTG_OK;
}

rtg_status rtg_model_run_ids(rtg_model* model, const int32_t* input_ids, const rtg_tensor* inputs, int32_t n_inputs,
                             const int32_t* output_ids, int32_t n_outputs, rtg_run** out) {
  Model* m = reinterpret_cast<Model*>(model);
  if (!m || !out || (n_inputs > 0 && (!inputs || !input_ids)) || (n_outputs > 0 && !output_ids))
    return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_CUDA(cudaSetDevice(m->ctx->device));
  static const int32_t none = 0;
  Run* run = nullptr;
  RTG_TRY(run_inputs(m, inputs, n_inputs, nullptr, &run, input_ids ? input_ids : &none, output_ids ? output_ids : &none,
                     n_outputs, false));
  *out = reinterpret_cast<rtg_run*>(run);
  return RTG_OK;
}

rtg_status rtg_model_partial_run(rtg_model* model, const int32_t* input_ids, const rtg_tensor* inputs, int32_t n_inputs,
                                 const int32_t* output_ids, int32_t n_outputs, rtg_run** out) {
  Model* m = reinterpret_cast<Model*>(model);
  if (!m || !out || (n_inputs > 0 && (!inputs || !input_ids)) || (n_outputs > 0 && !output_ids))
    return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_CUDA(cudaSetDevice(m->ctx->device));
  static const int32_t none = 0;
  Run* run = nullptr;
  RTG_TRY(run_inputs(m, inputs, n_inputs, nullptr, &run, input_ids ? input_ids : &none, output_ids ? output_ids : &none,
                     n_outputs, true));
  *out = reinterpret_cast<rtg_run*>(run);
  return RTG_OK;
}

rtg_status rtg_run_output_node(const rtg_run* run, int32_t i, int32_t* node_id) {
  const Run* r = reinterpret_cast<const Run*>(run);
  if (!r || !node_id || i < 0 || i >= static_cast<int32_t>(r->node_ids.size()))
    return set_error(RTG_ERR_INVALID_VALUE, "output index out of range");
  *node_id = r->node_ids[i];
  return RTG_OK;
}

rtg_status rtg_model_find_node(const rtg_model* model, const char* name, int32_t* id) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || !name || !id) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  for (size_t i = 0; i < m->graph.nodes.size(); ++i)
    if (m->graph.nodes[i].name == name) {
      *id = static_cast<int32_t>(i);
      return RTG_OK;
    }
  return set_errorf(RTG_ERR_INVALID_VALUE, "node \"%s\" not found", name);
}

rtg_status rtg_model_input_id(const rtg_model* model, int32_t i, int32_t* id) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || !id || i < 0 || i >= static_cast<int32_t>(m->graph.input_ids.size()))
    return set_error(RTG_ERR_INVALID_VALUE, "input index out of range");
  *id = m->graph.input_ids[i];
  return RTG_OK;
}

rtg_status rtg_model_output_id(const rtg_model* model, int32_t i, int32_t* id) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || !id || i < 0 || i >= static_cast<int32_t>(m->graph.output_ids.size()))
    return set_error(RTG_ERR_INVALID_VALUE, "output index out of range");
  *id = m->graph.output_ids[i];
  return RTG_OK;
}

rtg_status rtg_model_node_kind(const rtg_model* model, int32_t id, int32_t* kind) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || !kind || id < 0 || id >= static_cast<int32_t>(m->graph.nodes.size()))
    return set_error(RTG_ERR_INVALID_VALUE, "node id out of range");
  *kind = static_cast<int32_t>(m->graph.nodes[id].kind);
  return RTG_OK;
}

rtg_status rtg_model_source_op(const rtg_model* model, int32_t value_id, const char** op_name, const char** node_name) {
  const Model* m = reinterpret_cast<const Model*>(model);
  if (!m || value_id < 0 || value_id >= static_cast<int32_t>(m->graph.nodes.size()))
    return set_error(RTG_ERR_INVALID_VALUE, "node id out of range");
  const int32_t src = m->graph.source_node(value_id);
  if (src < 0) return set_error(RTG_ERR_INVALID_VALUE, "value has no source operator");
  if (op_name) *op_name = m->graph.nodes[src].op->name();
  if (node_name) *node_name = m->graph.nodes[src].name.c_str();
  return RTG_OK;
}

/* Asynchronous Model::run for host buffers (pinned memory for real overlap). */
rtg_status rtg_model_submit(rtg_model* model, const rtg_tensor* inputs, int32_t n_inputs, rtg_tensor* outputs,
                            int32_t n_outputs, rtg_ticket** out) {
  Model* m = reinterpret_cast<Model*>(model);
  if (!m || !out || (n_inputs > 0 && !inputs) || (n_outputs > 0 && !outputs))
    return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  Context* ctx = m->ctx;
  RTG_CUDA(cudaSetDevice(ctx->device));
  if (n_inputs != static_cast<int32_t>(m->graph.input_ids.size()))
    return set_errorf(RTG_ERR_MISSING_INPUTS, "model expects %zu inputs, got %d", m->graph.input_ids.size(), n_inputs);
  if (n_outputs != static_cast<int32_t>(m->graph.output_ids.size()))
    return set_errorf(RTG_ERR_INVALID_VALUE, "model has %zu outputs, got %d buffers", m->graph.output_ids.size(), n_outputs);
  for (int32_t i = 0; i < n_inputs; ++i) {
    rtg_tensor t = inputs[i];
    if (t.ndim < 0 || t.ndim > RTG_MAX_DIMS || (t.dtype != RTG_F32 && t.dtype != RTG_I32))
      return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
    if (t.device != RTG_HOST || !is_contiguous(t))
      return set_error(RTG_ERR_UNSUPPORTED_VALUE, "rtg_model_submit takes contiguous host inputs");
  }
  for (int32_t i = 0; i < n_outputs; ++i)
    if (outputs[i].device != RTG_HOST || !outputs[i].data)
      return set_error(RTG_ERR_UNSUPPORTED_VALUE, "rtg_model_submit writes to host output buffers");
  SubmitPipeline& pp = m->pipe;
  if (!pp.upload) RTG_CUDA(cudaStreamCreateWithFlags(&pp.upload, cudaStreamNonBlocking));
  SubmitPipeline::Slot& sl = pp.slots[pp.next++ & 1];
  if (!sl.uploaded) {
    RTG_CUDA(cudaEventCreateWithFlags(&sl.uploaded, cudaEventDisableTiming));
    RTG_CUDA(cudaEventCreateWithFlags(&sl.consumed, cudaEventDisableTiming));
  }
  sl.staging.resize(n_inputs, nullptr);
  sl.cap.resize(n_inputs, 0);
  // the slot's previous contents have been copied into that run's input tensors
  if (sl.in_use) RTG_CUDA(cudaStreamWaitEvent(pp.upload, sl.consumed, 0));
  std::vector<rtg_tensor> dev(inputs, inputs + n_inputs);
  for (int32_t i = 0; i < n_inputs; ++i) {
    const size_t bytes = static_cast<size_t>(numel(inputs[i])) * 4;
    if (bytes > sl.cap[i]) {
      if (sl.staging[i]) {
        RTG_CUDA(cudaStreamSynchronize(ctx->stream));
        RTG_CUDA(cudaFree(sl.staging[i]));
        sl.staging[i] = nullptr, sl.cap[i] = 0;
      }
      RTG_CUDA(cudaMalloc(&sl.staging[i], bytes));
      sl.cap[i] = bytes;
    }
    if (bytes) RTG_CUDA(cudaMemcpyAsync(sl.staging[i], inputs[i].data, bytes, cudaMemcpyHostToDevice, pp.upload));
    dev[i].data = sl.staging[i];
    dev[i].device = ctx->device;
  }
  RTG_CUDA(cudaEventRecord(sl.uploaded, pp.upload));
  RTG_CUDA(cudaStreamWaitEvent(ctx->stream, sl.uploaded, 0));
  Run* run = nullptr;
  rtg_status st = run_inputs(m, dev.data(), n_inputs, inputs, &run);
  // whatever happened, the staging reads that were enqueued are ordered before this point of the compute stream
  RTG_CUDA(cudaEventRecord(sl.consumed, ctx->stream));
  sl.in_use = true;
  if (st != RTG_OK) return st;
  std::unique_ptr<Run> guard(run);
  auto drop = [&]() {
    for (Value& v : run->outputs) release_value(ctx, &v);
  };
  for (int32_t i = 0; i < n_outputs; ++i) {
    Value& v = run->outputs[i];
    const size_t bytes = static_cast<size_t>(numel(v.t)) * 4;
    const size_t have = static_cast<size_t>(numel(outputs[i])) * 4;
    if (have < bytes) {
      drop();
      return set_errorf(RTG_ERR_INVALID_VALUE, "output buffer %d too small: %zu bytes for %zu", i, have, bytes);
    }
    outputs[i].ndim = v.t.ndim;
    outputs[i].dtype = v.t.dtype;
    for (int k = 0; k < v.t.ndim; ++k) outputs[i].shape[k] = v.t.shape[k];
    set_contiguous_strides(outputs[i]);
    if (bytes == 0) continue;
    Scratch tmp;
    const void* src = v.t.data;
    if (!is_contiguous(v.t)) {
      rtg_tensor c = v.t;
      st = tmp.alloc(ctx, bytes);
      if (st == RTG_OK) {
        c.data = tmp.p;
        set_contiguous_strides(c);
        st = copy_view(ctx, c, v.t);
      }
      if (st != RTG_OK) {
        drop();
        return st;
      }
      src = tmp.p;
    }
    cudaError_t e = cudaMemcpyAsync(outputs[i].data, src, bytes, cudaMemcpyDeviceToHost, ctx->stream);
    if (e != cudaSuccess) {
      drop();
      return set_errorf(RTG_ERR_CUDA, "output download failed: %s", cudaGetErrorString(e));
    }
  }
  drop();  // stream-ordered: the pool may hand these buffers to later work on the same stream
  Ticket* tk = new Ticket();
  tk->ctx = ctx;
  cudaError_t e = cudaEventCreateWithFlags(&tk->done, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventRecord(tk->done, ctx->stream);
  if (e != cudaSuccess) {
    delete tk;
    return set_errorf(RTG_ERR_CUDA, "ticket event failed: %s", cudaGetErrorString(e));
  }
  *out = reinterpret_cast<rtg_ticket*>(tk);
  return RTG_OK;
}

rtg_status rtg_ticket_wait(rtg_ticket* ticket) {
  Ticket* tk = reinterpret_cast<Ticket*>(ticket);
  if (!tk) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_CUDA(cudaSetDevice(tk->ctx->device));
  cudaError_t e = cudaEventSynchronize(tk->done);
  cudaEventDestroy(tk->done);
  delete tk;
  if (e != cudaSuccess) return set_errorf(RTG_ERR_CUDA, "run failed: %s", cudaGetErrorString(e));
  return RTG_OK;
}

rtg_status rtg_run_num_outputs(const rtg_run* run, int32_t* n) {
  const Run* r = reinterpret_cast<const Run*>(run);
  if (!r || !n) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  *n = static_cast<int32_t>(r->outputs.size());
  return RTG_OK;
}

rtg_status rtg_run_output_info(const rtg_run* run, int32_t i, int64_t* shape, int32_t* ndim, int32_t* dtype) {
  const Run* r = reinterpret_cast<const Run*>(run);
  if (!r || i < 0 || i >= static_cast<int32_t>(r->outputs.size()))
    return set_error(RTG_ERR_INVALID_VALUE, "output index out of range");
  const rtg_tensor& t = r->outputs[i].t;
  if (ndim) *ndim = t.ndim;
  if (dtype) *dtype = t.dtype;
  if (shape)
    for (int k = 0; k < t.ndim; ++k) shape[k] = t.shape[k];
  return RTG_OK;
}

rtg_status rtg_run_output_copy(rtg_run* run, int32_t i, void* dst, size_t dst_bytes, int32_t device) {
  Run* r = reinterpret_cast<Run*>(run);
  if (!r || i < 0 || i >= static_cast<int32_t>(r->outputs.size()))
    return set_error(RTG_ERR_INVALID_VALUE, "output index out of range");
  Context* ctx = r->ctx;
  RTG_CUDA(cudaSetDevice(ctx->device));
  Value& v = r->outputs[i];
  const size_t bytes = static_cast<size_t>(numel(v.t)) * 4;
  if (dst_bytes < bytes) return set_error(RTG_ERR_INVALID_VALUE, "destination buffer too small");
  if (bytes == 0) return RTG_OK;
  // results leave in the reference layout (contiguous, logical order)
  Scratch tmp;
  const void* src = v.t.data;
  if (!is_contiguous(v.t)) {
    RTG_TRY(tmp.alloc(ctx, bytes));
    rtg_tensor c = v.t;
    c.data = tmp.p;
    set_contiguous_strides(c);
    RTG_TRY(copy_view(ctx, c, v.t));
    src = tmp.p;
  }
  RTG_CUDA(cudaMemcpyAsync(dst, src, bytes, device == RTG_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice,
                           ctx->stream));
  RTG_CUDA(cudaStreamSynchronize(ctx->stream));
  return RTG_OK;
}

rtg_status rtg_run_free(rtg_run* run) {
  Run* r = reinterpret_cast<Run*>(run);
  if (!r) return RTG_OK;
  for (Value& v : r->outputs) release_value(r->ctx, &v);
  delete r;
  return RTG_OK;
}

}  // extern "C"
