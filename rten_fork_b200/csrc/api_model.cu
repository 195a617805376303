// C ABI: Model::load / load_file / run and RunOptions timing (reference: src/model.rs:214-263,580-646,
// src/graph.rs:467-483,1053-1135, src/timing.rs).
#include <stdio.h>

#include <algorithm>
#include <map>

#include "graph.h"

using namespace rtg;

namespace {

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
std::string format_timing(const std::vector<TimingRecord>& recs, const DevPool& pool) {
  std::map<std::string, std::pair<double, int>> by_op;
  double total = 0;
  for (const TimingRecord& r : recs) {
    by_op[r.op].first += r.ms;
    by_op[r.op].second++;
    total += r.ms;
  }
  std::vector<std::pair<std::string, std::pair<double, int>>> rows(by_op.begin(), by_op.end());
  std::sort(rows.begin(), rows.end(), [](auto& a, auto& b) { return a.second.first > b.second.first; });
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
  m->graph.release(m->ctx);
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
  Context* ctx = m->ctx;
  RTG_CUDA(cudaSetDevice(ctx->device));
  Graph& g = m->graph;
  if (n_inputs != static_cast<int32_t>(g.input_ids.size()))
    return set_errorf(RTG_ERR_MISSING_INPUTS, "model expects %zu inputs, got %d", g.input_ids.size(), n_inputs);
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
        st = copy_view(ctx, v.t, t);
        if (st != RTG_OK) {
          release_value(ctx, &v);
          cleanup();
          return st;
        }
      }
    }
    in.emplace_back(g.input_ids[i], v);
  }
  std::unique_ptr<Run> run(new Run());
  run->ctx = ctx;
  std::vector<TimingRecord> recs;
  rtg_status st = g.run(ctx, in, g.output_ids, &run->outputs, m->timing ? &recs : nullptr);
  if (st != RTG_OK) return st;  // Graph::run released the inputs
  if (m->timing) m->timing_report = format_timing(recs, ctx->pool);
  *out = reinterpret_cast<rtg_run*>(run.release());
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
