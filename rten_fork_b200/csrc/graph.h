// Host runtime mirroring the reference surface for the hot path:
//   Operator   src/ops/mod.rs:821-913        OpRegistry  src/op_registry.rs:31-62
//   Graph      src/graph.rs:430-1346         Model       src/model.rs:214-646
// Tensors live in device memory owned by the context's TensorPool; operators borrow input views and return owned
// outputs, exactly as in the reference.  New here: 4-D float activations may be NHWC-strided views of their logical
// NCHW shape (what the TMA im2col convolution produces), and a load-time pass fuses Conv/MatMul epilogues.
#pragma once
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "flatbuf.h"
#include "ops.h"

namespace rtg {

// ---------------------------------------------------------------- values
// A tensor flowing through the graph: a device view plus, for constants and shape-like results, a host copy that
// shape-consuming operators (Reshape, Slice, ...) read without a device round trip.
struct Value {
  rtg_tensor t = {};                         // device view (t.data may be null for an empty tensor)
  bool owned = false;                        // t.data came from the context pool and is returned when dead
  std::shared_ptr<std::vector<int32_t>> host_i32;  // host mirror of small int tensors
  std::shared_ptr<std::vector<float>> host_f32;    // host mirror of small float tensors (scalars for Clip etc.)
};

struct OpInputs {
  std::vector<const Value*> v;  // nullptr = missing optional input
  const Value* get(size_t i) const { return i < v.size() ? v[i] : nullptr; }
  rtg_status require(size_t i, const Value** out) const;
};

// Allocate an owned output of the given shape (contiguous) or NHWC-strided when `nhwc` (4-D only).
rtg_status alloc_value(Context* ctx, int dtype, int ndim, const int64_t* shape, bool nhwc, Value* out);
void release_value(Context* ctx, Value* v);
// Read a small int / float tensor on the host (uses the host mirror when present, else a synchronous copy).
rtg_status host_ints(Context* ctx, const Value& v, std::vector<int64_t>* out);
rtg_status host_floats(Context* ctx, const Value& v, std::vector<float>* out);

// ---------------------------------------------------------------- operators
class Operator {
 public:
  virtual ~Operator() {}
  virtual const char* name() const = 0;
  virtual rtg_status run(Context* ctx, const OpInputs& in, std::vector<Value>* out) = 0;
  virtual bool can_run_in_place() const { return false; }
  virtual bool is_commutative() const { return false; }
  // `io` is an owned temporary (input 0, or the chosen operand of a commutative op); `rest` are the other inputs.
  virtual rtg_status run_in_place(Context* ctx, Value* io, const OpInputs& rest) {
    (void)ctx, (void)io, (void)rest;
    return set_error(RTG_ERR_INTERNAL, "operator does not run in place");
  }
  // Called once at load for every operator with the constant inputs known so far (nullptr where not constant):
  // lets Conv/Gemm/MatMul pre-arrange weights on the device (the role of prepack_a/prepack_b).
  virtual rtg_status prepare(Context* ctx, const OpInputs& consts) {
    (void)ctx, (void)consts;
    return RTG_OK;
  }
  virtual void release(Context* ctx) { (void)ctx; }
};

// Attribute access for operator readers (one OperatorNode of the file).
struct OpNodeView {
  uint8_t type = 0;
  uint8_t attrs_type = 0;
  fb::Table attrs;
};
using ReadOpFn = std::function<rtg_status(const OpNodeView&, std::unique_ptr<Operator>*)>;

// OpRegistry (src/op_registry.rs:31-62): operator type -> reader.
class OpRegistry {
 public:
  void register_op(uint8_t type, ReadOpFn fn) { ops_[type] = std::move(fn); }
  rtg_status read_op(const OpNodeView& node, std::unique_ptr<Operator>* out) const;
  static const OpRegistry& with_all_ops();  // every operator this library implements
  static const char* type_name(uint8_t type);
  size_t size() const { return ops_.size(); }

 private:
  std::unordered_map<uint8_t, ReadOpFn> ops_;
};

// ---------------------------------------------------------------- graph
enum class NodeKind : uint8_t { Operator, Constant, Value };

struct DimSpec {
  int64_t value = -1;   // -1 => symbolic
  std::string name;
};

struct Node {
  NodeKind kind = NodeKind::Value;
  std::string name;
  // Operator
  std::unique_ptr<Operator> op;
  uint8_t op_type = 0;
  std::vector<int32_t> inputs, outputs;  // node ids; negative = absent
  // Constant
  Value constant;                        // device-resident (uploaded at load) + host mirror for small tensors
  std::vector<uint8_t> const_host;       // host bytes (kept for constant propagation / shape ops)
  // Value
  std::vector<DimSpec> shape;
  bool has_shape = false;
};

struct TimingRecord {
  std::string op, node;
  double ms = 0;
  std::string shapes;
};

class Graph {
 public:
  std::vector<Node> nodes;
  std::vector<int32_t> input_ids, output_ids;

  // Graph::run (src/graph.rs:733-1073): plan (cached), then execute with refcounts, in-place selection and pool reuse.
  // `inputs` / `outputs` address values by node id (Model::run, src/model.rs:580-592).
  rtg_status run(Context* ctx, const std::vector<std::pair<int32_t, Value>>& inputs, const std::vector<int32_t>& outputs,
                 std::vector<Value>* results, std::vector<TimingRecord>* timing);
  // Graph::partial_run (src/graph.rs:1147-1183): with an incomplete set of inputs, evaluates every operator of the
  // plan whose inputs are available and returns the leaves of that sub-plan (values that are requested outputs or
  // feed operators that could not run) as (node id, value).
  rtg_status partial_run(Context* ctx, const std::vector<std::pair<int32_t, Value>>& inputs,
                         const std::vector<int32_t>& outputs, std::vector<std::pair<int32_t, Value>>* leaves);
  // create_plan (src/graph.rs:1256-1346): operator node ids in execution order for the requested outputs.
  // `allow_missing`: PlanOptions::allow_missing_inputs — unresolvable values are tolerated (partial_run prunes).
  rtg_status create_plan(const std::vector<int32_t>& inputs, const std::vector<int32_t>& outputs,
                         std::vector<int32_t>* plan, bool allow_missing = false) const;
  // The operator producing value `id`, or -1 (Graph::get_source_node).
  int32_t source_node(int32_t id) const;
  int num_operators() const;
  void release(Context* ctx);

 private:
  rtg_status run_plan(Context* ctx, const std::vector<std::pair<int32_t, Value>>& inputs,
                      const std::vector<int32_t>& plan, const std::vector<int32_t>& outputs, std::vector<Value>* results,
                      std::vector<TimingRecord>* timing);
  // prune_plan (src/graph.rs:1185-1235)
  void prune_plan(const std::vector<int32_t>& plan, const std::vector<int32_t>& inputs,
                  const std::vector<int32_t>& outputs, std::vector<int32_t>* pruned, std::vector<int32_t>* leaves) const;
  // get_cached_plan (src/graph.rs:513,768-795): concurrent runs on one model are allowed, the cache is behind a mutex
  std::mutex plan_mu_;
  std::vector<int32_t> cached_inputs_, cached_outputs_;
  std::shared_ptr<const std::vector<int32_t>> cached_plan_;
};

// GraphOptimizer::propagate_constants (src/optimize.rs:301-327): a partial run with no inputs evaluates everything that
// depends only on constants; those values become constant nodes.  Returns how many values were replaced.
rtg_status propagate_constants(Context* ctx, Graph* g, int* replaced);

// GPU-side GraphOptimizer (src/optimize.rs:286-297 plus the new epilogue fusions); returns how many nodes it fused.
int optimize_graph(Graph* g);

// ---------------------------------------------------------------- model
// rtg_model_submit: host inputs are uploaded on a second stream into one of two staging slots, so the upload of run
// i + 1 overlaps the graph of run i; outputs go back to the caller's host buffers on the compute stream.
struct SubmitPipeline {
  struct Slot {
    std::vector<void*> staging;   // one device buffer per graph input
    std::vector<size_t> cap;
    cudaEvent_t uploaded = nullptr, consumed = nullptr;
    bool in_use = false;          // `consumed` has been recorded at least once
  };
  cudaStream_t upload = nullptr;
  Slot slots[2];
  uint64_t next = 0;
  void release();
};

// One cached plan captured as a CUDA graph (the loop it replaces is run_plan, src/graph.rs:883-1051).
// Key: the input / output node ids, the input shapes and the math mode.  The first run with a key executes eagerly
// (it sizes the pool and lets operators settle their caches); the second is stream-captured, instantiated and then
// replayed by every later run: ~70 launches and their tensor-map encodes become one cudaGraphLaunch.  Tensor maps are
// kernel parameters, so they are baked into the graph together with the pool addresses of every temporary.
struct CapturedPlan {
  std::vector<int32_t> in_ids, out_ids;
  std::vector<rtg_tensor> in_desc;   // dtype / ndim / shape per input
  int math = 0;
  int seen = 0;                      // eager runs so far
  bool failed = false;               // capture was tried and this graph cannot be captured (host round trips)
  cudaGraphExec_t exec = nullptr;
  std::vector<Value> in_bufs;        // persistent device inputs the captured kernels read
  std::vector<Value> outs;           // outputs of the captured run; every replay overwrites them
  uint64_t launches = 0;             // kernels in the graph (Context::launches advances by this per replay)
};

struct Model {
  Context* ctx = nullptr;
  Graph graph;
  bool timing = false;
  std::string timing_report;
  std::vector<std::string> name_store;
  SubmitPipeline pipe;
  // CUDA-graph replay of cached plans.  RTEN_GPU_GRAPH=0 (or rtg_model_set_graph_capture) turns it off.
  bool graph_capture = true;
  std::mutex cap_mu;                 // replays share the captured buffers: one at a time per model
  std::vector<std::unique_ptr<CapturedPlan>> captured;
  uint64_t replays = 0;
  void release_captured();
};
rtg_status load_model(Context* ctx, const uint8_t* data, size_t len, bool optimize, Model** out);

struct Run {
  Context* ctx = nullptr;
  std::vector<Value> outputs;
  std::vector<int32_t> node_ids;  // the value node each output belongs to (the leaves, for a partial run)
};

struct Ticket {  // one rtg_model_submit in flight
  Context* ctx = nullptr;
  cudaEvent_t done = nullptr;
};

}  // namespace rtg
