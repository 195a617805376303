// Host-side runtime shared by every translation unit of librten_gpu.so: status/error plumbing, the per-GPU context,
// the device TensorPool and the weight cache.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/rten_gpu.h"

namespace rtg {

// ---------------------------------------------------------------- errors
rtg_status set_error(rtg_status code, const char* msg);
rtg_status set_errorf(rtg_status code, const char* fmt, ...);
#define RTG_CUDA(expr)                                                                                   \
  do {                                                                                                   \
    cudaError_t _e = (expr);                                                                             \
    if (_e != cudaSuccess)                                                                               \
      return ::rtg::set_errorf(RTG_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)
#define RTG_TRY(expr)                 \
  do {                                \
    rtg_status _s = (expr);           \
    if (_s != RTG_OK) return _s;      \
  } while (0)

// ---------------------------------------------------------------- device TensorPool
// Reference: src/tensor_pool.rs:103-199.  Same policy: a free list of buffers, `alloc` takes the smallest free buffer
// whose capacity fits (best fit), otherwise allocates; `add` returns a buffer.  Reuse is safe without events because
// a pool belongs to one context and every consumer is ordered on that context's stream.
class DevPool {
 public:
  ~DevPool();
  rtg_status alloc(size_t bytes, void** out);
  rtg_status free(void* p);
  void release_all();
  uint64_t alloc_count = 0, hit_count = 0;
  size_t len() const { return free_.size(); }
  size_t bytes_reserved = 0;
  // > 0 while captured CUDA graphs reference addresses of this pool (Model graph capture): pooled buffers are then
  // never given back to the driver, their addresses are baked into the graphs
  int pinned_addresses = 0;
  // RTEN_USE_POOL=0 (src/graph.rs:871 env_flag("RTEN_USE_POOL", true)): released buffers go back to the driver instead
  // of the pool, every allocation is a cudaMalloc.  A debugging aid, as in the reference; it also turns graph capture off.
  bool use_pool = true;

 private:
  struct Slot {
    size_t cap;
    bool pooled;  // sitting in free_ (a second free of the same pointer is an error, not a silent double entry)
  };
  std::multimap<size_t, void*> free_;          // capacity -> buffer
  std::unordered_map<void*, Slot> capacity_;   // every live or pooled buffer
};

// ---------------------------------------------------------------- context
struct CachedWeight {
  void* dev = nullptr;
  size_t bytes = 0;
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct Context {
  int device = 0;
  int num_sms = 148;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int math = RTG_MATH_3XTF32;
  int debug_flags = 0;  // rtg_ctx_set_debug_flags (developer / test switches), OR-ed with RTG_DEBUG_FLAGS
  DevPool pool;
  uint64_t launches = 0;
  std::string last_kernel;
  EncodeTiledFn encode_tiled = nullptr;
  EncodeIm2colFn encode_im2col = nullptr;
  int driver_version = 0;
  // Optional per-launch timing of the contraction kernel (bench.py's roofline): CUDA events recorded on `stream`
  // around every gemm_tcgen05 launch together with that launch's algorithmic flops / bytes.
  struct KernelRec {
    cudaEvent_t a, b;
    double flops, bytes;
  };
  bool ktiming = false;
  std::vector<KernelRec> krecs;
  double pending_flops = 0, pending_bytes = 0;  // set by the caller of the next contraction launch
  // Pre-arranged weights, keyed by (source pointer, tag).  The role of PackedAMatrix / PackedBMatrix
  // (src/gemm.rs:26-92), except that entries persist across calls for constant operands.
  std::mutex cache_mu;
  std::map<std::pair<const void*, uint64_t>, CachedWeight> weight_cache;
  // pinned staging for host tensors
  void* pinned = nullptr;
  size_t pinned_bytes = 0;
  rtg_status ensure_pinned(size_t bytes);
};

// RAII scratch buffer from the context pool.
struct Scratch {
  Context* ctx = nullptr;
  void* p = nullptr;
  Scratch() {}
  Scratch(const Scratch&) = delete;
  Scratch& operator=(const Scratch&) = delete;
  ~Scratch() { reset(); }
  rtg_status alloc(Context* c, size_t bytes) {
    reset();
    ctx = c;
    return c->pool.alloc(bytes, &p);
  }
  void reset() {
    if (p) ctx->pool.free(p);
    p = nullptr;
  }
  float* f32() const { return static_cast<float*>(p); }
};

inline int64_t div_ceil(int64_t a, int64_t b) { return (a + b - 1) / b; }

}  // namespace rtg
