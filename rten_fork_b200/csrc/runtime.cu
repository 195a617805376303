// Context, device TensorPool, error strings, host<->device staging.
#include "runtime.h"
#include <string>
#include <algorithm>

#include <stdarg.h>
#include <stdio.h>
#include <string.h>

namespace rtg {

static thread_local char g_err[512] = "";

rtg_status set_error(rtg_status code, const char* msg) {
  snprintf(g_err, sizeof(g_err), "%s", msg);
  return code;
}
rtg_status set_errorf(rtg_status code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

// ---------------------------------------------------------------- DevPool
DevPool::~DevPool() { release_all(); }

void DevPool::release_all() {
  for (auto& kv : capacity_) cudaFree(kv.first);
  capacity_.clear();
  free_.clear();
  bytes_reserved = 0;
}

rtg_status DevPool::alloc(size_t bytes, void** out) {
  alloc_count++;
  if (bytes == 0) bytes = 256;
  // Best fit, like TensorPool::alloc (src/tensor_pool.rs:129-163): the smallest pooled buffer that is large enough.
  // Unlike the reference we refuse a buffer more than 4x too large so that a 100 MB activation buffer is not burnt
  // on a 4 KB request while a long graph is running.
  auto it = free_.lower_bound(bytes);
  if (it != free_.end() && it->first <= bytes * 4 + (1 << 20)) {
    *out = it->second;
    capacity_[it->second].pooled = false;
    free_.erase(it);
    hit_count++;
    return RTG_OK;
  }
  size_t cap = (bytes + 511) & ~size_t(511);
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, cap);
  if (e != cudaSuccess && pinned_addresses > 0) {
    cudaGetLastError();
    return set_errorf(RTG_ERR_CUDA, "cudaMalloc(%zu) failed: %s (pooled buffers are pinned by captured graphs)", cap,
                      cudaGetErrorString(e));
  }
  if (e != cudaSuccess) {
    // Give pooled memory back to the driver and retry once.
    for (auto& kv : free_) {
      cudaFree(kv.second);
      bytes_reserved -= kv.first;
      capacity_.erase(kv.second);
    }
    free_.clear();
    cudaGetLastError();
    e = cudaMalloc(&p, cap);
    if (e != cudaSuccess) return set_errorf(RTG_ERR_CUDA, "cudaMalloc(%zu) failed: %s", cap, cudaGetErrorString(e));
  }
  capacity_[p] = Slot{cap, false};
  bytes_reserved += cap;
  *out = p;
  return RTG_OK;
}

rtg_status DevPool::free(void* p) {
  if (!p) return RTG_OK;
  auto it = capacity_.find(p);
  if (it == capacity_.end()) return set_error(RTG_ERR_INVALID_VALUE, "pointer was not allocated from this pool");
  if (it->second.pooled) return set_error(RTG_ERR_INVALID_VALUE, "pointer was already returned to the pool (double free)");
  if (!use_pool && pinned_addresses == 0) {
    bytes_reserved -= it->second.cap;
    capacity_.erase(it);
    RTG_CUDA(cudaFree(p));  // waits for the device: kernels still reading the buffer finish first
    return RTG_OK;
  }
  it->second.pooled = true;
  free_.emplace(it->second.cap, p);
  return RTG_OK;
}

rtg_status Context::ensure_pinned(size_t bytes) {
  if (bytes <= pinned_bytes) return RTG_OK;
  if (pinned) cudaFreeHost(pinned);
  pinned = nullptr;
  pinned_bytes = 0;
  size_t cap = (bytes + (1 << 20)) & ~size_t((1 << 20) - 1);
  RTG_CUDA(cudaMallocHost(&pinned, cap));
  pinned_bytes = cap;
  return RTG_OK;
}

}  // namespace rtg

using namespace rtg;

extern "C" {

const char* rtg_last_error(void) { return g_err; }

rtg_status rtg_ctx_create(int32_t device, rtg_ctx** out) {
  if (!out) return set_error(RTG_ERR_INVALID_VALUE, "out is NULL");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return set_errorf(RTG_ERR_CUDA, "no CUDA device available (%s); librten_gpu has no CPU fallback",
                      e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  if (device < 0 || device >= count) return set_errorf(RTG_ERR_INVALID_VALUE, "device %d out of range", device);
  RTG_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  RTG_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return set_errorf(RTG_ERR_CUDA, "device %d is sm_%d%d; librten_gpu is built for sm_100a only", device, prop.major,
                      prop.minor);
  Context* c = new Context();
  c->device = device;
  c->num_sms = prop.multiProcessorCount;
  // Environment, read once per context like the reference reads its own (src/threading.rs:34-60, src/graph.rs:871):
  //   RTEN_GPU_SMS    n  the share of the machine kernels size their grids for -- the counterpart of RTEN_NUM_THREADS
  //                      (the reference's thread pool size); persistent kernels launch at most n CTAs
  //   RTEN_GPU_MATH   tf32 | 3xtf32 | tf32_bf16x2   initial math mode of the context (rtg_ctx_set_math overrides it)
  //   RTEN_USE_POOL   0  no buffer reuse (see DevPool::use_pool)
  if (const char* e = getenv("RTEN_GPU_SMS")) {
    const int n = atoi(e);
    if (n <= 0) {
      delete c;
      return set_errorf(RTG_ERR_INVALID_VALUE, "RTEN_GPU_SMS must be a positive number, got \"%s\"", e);
    }
    c->num_sms = std::min(c->num_sms, n);
  }
  if (const char* e = getenv("RTEN_GPU_MATH")) {
    const std::string v(e);
    if (v == "tf32") c->math = RTG_MATH_TF32;
    else if (v == "3xtf32") c->math = RTG_MATH_3XTF32;
    else if (v == "tf32_bf16x2") c->math = RTG_MATH_TF32_BF16X2;
    else {
      delete c;
      return set_errorf(RTG_ERR_INVALID_VALUE, "RTEN_GPU_MATH must be tf32, 3xtf32 or tf32_bf16x2, got \"%s\"", e);
    }
  }
  if (const char* e = getenv("RTEN_USE_POOL")) {
    const std::string v(e);
    c->pool.use_pool = !(v == "0" || v == "false" || v == "no" || v == "off");  // str_as_bool, src/env.rs
  }
  RTG_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  c->own_stream = true;
  cudaDriverGetVersion(&c->driver_version);
  cudaDriverEntryPointQueryResult qres;
  void* fn = nullptr;
  RTG_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  c->encode_tiled = reinterpret_cast<EncodeTiledFn>(fn);
  fn = nullptr;
  RTG_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &qres));
  c->encode_im2col = reinterpret_cast<EncodeIm2colFn>(fn);
  if (!c->encode_tiled || !c->encode_im2col) {
    delete c;
    return set_error(RTG_ERR_CUDA, "driver does not export cuTensorMapEncodeTiled/Im2col");
  }
  *out = reinterpret_cast<rtg_ctx*>(c);
  return RTG_OK;
}

rtg_status rtg_ctx_set_debug_flags(rtg_ctx* ctx, int32_t flags) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  c->debug_flags = flags;
  return RTG_OK;
}

rtg_status rtg_ctx_destroy(rtg_ctx* ctx) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return RTG_OK;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  for (auto& kv : c->weight_cache) c->pool.free(kv.second.dev);
  c->weight_cache.clear();
  c->pool.release_all();
  if (c->pinned) cudaFreeHost(c->pinned);
  if (c->own_stream) cudaStreamDestroy(c->stream);
  delete c;
  return RTG_OK;
}

rtg_status rtg_ctx_set_stream(rtg_ctx* ctx, void* cuda_stream) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  if (c->own_stream) {
    cudaStreamSynchronize(c->stream);
    cudaStreamDestroy(c->stream);
    c->own_stream = false;
  }
  c->stream = static_cast<cudaStream_t>(cuda_stream);
  return RTG_OK;
}

rtg_status rtg_ctx_set_math(rtg_ctx* ctx, int32_t mode) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  if (mode != RTG_MATH_TF32 && mode != RTG_MATH_3XTF32 && mode != RTG_MATH_TF32_BF16X2) return set_error(RTG_ERR_INVALID_VALUE, "unknown math mode");
  c->math = mode;
  return RTG_OK;
}
rtg_status rtg_ctx_get_math(rtg_ctx* ctx, int32_t* mode) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c || !mode) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  *mode = c->math;
  return RTG_OK;
}

rtg_status rtg_ctx_sync(rtg_ctx* ctx) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  RTG_CUDA(cudaStreamSynchronize(c->stream));
  return RTG_OK;
}

rtg_status rtg_ctx_launch_count(rtg_ctx* ctx, uint64_t* count) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c || !count) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  *count = c->launches;
  return RTG_OK;
}

rtg_status rtg_ctx_set_kernel_timing(rtg_ctx* ctx, int32_t enabled) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  c->ktiming = enabled != 0;
  return RTG_OK;
}

rtg_status rtg_ctx_kernel_timing(rtg_ctx* ctx, uint64_t* launches, double* total_ms, double* flops, double* bytes) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_CUDA(cudaSetDevice(c->device));
  RTG_CUDA(cudaStreamSynchronize(c->stream));
  double ms = 0, fl = 0, by = 0;
  for (auto& r : c->krecs) {
    float t = 0;
    cudaEventElapsedTime(&t, r.a, r.b);
    ms += t, fl += r.flops, by += r.bytes;
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  if (launches) *launches = c->krecs.size();
  if (total_ms) *total_ms = ms;
  if (flops) *flops = fl;
  if (bytes) *bytes = by;
  c->krecs.clear();
  return RTG_OK;
}

const char* rtg_ctx_last_kernel(rtg_ctx* ctx) {
  Context* c = reinterpret_cast<Context*>(ctx);
  return c ? c->last_kernel.c_str() : "";
}

rtg_status rtg_pool_alloc(rtg_ctx* ctx, size_t bytes, void** dptr) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c || !dptr) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_CUDA(cudaSetDevice(c->device));
  return c->pool.alloc(bytes, dptr);
}
rtg_status rtg_pool_free(rtg_ctx* ctx, void* dptr) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  return c->pool.free(dptr);
}
rtg_status rtg_pool_stats(rtg_ctx* ctx, uint64_t* alloc_count, uint64_t* hit_count, uint64_t* len) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  if (alloc_count) *alloc_count = c->pool.alloc_count;
  if (hit_count) *hit_count = c->pool.hit_count;
  if (len) *len = c->pool.len();
  return RTG_OK;
}

rtg_status rtg_upload(rtg_ctx* ctx, void* dst_device, const void* src_host, size_t bytes) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  RTG_CUDA(cudaMemcpyAsync(dst_device, src_host, bytes, cudaMemcpyHostToDevice, c->stream));
  RTG_CUDA(cudaStreamSynchronize(c->stream));
  return RTG_OK;
}
rtg_status rtg_download(rtg_ctx* ctx, void* dst_host, const void* src_device, size_t bytes) {
  Context* c = reinterpret_cast<Context*>(ctx);
  if (!c) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  RTG_CUDA(cudaMemcpyAsync(dst_host, src_device, bytes, cudaMemcpyDeviceToHost, c->stream));
  RTG_CUDA(cudaStreamSynchronize(c->stream));
  return RTG_OK;
}

}  // extern "C"
