// C ABI: rtg_gemm — the GemmExecutor entry point (reference: src/gemm.rs:448-542, 733-765).
#include "gemm_host.h"

#include <string.h>

using namespace rtg;

namespace rtg {

// Elements spanned by a [rows, cols] view with the given strides.
static size_t span_elems(size_t rows, size_t cols, ptrdiff_t rs, ptrdiff_t cs) {
  if (rows == 0 || cols == 0) return 0;
  return (rows - 1) * static_cast<size_t>(rs) + (cols - 1) * static_cast<size_t>(cs) + 1;
}

}  // namespace rtg

extern "C" rtg_status rtg_gemm(rtg_ctx* ctx_, int32_t device, float* out, size_t out_len, size_t out_row_stride,
                               const float* a, size_t m, size_t k, ptrdiff_t a_rs, ptrdiff_t a_cs, const float* b,
                               size_t b_rows, size_t n, ptrdiff_t b_rs, ptrdiff_t b_cs, float alpha, float beta,
                               const float* bias, size_t bias_len) {
  Context* ctx = reinterpret_cast<Context*>(ctx_);
  if (!ctx) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  // The reference asserts these (src/gemm.rs:743-750); a panic becomes a status code.
  if (k != b_rows) return set_error(RTG_ERR_PANIC, "Columns of matrix `a` must match rows of matrix `b`");
  if (bias && bias_len != m) return set_error(RTG_ERR_PANIC, "Bias vector length must match rows of matrix `a`");
  if (a_rs < 0 || a_cs < 0 || b_rs < 0 || b_cs < 0) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "negative strides");
  if (m == 0 || n == 0) return RTG_OK;  // src/gemm.rs:753-755
  RTG_CUDA(cudaSetDevice(ctx->device));
  const bool host = device == RTG_HOST;
  if (!host && device != ctx->device) return set_error(RTG_ERR_INVALID_VALUE, "tensor device does not match context");

  if (k == 0) {
    // src/gemm.rs:757-765: the whole output slice becomes beta * out (0 when beta == 0); bias is not applied.
    if (host) {
      for (size_t i = 0; i < out_len; ++i) out[i] = beta == 0.f ? 0.f : beta * out[i];
      return RTG_OK;
    }
    return scale_buffer(ctx, out, out_len, beta);
  }
  if (out_row_stride < n || (m - 1) * out_row_stride + n > out_len)
    return set_error(RTG_ERR_PANIC, "Output buffer should be large enough");

  Scratch da, db, dout, dbias;
  const float *pa = a, *pb = b, *pbias = bias;
  float* pout = out;
  const size_t out_span = (m - 1) * out_row_stride + n;
  if (host) {
    const size_t na = span_elems(m, k, a_rs, a_cs), nb = span_elems(k, n, b_rs, b_cs);
    RTG_TRY(da.alloc(ctx, na * 4));
    RTG_TRY(db.alloc(ctx, nb * 4));
    RTG_TRY(dout.alloc(ctx, out_span * 4));
    RTG_CUDA(cudaMemcpyAsync(da.p, a, na * 4, cudaMemcpyHostToDevice, ctx->stream));
    RTG_CUDA(cudaMemcpyAsync(db.p, b, nb * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (beta != 0.f) RTG_CUDA(cudaMemcpyAsync(dout.p, out, out_span * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (bias) {
      RTG_TRY(dbias.alloc(ctx, m * 4));
      RTG_CUDA(cudaMemcpyAsync(dbias.p, bias, m * 4, cudaMemcpyHostToDevice, ctx->stream));
      pbias = dbias.f32();
    }
    pa = da.f32();
    pb = db.f32();
    pout = dout.f32();
  }

  Epilogue ep;
  ep.alpha = alpha;
  ep.beta = beta;
  ep.bias = pbias;
  ep.bias_mode = BIAS_ROW;
  ep.R = beta != 0.f ? pout : nullptr;
  OutMap om = {static_cast<int64_t>(out_row_stride), 0, 0, 0};
  ep.rmap = om;
  RTG_TRY(gemm_device(ctx, MatView{pa, (int64_t)m, (int64_t)k, a_rs, a_cs}, MatView{pb, (int64_t)k, (int64_t)n, b_rs, b_cs},
                      pout, om, ep));
  if (host) {
    if (out_row_stride == n) {
      RTG_CUDA(cudaMemcpyAsync(out, pout, out_span * 4, cudaMemcpyDeviceToHost, ctx->stream));
    } else {
      // only the [m, n] window belongs to the result; the gaps between rows stay untouched
      RTG_CUDA(cudaMemcpy2DAsync(out, out_row_stride * 4, pout, out_row_stride * 4, n * 4, m, cudaMemcpyDeviceToHost,
                                 ctx->stream));
    }
    RTG_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  return RTG_OK;
}

// ---------------------------------------------------------------- split-K GEMM on static scratch (allocates nothing)
extern "C" rtg_status rtg_gemm_static_scratch(rtg_ctx* ctx_, float* out, size_t out_len, size_t out_row_stride,
                                              const float* a, size_t m, size_t k, const float* b, size_t n, float alpha,
                                              float beta, const float* bias, int32_t splits) {
  Context* ctx = reinterpret_cast<Context*>(ctx_);
  if (!ctx) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  if (!out || !a || !b) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  if (m == 0 || n == 0) return RTG_OK;
  if (out_row_stride < n || (m - 1) * out_row_stride + n > out_len)
    return set_error(RTG_ERR_PANIC, "Output buffer should be large enough");
  RTG_CUDA(cudaSetDevice(ctx->device));
  return gemm_splitk_static(ctx, a, static_cast<int64_t>(m), static_cast<int64_t>(k), b, static_cast<int64_t>(n), out,
                            static_cast<int64_t>(out_row_stride), alpha, beta, bias, splits);
}

// ---------------------------------------------------------------- N-split GEMM with peer stores
extern "C" rtg_status rtg_symm_alloc(rtg_ctx* ctx_, size_t bytes, void** dptr, uint8_t handle[RTG_SYMM_HANDLE_BYTES]) {
  Context* ctx = reinterpret_cast<Context*>(ctx_);
  if (!ctx || !dptr || !handle) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == RTG_SYMM_HANDLE_BYTES, "handle size");
  RTG_CUDA(cudaSetDevice(ctx->device));
  // a dedicated allocation: IPC handles name whole cudaMalloc regions, not pool slices
  RTG_CUDA(cudaMalloc(dptr, bytes ? bytes : 16));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, *dptr);
  if (e != cudaSuccess) {
    cudaFree(*dptr);
    *dptr = nullptr;
    return set_errorf(RTG_ERR_CUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  memcpy(handle, &h, sizeof(h));
  return RTG_OK;
}

extern "C" rtg_status rtg_symm_free(rtg_ctx* ctx_, void* dptr) {
  Context* ctx = reinterpret_cast<Context*>(ctx_);
  if (!ctx) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  RTG_CUDA(cudaSetDevice(ctx->device));
  RTG_CUDA(cudaStreamSynchronize(ctx->stream));
  RTG_CUDA(cudaFree(dptr));
  return RTG_OK;
}

extern "C" rtg_status rtg_symm_open(rtg_ctx* ctx_, const uint8_t handle[RTG_SYMM_HANDLE_BYTES], void** peer_dptr) {
  Context* ctx = reinterpret_cast<Context*>(ctx_);
  if (!ctx || !handle || !peer_dptr) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  RTG_CUDA(cudaSetDevice(ctx->device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  // maps the peer's allocation into this process and enables peer access to its device on first use
  RTG_CUDA(cudaIpcOpenMemHandle(peer_dptr, h, cudaIpcMemLazyEnablePeerAccess));
  return RTG_OK;
}

extern "C" rtg_status rtg_symm_close(rtg_ctx* ctx_, void* peer_dptr) {
  Context* ctx = reinterpret_cast<Context*>(ctx_);
  if (!ctx) return set_error(RTG_ERR_INVALID_VALUE, "ctx is NULL");
  RTG_CUDA(cudaSetDevice(ctx->device));
  RTG_CUDA(cudaStreamSynchronize(ctx->stream));
  RTG_CUDA(cudaIpcCloseMemHandle(peer_dptr));
  return RTG_OK;
}

extern "C" rtg_status rtg_gemm_scatter(rtg_ctx* ctx_, const float* a, size_t m, size_t k, ptrdiff_t a_rs,
                                       ptrdiff_t a_cs, const float* b, size_t n_block, ptrdiff_t b_rs, ptrdiff_t b_cs,
                                       size_t col_offset, size_t n_total, float* const* dst, int32_t n_dst) {
  Context* ctx = reinterpret_cast<Context*>(ctx_);
  if (!ctx || !dst) return set_error(RTG_ERR_INVALID_VALUE, "NULL argument");
  if (n_dst < 1 || n_dst > 8) return set_error(RTG_ERR_INVALID_VALUE, "n_dst must be in 1..8");
  if (a_rs < 0 || a_cs < 0 || b_rs < 0 || b_cs < 0) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "negative strides");
  if (col_offset + n_block > n_total) return set_error(RTG_ERR_PANIC, "Output buffer should be large enough");
  if (col_offset % 4 != 0 || n_total % 4 != 0)
    return set_error(RTG_ERR_UNSUPPORTED_VALUE, "col_offset and n_total must be multiples of 4");
  if (m == 0 || n_block == 0) return RTG_OK;
  if (k == 0) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "k == 0");
  for (int i = 0; i < n_dst; ++i)
    if (!dst[i] || (reinterpret_cast<uintptr_t>(dst[i]) & 15) != 0)
      return set_error(RTG_ERR_INVALID_VALUE, "dst buffers must be non-NULL and 16-byte aligned");
  RTG_CUDA(cudaSetDevice(ctx->device));
  Epilogue ep;
  ep.n_peers = n_dst - 1;
  for (int i = 1; i < n_dst; ++i) ep.peer_D[i - 1] = dst[i];
  OutMap om = {static_cast<int64_t>(n_total), 0, 0, static_cast<int32_t>(col_offset)};
  return gemm_device(ctx, MatView{a, (int64_t)m, (int64_t)k, a_rs, a_cs},
                     MatView{b, (int64_t)k, (int64_t)n_block, b_rs, b_cs}, dst[0], om, ep);
}
