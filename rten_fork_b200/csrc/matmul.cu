// MatMul and Gemm operators on top of the contraction engine (reference: src/ops/matmul.rs:27-81 gemm_op,
// :119-239 matmul_impl).
#include "ops.h"

namespace rtg {

rtg_status matmul_out_shape(const rtg_tensor* a, const rtg_tensor* b, int64_t* shape, int32_t* ndim) {
  if (!a || !b) return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  if (a->dtype != RTG_F32 || b->dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  if (a->ndim < 2 || b->ndim < 2) return set_error(RTG_ERR_INVALID_VALUE, "Inputs must have >= 2 dimensions");
  if (a->shape[a->ndim - 1] != b->shape[b->ndim - 2])
    return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Columns of first matrix does not match rows of second matrix");
  const int pa = a->ndim - 2, pb = b->ndim - 2, pn = std::max(pa, pb);
  if (pn + 2 > RTG_MAX_DIMS) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "too many dimensions");
  for (int i = 0; i < pn; ++i) {
    const int ia = i - (pn - pa), ib = i - (pn - pb);
    const int64_t da = ia >= 0 ? a->shape[ia] : 1, db = ib >= 0 ? b->shape[ib] : 1;
    if (da == db) shape[i] = da;
    else if (da == 1) shape[i] = db;
    else if (db == 1) shape[i] = da;
    else return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Cannot broadcast shapes");
  }
  shape[pn] = a->shape[a->ndim - 2];
  shape[pn + 1] = b->shape[b->ndim - 1];
  *ndim = pn + 2;
  return RTG_OK;
}

rtg_status matmul_forward(Context* ctx, const rtg_tensor& a, const rtg_tensor& b, const rtg_tensor& out,
                          const Epilogue* fused) {
  int64_t oshape[RTG_MAX_DIMS];
  int32_t ond = 0;
  RTG_TRY(matmul_out_shape(&a, &b, oshape, &ond));
  const int64_t M = a.shape[a.ndim - 2], K = a.shape[a.ndim - 1], N = b.shape[b.ndim - 1];
  int64_t total = 1, num_a = 1, num_b = 1;
  for (int i = 0; i < ond; ++i) total *= oshape[i];
  for (int i = 0; i < a.ndim - 2; ++i) num_a *= a.shape[i];
  for (int i = 0; i < b.ndim - 2; ++i) num_b *= b.shape[i];
  if (total == 0) return RTG_OK;
  if (!is_contiguous(out)) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "matmul output must be contiguous");
  float* D = static_cast<float*>(out.data);
  Epilogue ep;
  if (fused) ep = *fused;
  if (K == 0) {
    if (fused && (fused->bias || fused->act != ACT_NONE || fused->alpha != 1.f))
      return set_error(RTG_ERR_UNSUPPORTED_VALUE, "fused epilogue with K == 0");
    RTG_CUDA(cudaMemsetAsync(D, 0, static_cast<size_t>(total) * 4, ctx->stream));
    return RTG_OK;
  }
  const float* pa = static_cast<const float*>(a.data);
  const float* pb = static_cast<const float*>(b.data);
  const int64_t a_rs = a.strides[a.ndim - 2], a_cs = a.strides[a.ndim - 1];
  const int64_t b_rs = b.strides[b.ndim - 2], b_cs = b.strides[b.ndim - 1];

  // [A.., M, K] x [K, N] collapses to one [A*M, K] x [K, N] product (src/ops/matmul.rs:155-169)
  if (num_a > 1 && num_b == 1) {
    Scratch tmp;
    int64_t rs = a_rs;
    bool collapsible = a_cs == 1 || K == 1;
    // rows of consecutive matrices must be equally spaced: stride[i] == shape[i+1] * stride[i+1] up the prefix
    for (int i = a.ndim - 3; i >= 0 && collapsible; --i)
      if (a.shape[i] != 1 && a.strides[i] != a.shape[i + 1] * a.strides[i + 1]) collapsible = false;
    if (!collapsible) {
      RTG_TRY(tmp.alloc(ctx, static_cast<size_t>(numel(a)) * 4));
      rtg_tensor ac = a;
      ac.data = tmp.p;
      set_contiguous_strides(ac);
      RTG_TRY(copy_view(ctx, ac, a));
      pa = tmp.f32();
      rs = K;
      return gemm_device(ctx, MatView{pa, num_a * M, K, rs, 1}, MatView{pb, K, N, b_rs, b_cs}, D, OutMap{N, 0, 0, 0}, ep);
    }
    return gemm_device(ctx, MatView{pa, num_a * M, K, rs, a_cs}, MatView{pb, K, N, b_rs, b_cs}, D, OutMap{N, 0, 0, 0},
                       ep);
  }

  // general case (src/ops/matmul.rs:171-238): one product per broadcast batch index.  Batch dimensions whose strides
  // nest are merged; up to two remaining dimensions run as ONE batched launch (TMA walks the batch coordinates).
  const int pn = ond - 2;
  {
    int64_t size[RTG_MAX_DIMS], sa[RTG_MAX_DIMS], sb[RTG_MAX_DIMS];
    int nd = 0;  // outermost first
    for (int i = 0; i < pn; ++i) {
      if (oshape[i] == 1) continue;
      const int ia = i - (pn - (a.ndim - 2)), ib = i - (pn - (b.ndim - 2));
      const int64_t ta = (ia >= 0 && a.shape[ia] != 1) ? a.strides[ia] : 0;
      const int64_t tb = (ib >= 0 && b.shape[ib] != 1) ? b.strides[ib] : 0;
      if (nd > 0 && sa[nd - 1] == oshape[i] * ta && sb[nd - 1] == oshape[i] * tb) {
        size[nd - 1] *= oshape[i], sa[nd - 1] = ta, sb[nd - 1] = tb;  // merge with the dimension outside it
      } else {
        size[nd] = oshape[i], sa[nd] = ta, sb[nd] = tb, ++nd;
      }
    }
    if (nd >= 1 && nd <= 2) {
      const int64_t n0 = size[nd - 1], n1 = nd == 2 ? size[0] : 1;
      BatchedView va{pa, M, K, a_rs, a_cs, sa[nd - 1], nd == 2 ? sa[0] : 0};
      BatchedView vb{pb, K, N, b_rs, b_cs, sb[nd - 1], nd == 2 ? sb[0] : 0};
      if (gemm_batched_supported(va, vb, n0, n1, N)) {
        ep.rmap = OutMap{N, 0, 0, 0};
        return gemm_batched_device(ctx, va, vb, n0, n1, D, ep);
      }
    }
  }
  int64_t idx[RTG_MAX_DIMS] = {0};
  const int64_t batches = total / (M * N);
  for (int64_t it = 0; it < batches; ++it) {
    int64_t off_a = 0, off_b = 0;
    for (int i = 0; i < pn; ++i) {
      const int ia = i - (pn - (a.ndim - 2)), ib = i - (pn - (b.ndim - 2));
      if (ia >= 0 && a.shape[ia] != 1) off_a += idx[i] * a.strides[ia];
      if (ib >= 0 && b.shape[ib] != 1) off_b += idx[i] * b.strides[ib];
    }
    RTG_TRY(gemm_device(ctx, MatView{pa + off_a, M, K, a_rs, a_cs}, MatView{pb + off_b, K, N, b_rs, b_cs},
                        D + it * M * N, OutMap{N, 0, 0, 0}, ep));
    for (int i = pn - 1; i >= 0; --i) {
      if (++idx[i] < oshape[i]) break;
      idx[i] = 0;
    }
  }
  return RTG_OK;
}

rtg_status gemm_op_out_shape(const rtg_tensor* a, const rtg_tensor* b, bool ta, bool tb, int64_t* shape) {
  if (!a || !b) return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  if (a->dtype != RTG_F32 || b->dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  if (a->ndim != 2) return set_error(RTG_ERR_INVALID_VALUE, "a must have 2 dims");
  if (b->ndim != 2) return set_error(RTG_ERR_INVALID_VALUE, "b must have 2 dims");
  shape[0] = ta ? a->shape[1] : a->shape[0];
  shape[1] = tb ? b->shape[0] : b->shape[1];
  return RTG_OK;
}

rtg_status gemm_op_forward(Context* ctx, const rtg_tensor& a, const rtg_tensor& b, const rtg_tensor* c, float alpha,
                           float beta, bool ta, bool tb, const rtg_tensor& out, const BPack* b_pack) {
  int64_t os[2];
  RTG_TRY(gemm_op_out_shape(&a, &b, ta, tb, os));
  const int64_t M = os[0], N = os[1];
  const int64_t K = ta ? a.shape[0] : a.shape[1], Kb = tb ? b.shape[1] : b.shape[0];
  // transposition is a view change (src/ops/matmul.rs:39-40)
  const int64_t a_rs = ta ? a.strides[1] : a.strides[0], a_cs = ta ? a.strides[0] : a.strides[1];
  const int64_t b_rs = tb ? b.strides[1] : b.strides[0], b_cs = tb ? b.strides[0] : b.strides[1];
  const bool use_c = c != nullptr && beta != 0.f;
  if (use_c) {
    // can_broadcast_to(out_shape)
    if (c->ndim > 2) return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Cannot broadcast c to output shape");
    for (int i = 0; i < c->ndim; ++i) {
      const int64_t d = c->shape[c->ndim - 1 - i], t = os[1 - i];
      if (d != 1 && d != t) return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "Cannot broadcast c to output shape");
    }
  }
  if (K != Kb) return set_error(RTG_ERR_PANIC, "Columns of matrix `a` must match rows of matrix `b`");
  if (M * N == 0) return RTG_OK;
  if (!is_contiguous(out)) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "gemm output must be contiguous");
  float* D = static_cast<float*>(out.data);
  Epilogue ep;
  ep.alpha = alpha;
  if (b_pack && b_pack->p) ep.b_pack = *b_pack, ep.b_pack_const = true;  // packed at load time (GemmOp::prepare)
  // c that is one value per output column (the bias of a linear layer) with beta = 1: alpha * acc + 1.0 * c[n] is the
  // epilogue's per-column bias term exactly, without first expanding c into the output and reading it back
  const bool c_is_row = use_c && beta == 1.f && c->device != RTG_HOST && c->shape[c->ndim > 0 ? c->ndim - 1 : 0] == N &&
                        (c->ndim == 1 || (c->ndim == 2 && c->shape[0] == 1)) && c->strides[c->ndim - 1] == 1;
  if (c_is_row) {
    ep.bias = static_cast<const float*>(c->data);
    ep.bias_mode = BIAS_COL;
  } else if (use_c) {
    // expand c into the output, then accumulate with beta (src/ops/matmul.rs:46-64)
    rtg_tensor cb = {};
    cb.data = c->data, cb.dtype = RTG_F32, cb.ndim = 2, cb.device = c->device;
    cb.shape[0] = M, cb.shape[1] = N;
    for (int i = 0; i < 2; ++i) {
      const int ic = i - (2 - c->ndim);
      cb.strides[i] = (ic >= 0 && c->shape[ic] != 1) ? c->strides[ic] : 0;
    }
    RTG_TRY(copy_view(ctx, out, cb));
    ep.beta = beta;
    ep.R = D;
    ep.rmap = OutMap{N, 0, 0, 0};
  }
  if (K == 0) {
    if (use_c) return scale_buffer(ctx, D, static_cast<size_t>(M * N), beta);
    RTG_CUDA(cudaMemsetAsync(D, 0, static_cast<size_t>(M * N) * 4, ctx->stream));
    return RTG_OK;
  }
  return gemm_device(ctx, MatView{static_cast<const float*>(a.data), M, K, a_rs, a_cs},
                     MatView{static_cast<const float*>(b.data), K, N, b_rs, b_cs}, D, OutMap{N, 0, 0, 0}, ep);
}

}  // namespace rtg
