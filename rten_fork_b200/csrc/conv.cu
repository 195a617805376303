// Conv operator on the device (reference: src/ops/conv.rs:86-311 conv(), conv_2d_pointwise :24-68,
// src/ops/conv/depthwise.rs:127-203, src/ops/conv/im2col.rs).  Dense and grouped convolutions run as implicit GEMM on
// the tcgen05 kernel with TMA im2col loads over NHWC activations (pixels are GEMM rows, output channels GEMM columns,
// batch folded into M); depthwise runs as a direct CUDA-core kernel.
#include "conv_kernels.cuh"
#include "depthwise_tma.cuh"
#include "ops.h"

namespace rtg {

// ---------------------------------------------------------------- shape / attribute checks
rtg_status conv_plan(const rtg_tensor* input, const rtg_tensor* weight, const rtg_tensor* bias,
                     const rtg_conv_attrs* a, ConvPlan* p) {
  if (!input || !weight) return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  if (input->dtype != RTG_F32 || weight->dtype != RTG_F32 || (bias && bias->dtype != RTG_F32))
    return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  int64_t pads[4] = {0, 0, 0, 0};
  int64_t strides[2], dils[2];
  if (input->ndim == 3) {
    // 1-D convolution is lifted to 2-D (src/ops/conv.rs:98-139)
    if (weight->ndim != 3) return set_error(RTG_ERR_INVALID_VALUE, "kernel must have 3 dims (OCW)");
    p->is_1d = true;
    p->n = input->shape[0], p->c = input->shape[1], p->h = 1, p->w = input->shape[2];
    p->o = weight->shape[0], p->cg = weight->shape[1], p->kh = 1, p->kw = weight->shape[2];
    if (a->pad_mode == RTG_PAD_FIXED) {
      if (a->n_pads != 2) return set_error(RTG_ERR_INVALID_VALUE, "expected 2 pad values");
      pads[1] = a->pads[0], pads[3] = a->pads[1];
    }
    if (a->n_strides != 1) return set_error(RTG_ERR_INVALID_VALUE, "expected 1 stride value");
    if (a->n_dilations != 1) return set_error(RTG_ERR_INVALID_VALUE, "expected 1 dilation value");
    strides[0] = 1, strides[1] = a->strides[0];
    dils[0] = 1, dils[1] = a->dilations[0];
  } else {
    if (input->ndim != 4) return set_error(RTG_ERR_INVALID_VALUE, "input must have 4 dims (NCHW)");
    if (weight->ndim != 4) return set_error(RTG_ERR_INVALID_VALUE, "kernel must have 4 dims (OCHW)");
    p->is_1d = false;
    p->n = input->shape[0], p->c = input->shape[1], p->h = input->shape[2], p->w = input->shape[3];
    p->o = weight->shape[0], p->cg = weight->shape[1], p->kh = weight->shape[2], p->kw = weight->shape[3];
    if (a->n_strides != 2) return set_error(RTG_ERR_INVALID_VALUE, "expected 2 stride values");
    if (a->n_dilations != 2) return set_error(RTG_ERR_INVALID_VALUE, "expected 2 dilation values");
    strides[0] = a->strides[0], strides[1] = a->strides[1];
    dils[0] = a->dilations[0], dils[1] = a->dilations[1];
    if (a->pad_mode == RTG_PAD_FIXED) {
      if (a->n_pads != 4) return set_error(RTG_ERR_INVALID_VALUE, "Expected 4 padding values");
      for (int i = 0; i < 4; ++i) pads[i] = a->pads[i];
    }
  }
  if (bias && bias->ndim != 1) return set_error(RTG_ERR_INVALID_VALUE, "bias must have 1 dims");

  // calc_output_size_and_padding (src/ops/pooling.rs:27-89)
  if (dils[0] <= 0 || dils[1] <= 0) return set_error(RTG_ERR_INVALID_VALUE, "Dilations must be > 0");
  if (strides[0] <= 0 || strides[1] <= 0) return set_error(RTG_ERR_INVALID_VALUE, "Strides must be > 0");
  if (p->kh <= 0 || p->kw <= 0) return set_error(RTG_ERR_PANIC, "kernel size must be > 0");
  if (a->pad_mode == RTG_PAD_SAME) {
    p->oh = div_ceil(p->h, strides[0]);
    p->ow = div_ceil(p->w, strides[1]);
    if (p->oh == 0 || p->ow == 0) return set_error(RTG_ERR_PANIC, "attempt to subtract with overflow");
    const int64_t th = std::max<int64_t>(0, (p->oh - 1) * strides[0] + (p->kh - 1) * dils[0] + 1 - p->h);
    const int64_t tw = std::max<int64_t>(0, (p->ow - 1) * strides[1] + (p->kw - 1) * dils[1] + 1 - p->w);
    pads[0] = th / 2, pads[1] = tw / 2, pads[2] = div_ceil(th, 2), pads[3] = div_ceil(tw, 2);  // SAME_UPPER
  } else {
    for (int i = 0; i < 4; ++i)
      if (pads[i] < 0) return set_error(RTG_ERR_INVALID_VALUE, "padding must be >= 0");
    const int64_t ph = p->h + pads[0] + pads[2], pw = p->w + pads[1] + pads[3];
    const int64_t dkh = p->kh + (p->kh - 1) * (dils[0] - 1), dkw = p->kw + (p->kw - 1) * (dils[1] - 1);
    if (ph < dkh || pw < dkw) return set_error(RTG_ERR_INVALID_VALUE, "Input too small for kernel size");
    p->oh = (ph - dils[0] * (p->kh - 1) - 1) / strides[0] + 1;
    p->ow = (pw - dils[1] * (p->kw - 1) - 1) / strides[1] + 1;
  }
  for (int i = 0; i < 4; ++i) p->pads[i] = pads[i];
  p->stride_h = strides[0], p->stride_w = strides[1], p->dil_h = dils[0], p->dil_w = dils[1];
  p->groups = a->groups;

  const bool has_padding = pads[0] > 0 || pads[1] > 0 || pads[2] > 0 || pads[3] > 0;
  p->pointwise = p->kh == 1 && p->kw == 1 && !has_padding && p->groups == 1 && strides[0] == 1 && strides[1] == 1 &&
                 dils[0] == 1 && dils[1] == 1;
  p->depthwise = false;
  if (p->pointwise) {
    // The reference goes straight to the GEMM here, whose assert fires on a channel mismatch (src/gemm.rs:743).
    if (p->cg != p->c) return set_error(RTG_ERR_PANIC, "Columns of matrix `a` must match rows of matrix `b`");
  } else {
    if (p->groups == 0) return set_error(RTG_ERR_PANIC, "attempt to divide by zero");
    if (p->c / p->groups != p->cg)
      return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES,
                       "Input channels (per group) does not match kernel input channels");
    if (p->c % p->groups != 0 || p->o % p->groups != 0)
      return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES,
                       "Input channels and output channels must be divisible by group count");
    p->depthwise = p->c == p->o && p->groups == p->c;
  }
  if (bias && bias->shape[0] != p->o) return set_error(RTG_ERR_PANIC, "Bias vector length must match rows of matrix `a`");
  return RTG_OK;
}

// Row-window path: dense convolution whose kernel row (kw taps x channels padded to 4) fits one 32-float k-block.
// With so few channels the im2col path would fill at most a quarter of every per-tap channel block.
bool conv_window_supported(const ConvPlan& p) {
  if (p.pointwise || p.depthwise || p.groups != 1 || p.dil_w != 1) return false;
  const int64_t cp = div_ceil(p.c, 4) * 4;
  if (p.c > 8 || p.kw * cp > kBK || p.kh > 32) return false;
  if (p.stride_w < 1 || p.stride_h < 1) return false;
  return static_cast<int64_t>(p.stride_w) * cp * 4 % 16 == 0;
}

// ---------------------------------------------------------------- weights
// Packed copy (BPack) of the prepared K-major weight matrix [o][k] (row pitch k): computed once here, so that
// RTG_MATH_TF32_BF16X2 launches load both of its planes by TMA instead of splitting the weights again in every CTA.
static rtg_status prepare_wpack(Context* ctx, ConvWeights* out, int64_t o, int64_t k) {
  const int64_t ld = b_pack_ld(k);
  void* d = nullptr;
  RTG_TRY(ctx->pool.alloc(static_cast<size_t>(o) * ld * 4 * 2, &d));
  out->wpack = BPack{static_cast<const float*>(d), ld, o * ld};
  return make_b_pack(ctx, MatView{out->wmat, k, o, 1, k}, static_cast<float*>(d), o * ld);
}

rtg_status conv_prepare_weights(Context* ctx, const ConvPlan& p, const float* w_dev, ConvWeights* out) {
  out->wmat = nullptr;
  out->wpack = BPack{};
  out->direct = false;
  out->window = false;
  const int khw = static_cast<int>(p.kh * p.kw);
  if (p.o == 0 || p.cg == 0) return RTG_OK;
  if (p.depthwise) {
    void* d = nullptr;
    RTG_TRY(ctx->pool.alloc(static_cast<size_t>(p.c) * khw * 4, &d));
    const int total = static_cast<int>(p.c) * khw;
    transpose_dw_weights_kernel<<<(total + 255) / 256, 256, 0, ctx->stream>>>(static_cast<float*>(d), w_dev,
                                                                            static_cast<int>(p.c), khw);
    ctx->launches++;
    RTG_CUDA(cudaGetLastError());
    out->wmat = static_cast<float*>(d);
    out->kpt = 0;
    return RTG_OK;
  }
  if (p.pointwise && p.c % 4 == 0 && (reinterpret_cast<uintptr_t>(w_dev) & 15) == 0) {
    out->wmat = const_cast<float*>(w_dev);  // [O][C] is already the K-major B operand
    out->direct = true;
    out->kpt = static_cast<int>(p.c);
    return prepare_wpack(ctx, out, p.o, p.c);
  }
  if (conv_window_supported(p) && !(ctx->debug_flags & kForceIm2colFallback)) {
    const int cp = static_cast<int>(div_ceil(p.c, 4) * 4);
    const int64_t total = p.o * p.kh * kBK;
    void* d = nullptr;
    RTG_TRY(ctx->pool.alloc(static_cast<size_t>(total) * 4, &d));
    const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(total, 256), 148 * 16));
    arrange_window_weights_kernel<<<blocks, 256, 0, ctx->stream>>>(static_cast<float*>(d), w_dev, static_cast<int>(p.o),
                                                                   static_cast<int>(p.c), static_cast<int>(p.kh),
                                                                   static_cast<int>(p.kw), cp);
    ctx->launches++;
    RTG_CUDA(cudaGetLastError());
    out->wmat = static_cast<float*>(d);
    out->kpt = kBK;
    out->window = true;
    return prepare_wpack(ctx, out, p.o, p.kh * kBK);
  }
  const int kpt = static_cast<int>(div_ceil(p.cg, kBK) * kBK);
  const int64_t total = p.o * khw * kpt;
  void* d = nullptr;
  RTG_TRY(ctx->pool.alloc(static_cast<size_t>(total) * 4, &d));
  const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(total, 256), 148 * 16));
  arrange_conv_weights_kernel<<<blocks, 256, 0, ctx->stream>>>(static_cast<float*>(d), w_dev, static_cast<int>(p.o),
                                                               static_cast<int>(p.cg), khw, kpt);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  out->wmat = static_cast<float*>(d);
  out->kpt = kpt;
  return prepare_wpack(ctx, out, p.o, static_cast<int64_t>(khw) * kpt);
}

void conv_release_weights(Context* ctx, ConvWeights* w) {
  if (w->wmat && !w->direct) ctx->pool.free(w->wmat);
  if (w->wpack.p) ctx->pool.free(const_cast<float*>(w->wpack.p));
  w->wmat = nullptr;
  w->wpack = BPack{};
}

// ---------------------------------------------------------------- forward
namespace {

// Is the (H, W) plane of a 4-D view addressable as one dimension of H*W elements with a single stride?
bool plane_collapsible(const rtg_tensor& t) { return t.shape[2] == 1 || t.strides[2] == t.shape[3] * t.strides[3]; }
int64_t plane_stride(const rtg_tensor& t) { return t.shape[3] == 1 && t.shape[2] > 1 ? t.strides[2] : t.strides[3]; }

// Gather channels [c_off, c_off+c) of x into a dense NHWC buffer with channel pitch cp (zero padded).
rtg_status to_nhwc(Context* ctx, const rtg_tensor& x, int c_off, int c, int cp, float* dst) {
  const int n = static_cast<int>(x.shape[0]);
  const int hw = static_cast<int>(x.shape[2] * x.shape[3]);
  if (n == 0 || hw == 0 || cp == 0) return RTG_OK;
  dim3 grid(static_cast<unsigned>(div_ceil(hw, 32)), static_cast<unsigned>(div_ceil(cp, 32)), n);
  if (grid.y > 65535 || grid.z > 65535) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "conv: batch or channels too large");
  nchw_to_nhwc_kernel<<<grid, dim3(32, 8), 0, ctx->stream>>>(dst, static_cast<const float*>(x.data), c, hw, cp,
                                                            x.strides[0], x.strides[1], plane_stride(x), c_off);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

rtg_status from_nhwc(Context* ctx, const float* src, int c, int cp, const rtg_tensor& out, int c_off) {
  const int n = static_cast<int>(out.shape[0]);
  const int hw = static_cast<int>(out.shape[2] * out.shape[3]);
  if (n == 0 || hw == 0 || c == 0) return RTG_OK;
  dim3 grid(static_cast<unsigned>(div_ceil(hw, 32)), static_cast<unsigned>(div_ceil(c, 32)), n);
  nhwc_to_nchw_kernel<<<grid, dim3(32, 8), 0, ctx->stream>>>(static_cast<float*>(out.data), src, c, hw, cp,
                                                            out.strides[0], out.strides[1], plane_stride(out), c_off);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

}  // namespace

// Tile shape of the TMA-staged depthwise kernel: 128 threads = (TW / 4 strips) x (CB / 4 channel vectors).  Picks the
// shape that wastes the fewest threads on channels beyond C and columns beyond OW, then cuts the rows into bands so
// that there are several CTAs per SM.
static rtg_status launch_depthwise3x3_tma(Context* ctx, float* out, const float* x, const float* wt, const float* bias,
                                          const DwGeom& g) {
  const int stride = g.stride_h;
  int best_tw = 0;
  double best_waste = 1e30;
  for (int tw = 64; tw >= 8; tw >>= 1) {
    const int cb = (kDwThreads / (tw / kDwTX)) * 4;
    const double waste = (static_cast<double>(div_ceil(g.c, cb)) * cb / g.c) * (static_cast<double>(div_ceil(g.ow, tw)) * tw / g.ow);
    if (waste < best_waste - 1e-9) best_waste = waste, best_tw = tw;
  }
  DwTile t = {};
  t.tw = best_tw;
  t.cvs = kDwThreads / (t.tw / kDwTX);
  const int cb = t.cvs * 4;
  t.iwb = (t.tw - 1) * stride + 3;
  t.box_bytes = t.iwb * cb * 4;
  t.row_bytes = static_cast<int>(div_ceil(t.box_bytes, 128) * 128);
  t.ch_blocks = static_cast<int>(div_ceil(g.c, cb));
  t.bands = static_cast<int>(div_ceil(g.ow, t.tw));
  const int64_t columns = static_cast<int64_t>(g.n) * t.ch_blocks * t.bands;
  // Row bands.  CTAs share an SM's issue slots and HBM bandwidth, so a launch takes about (jobs an SM is handed) x (rows a
  // job streams: its output rows x stride, 2 halo rows, ~3 rows' worth of start-up); 4 CTAs fit an SM (126 registers x
  // 128 threads; the ring is <= 51 KB).  Pick the band count with the smallest estimate, at least 4 output rows per job.
  const int smem_bytes = kDwRing * t.row_bytes;
  const int64_t per_sm = std::max<int64_t>(1, std::min<int64_t>(4, (227 * 1024) / (smem_bytes + 1024)));
  static const int rows_env = [] { const char* e = std::getenv("RTG_DW_ROWS"); return e ? std::atoi(e) : 0; }();
  static const int pred_env = [] { const char* e = std::getenv("RTG_DW_PRED"); return e ? std::atoi(e) : 0; }();
  t.force_pred = pred_env;
  t.one2 = 0x3f8000003f800000ull;
  int best_rows = g.oh;
  double best_cost = 1e300;
  for (int64_t rb = 1; rb <= g.oh; ++rb) {
    const int rows = static_cast<int>(div_ceil(g.oh, rb));  // the most even cut into rb bands
    if (rows < std::min(4, g.oh)) break;
    if (div_ceil(g.oh, rows) != rb) continue;  // the same cut as a smaller band count
    const int64_t jobs_r = columns * rb;
    const double rounds = static_cast<double>(div_ceil(jobs_r, static_cast<int64_t>(ctx->num_sms)));
    // fewer resident CTAs than an SM can hold leave the per-row latency exposed
    const double thin = jobs_r < ctx->num_sms * per_sm ? static_cast<double>(ctx->num_sms * per_sm) / jobs_r : 1.0;
    const double cost = rounds * (rows * stride + 2 + 3) * std::min(thin, 2.0);
    if (cost < best_cost - 1e-9) best_cost = cost, best_rows = rows;
  }
  t.rows_per_job = rows_env > 0 ? std::min(rows_env, g.oh) : best_rows;
  t.row_bands = static_cast<int>(div_ceil(g.oh, t.rows_per_job));
  const int64_t jobs = columns * t.row_bands;
  if (jobs > INT32_MAX) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "depthwise convolution: too many tiles");

  CUtensorMap map;
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(g.c), static_cast<cuuint64_t>(g.w), static_cast<cuuint64_t>(g.h),
                        static_cast<cuuint64_t>(g.n)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(g.cp_in) * 4, static_cast<cuuint64_t>(g.w) * g.cp_in * 4,
                           static_cast<cuuint64_t>(g.h) * g.w * g.cp_in * 4};
  cuuint32_t box[4] = {static_cast<cuuint32_t>(cb), static_cast<cuuint32_t>(t.iwb), 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = ctx->encode_tiled(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x), dims, strides, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_errorf(RTG_ERR_CUDA, "cuTensorMapEncodeTiled (depthwise input) failed (%d): dims [%d,%d,%d,%d] box [%d,%d]",
                      (int)r, g.c, g.w, g.h, g.n, cb, t.iwb);
  const int smem = smem_bytes;
  const int ac = g.act == ACT_NONE ? 0 : g.act == ACT_RELU ? 1 : g.act == ACT_CLIP ? 2 : 3;
  using Kern = void (*)(const CUtensorMap, float*, const float*, const float*, const DwGeom, const DwTile);
#define RTG_DW_ROW(S, A) {depthwise3x3_tma_kernel<S, A, 8>, depthwise3x3_tma_kernel<S, A, 16>, depthwise3x3_tma_kernel<S, A, 32>, depthwise3x3_tma_kernel<S, A, 64>}
  static const Kern kerns[2][4][4] = {{RTG_DW_ROW(1, 0), RTG_DW_ROW(1, 1), RTG_DW_ROW(1, 2), RTG_DW_ROW(1, 3)},
                                      {RTG_DW_ROW(2, 0), RTG_DW_ROW(2, 1), RTG_DW_ROW(2, 2), RTG_DW_ROW(2, 3)}};
#undef RTG_DW_ROW
  const int ci = t.cvs == 8 ? 0 : t.cvs == 16 ? 1 : t.cvs == 32 ? 2 : 3;
  Kern kern = kerns[stride - 1][ac][ci];
  static bool configured[2][4][4][16] = {};
  if (!configured[stride - 1][ac][ci][ctx->device & 15]) {
    RTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    configured[stride - 1][ac][ci][ctx->device & 15] = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(jobs));
  cfg.blockDim = dim3(kDwThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = ctx->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  RTG_CUDA(cudaLaunchKernelEx(&cfg, kern, map, out, wt, bias, g, t));
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

rtg_status conv_forward(Context* ctx, const ConvPlan& p, const rtg_tensor& x_in, const ConvWeights& w,
                        const float* bias, const rtg_tensor* residual_in, int act, float act_lo, float act_hi,
                        const rtg_tensor& out_in) {
  if (p.n == 0 || p.o == 0 || p.oh == 0 || p.ow == 0) return RTG_OK;
  if (p.n > INT32_MAX || p.c > INT32_MAX || p.h > INT32_MAX || p.w > INT32_MAX || p.o > INT32_MAX)
    return set_error(RTG_ERR_UNSUPPORTED_VALUE, "conv dims exceed i32");
  // 4-D views of input / output (1-D convolutions carry H = 1)
  rtg_tensor x = x_in, out = out_in;
  rtg_tensor res_store;
  const rtg_tensor* residual = residual_in;
  auto lift = [&](rtg_tensor& t) {
    if (t.ndim == 3) {
      t.ndim = 4;
      t.shape[3] = t.shape[2], t.strides[3] = t.strides[2];
      t.shape[2] = 1, t.strides[2] = t.shape[3] * t.strides[3];
    }
  };
  lift(x);
  lift(out);
  if (residual) {
    res_store = *residual;
    lift(res_store);
    residual = &res_store;
  }

  // ---- output layout: NHWC-dense and NCHW-dense are written directly by the kernels
  Scratch out_tmp, res_tmp;
  rtg_tensor out_final = out;
  const bool need_out_tmp = !is_nhwc_dense(out) && !is_contiguous(out);
  if (need_out_tmp) {
    RTG_TRY(out_tmp.alloc(ctx, static_cast<size_t>(numel(out)) * 4));
    out.data = out_tmp.p;
    set_contiguous_strides(out);
  }
  const bool out_nhwc = is_nhwc_dense(out);
  if (residual) {
    bool same = true;
    for (int i = 0; i < 4; ++i) same = same && (residual->shape[i] == out.shape[i]);
    if (!same) return set_error(RTG_ERR_INCOMPATIBLE_INPUT_SHAPES, "residual shape must match the convolution output");
    bool same_strides = true;
    for (int i = 0; i < 4; ++i) same_strides = same_strides && (out.shape[i] == 1 || residual->strides[i] == out.strides[i]);
    if (!same_strides) {
      RTG_TRY(res_tmp.alloc(ctx, static_cast<size_t>(numel(out)) * 4));
      rtg_tensor r2 = out;
      r2.data = res_tmp.p;
      RTG_TRY(copy_view(ctx, r2, *residual));
      res_store = r2;
      residual = &res_store;
    }
  }

  const int N = static_cast<int>(p.n), C = static_cast<int>(p.c), H = static_cast<int>(p.h), W = static_cast<int>(p.w);
  const int O = static_cast<int>(p.o), OH = static_cast<int>(p.oh), OW = static_cast<int>(p.ow);
  const int G = static_cast<int>(p.groups), CG = static_cast<int>(p.cg), OG = O / std::max(G, 1);

  // ---- input as dense NHWC.  Sources that are neither NHWC-dense nor plane-collapsible are made contiguous first.
  Scratch x_cont, x_nhwc;
  if (!is_nhwc_dense(x) && !plane_collapsible(x)) {
    RTG_TRY(x_cont.alloc(ctx, static_cast<size_t>(numel(x)) * 4));
    rtg_tensor xc = x;
    xc.data = x_cont.p;
    set_contiguous_strides(xc);
    RTG_TRY(copy_view(ctx, xc, x));
    x = xc;
  }

  if (p.depthwise) {
    if (residual) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "fused residual is not supported for depthwise convolution");
    const float* xp = static_cast<const float*>(x.data);
    int cp_in = C;
    if (!is_nhwc_dense(x)) {
      RTG_TRY(x_nhwc.alloc(ctx, static_cast<size_t>(N) * H * W * C * 4));
      RTG_TRY(to_nhwc(ctx, x, 0, C, C, x_nhwc.f32()));
      xp = x_nhwc.f32();
    }
    Scratch o_nhwc;
    float* op = static_cast<float*>(out.data);
    if (!out_nhwc) {
      RTG_TRY(o_nhwc.alloc(ctx, static_cast<size_t>(N) * OH * OW * C * 4));
      op = o_nhwc.f32();
    }
    DwGeom g = {N, H, W, C, cp_in, C, static_cast<int>(p.kh), static_cast<int>(p.kw), OH, OW,
                static_cast<int>(p.stride_h), static_cast<int>(p.stride_w), static_cast<int>(p.pads[0]),
                static_cast<int>(p.pads[1]), static_cast<int>(p.dil_h), static_cast<int>(p.dil_w), act, act_lo, act_hi};
    const bool vec4 = C % 4 == 0 && (reinterpret_cast<uintptr_t>(xp) & 15) == 0 &&
                      (reinterpret_cast<uintptr_t>(op) & 15) == 0 && (reinterpret_cast<uintptr_t>(w.wmat) & 15) == 0;
    const int64_t total = static_cast<int64_t>(N) * OH * OW * (vec4 ? C / 4 : C);
    const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(total, 256), static_cast<int64_t>(ctx->num_sms) * 32));
    const bool k3 = vec4 && p.kh == 3 && p.kw == 3 && p.dil_h == 1 && p.dil_w == 1 && p.stride_h == p.stride_w &&
                    (p.stride_h == 1 || p.stride_h == 2) && cp_in % 4 == 0;
    static const bool dw_tma = []() {
      const char* e = getenv("RTG_DW_TMA");  // developer switch: 0 selects the register-window kernel
      return !e || atoi(e) != 0;
    }();
    if (k3 && dw_tma && W + 2 < (1 << 20)) {
      RTG_TRY(launch_depthwise3x3_tma(ctx, op, xp, w.wmat, bias, g));
      ctx->last_kernel = p.stride_h == 1 ? "depthwise3x3_tma<s1>" : "depthwise3x3_tma<s2>";
      if (!out_nhwc) RTG_TRY(from_nhwc(ctx, op, C, C, out, 0));
      return RTG_OK;
    }
    if (k3) {
      // rows per thread: as many as still leave ~512 threads per SM to hide the load latency
      constexpr int kTX = 4;
      const int64_t strips = static_cast<int64_t>(N) * div_ceil(OW, kTX) * (C / 4);
      int ty = 8;
      while (ty > 1 && strips * div_ceil(OH, ty) < static_cast<int64_t>(ctx->num_sms) * 768) ty >>= 1;
      const int64_t threads = strips * div_ceil(OH, ty);
      const int b3 = static_cast<int>(std::min<int64_t>(div_ceil(threads, 128), static_cast<int64_t>(ctx->num_sms) * 64));
      if (p.stride_h == 1)
        depthwise3x3_nhwc_kernel<1, kTX><<<b3, 128, 0, ctx->stream>>>(op, xp, w.wmat, bias, g, ty);
      else
        depthwise3x3_nhwc_kernel<2, kTX><<<b3, 128, 0, ctx->stream>>>(op, xp, w.wmat, bias, g, ty);
    } else if (vec4) {
      depthwise_nhwc_kernel<4><<<blocks, 256, 0, ctx->stream>>>(op, xp, w.wmat, bias, g);
    } else {
      depthwise_nhwc_kernel<1><<<blocks, 256, 0, ctx->stream>>>(op, xp, w.wmat, bias, g);
    }
    ctx->launches++;
    ctx->last_kernel = k3 ? (p.stride_h == 1 ? "depthwise3x3_nhwc<s1>" : "depthwise3x3_nhwc<s2>")
                          : vec4 ? "depthwise_nhwc<4>" : "depthwise_nhwc<1>";
    RTG_CUDA(cudaGetLastError());
    if (!out_nhwc) RTG_TRY(from_nhwc(ctx, op, C, C, out, 0));
  } else {
    if (w.window) {
      // zero-padded NHWC image with channels rounded up to 4; every 32-float window stays inside its row
      const int cpw = static_cast<int>(div_ceil(C, 4) * 4);
      const int hp = H + static_cast<int>(p.pads[0] + p.pads[2]);
      const int wp = std::max(W + static_cast<int>(p.pads[1] + p.pads[3]),
                              (OW - 1) * static_cast<int>(p.stride_w) + kBK / cpw);
      Scratch x_pad;
      RTG_TRY(x_pad.alloc(ctx, static_cast<size_t>(N) * hp * wp * cpw * 4));
      const int64_t pixels = static_cast<int64_t>(N) * hp * wp;
      const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(pixels, 256), static_cast<int64_t>(ctx->num_sms) * 16));
      pad_to_window_layout_kernel<<<blocks, 256, 0, ctx->stream>>>(
          x_pad.f32(), static_cast<const float*>(x.data), N, C, H, W, cpw, hp, wp, static_cast<int>(p.pads[0]),
          static_cast<int>(p.pads[1]), x.strides[0], x.strides[1], x.strides[2], x.strides[3]);
      ctx->launches++;
      RTG_CUDA(cudaGetLastError());
      Epilogue ep;
      ep.bias = bias;
      ep.bias_mode = BIAS_COL;
      ep.act = act, ep.act_lo = act_lo, ep.act_hi = act_hi;
      OutMap dm = out_nhwc ? OutMap{O, 0, 0, 0} : OutMap{0, OH * OW, O, 0};
      if (residual) {
        ep.R = static_cast<const float*>(residual->data);
        ep.beta = 1.f;
        ep.rmap = dm;
      }
      WindowGeom wg = {N, hp, wp, cpw, static_cast<int>(p.kh), static_cast<int>(p.kw), OH, OW,
                       static_cast<int>(p.stride_h), static_cast<int>(p.stride_w), static_cast<int>(p.dil_h), C};
      ep.b_pack = w.wpack, ep.b_pack_const = w.persistent;
      RTG_TRY(conv_window_gemm_device(ctx, x_pad.f32(), wg, w.wmat, O, static_cast<float*>(out.data), dm, ep));
      if (need_out_tmp) RTG_TRY(copy_view(ctx, out_final, out));
      return RTG_OK;
    }
    // Groups whose channel slice is not 16-byte aligned in the shared NHWC tensor get their own gathered copy.
    const bool per_group = G > 1 && (CG % 4 != 0);
    const int c_src = per_group ? CG : C;
    const int cp = static_cast<int>(div_ceil(c_src, 4) * 4);
    const bool direct_x = !per_group && is_nhwc_dense(x) && cp == C && (reinterpret_cast<uintptr_t>(x.data) & 15) == 0;
    const float* xp = static_cast<const float*>(x.data);
    if (!direct_x) {
      RTG_TRY(x_nhwc.alloc(ctx, static_cast<size_t>(N) * H * W * cp * 4));
      xp = x_nhwc.f32();
      if (!per_group) RTG_TRY(to_nhwc(ctx, x, 0, C, cp, x_nhwc.f32()));
    }
    const int64_t M = static_cast<int64_t>(N) * OH * OW;
    const int khw = static_cast<int>(p.kh * p.kw);
    for (int g = 0; g < G; ++g) {
      if (per_group) RTG_TRY(to_nhwc(ctx, x, g * CG, CG, cp, x_nhwc.f32()));
      Epilogue ep;
      ep.bias = bias ? bias + g * OG : nullptr;
      ep.bias_mode = BIAS_COL;
      ep.act = act, ep.act_lo = act_lo, ep.act_hi = act_hi;
      OutMap dm;
      if (out_nhwc) dm = OutMap{O, 0, 0, g * OG};
      else dm = OutMap{0, OH * OW, O, g * OG};
      if (residual) {
        ep.R = static_cast<const float*>(residual->data);
        ep.beta = 1.f;
        ep.rmap = dm;
      }
      float* D = static_cast<float*>(out.data);
      if (w.direct) {
        // pointwise: A = activations [M, C] (pitch cp), B = W [O][C]
        ep.b_pack = w.wpack, ep.b_pack_const = w.persistent;
        RTG_TRY(gemm_device(ctx, MatView{xp, M, C, cp, 1}, MatView{w.wmat, C, O, 1, C}, D, dm, ep));
        continue;
      }
      const int64_t kmat = static_cast<int64_t>(khw) * w.kpt;
      const float* wg = w.wmat + static_cast<int64_t>(g) * OG * kmat;
      ConvGeom cg = {N, H, W, cp, per_group ? 0 : g * CG, CG, static_cast<int>(p.kh), static_cast<int>(p.kw), OH, OW,
                     static_cast<int>(p.stride_h), static_cast<int>(p.stride_w), static_cast<int>(p.pads[0]),
                     static_cast<int>(p.pads[1]), static_cast<int>(p.pads[2]), static_cast<int>(p.pads[3]),
                     static_cast<int>(p.dil_h), static_cast<int>(p.dil_w)};
      if (conv_im2col_supported(cg) && !(ctx->debug_flags & kForceIm2colFallback)) {
        ep.b_pack = w.wpack, ep.b_pack_const = w.persistent;  // rows [g * OG, +OG) of both planes
        if (ep.b_pack.p) ep.b_pack.p += static_cast<int64_t>(g) * OG * w.wpack.ld;
        RTG_TRY(conv_gemm_device(ctx, xp, cg, wg, OG, D, dm, ep));
        ep.b_pack = BPack{};
      } else {
        // outside the TMA im2col limits: materialise the patch matrix, then a plain GEMM
        Scratch cols;
        RTG_TRY(cols.alloc(ctx, static_cast<size_t>(M) * kmat * 4));
        Im2colGeom ig = {N, H, W, cp, cg.c_off, CG, cg.kh, cg.kw, OH, OW, cg.stride_h, cg.stride_w,
                         cg.pad_t, cg.pad_l, cg.dil_h, cg.dil_w, w.kpt};
        const int64_t total = M * kmat;
        const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(total, 256), static_cast<int64_t>(ctx->num_sms) * 32));
        im2col_nhwc_kernel<<<blocks, 256, 0, ctx->stream>>>(cols.f32(), xp, ig);
        ctx->launches++;
        RTG_CUDA(cudaGetLastError());
        RTG_TRY(gemm_device(ctx, MatView{cols.f32(), M, kmat, kmat, 1}, MatView{wg, kmat, OG, 1, kmat}, D, dm, ep));
      }
    }
  }
  if (need_out_tmp) RTG_TRY(copy_view(ctx, out_final, out));
  return RTG_OK;
}

rtg_status conv_patches(Context* ctx, int kind, const ConvPlan& p, const rtg_tensor& x_in, int group, float* dst,
                        size_t dst_floats, int64_t dims[4]) {
  rtg_tensor x = x_in;
  if (x.ndim == 3) {  // 1-D convolution: H = 1
    x.ndim = 4;
    x.shape[3] = x.shape[2], x.strides[3] = x.strides[2];
    x.shape[2] = 1, x.strides[2] = x.shape[3] * x.strides[3];
  }
  const int N = static_cast<int>(p.n), C = static_cast<int>(p.c), H = static_cast<int>(p.h), W = static_cast<int>(p.w);
  const int OH = static_cast<int>(p.oh), OW = static_cast<int>(p.ow), CG = static_cast<int>(p.cg);
  if (group < 0 || group >= p.groups) return set_error(RTG_ERR_INVALID_VALUE, "group out of range");
  Scratch x_cont;
  if (!is_nhwc_dense(x) && !plane_collapsible(x)) {
    RTG_TRY(x_cont.alloc(ctx, static_cast<size_t>(numel(x)) * 4));
    rtg_tensor xc = x;
    xc.data = x_cont.p;
    set_contiguous_strides(xc);
    RTG_TRY(copy_view(ctx, xc, x));
    x = xc;
  }
  auto need = [&](int64_t a, int64_t b, int64_t c, int64_t d) {
    dims[0] = a, dims[1] = b, dims[2] = c, dims[3] = d;
    return static_cast<size_t>(a * b * c * d) <= dst_floats
               ? RTG_OK
               : set_error(RTG_ERR_INVALID_VALUE, "destination buffer too small");
  };
  const int cp = static_cast<int>(div_ceil(C, 4) * 4);
  if (kind == 1) {
    const int hp = H + static_cast<int>(p.pads[0] + p.pads[2]);
    const int wp = std::max(W + static_cast<int>(p.pads[1] + p.pads[3]), (OW - 1) * static_cast<int>(p.stride_w) + kBK / cp);
    RTG_TRY(need(N, hp, wp, cp));
    const int64_t pixels = static_cast<int64_t>(N) * hp * wp;
    if (pixels == 0) return RTG_OK;
    const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(pixels, 256), static_cast<int64_t>(ctx->num_sms) * 16));
    pad_to_window_layout_kernel<<<blocks, 256, 0, ctx->stream>>>(dst, static_cast<const float*>(x.data), N, C, H, W, cp,
                                                                 hp, wp, static_cast<int>(p.pads[0]),
                                                                 static_cast<int>(p.pads[1]), x.strides[0], x.strides[1],
                                                                 x.strides[2], x.strides[3]);
    ctx->launches++;
    RTG_CUDA(cudaGetLastError());
    return RTG_OK;
  }
  if (kind == 2) {
    RTG_TRY(need(N, H * W, cp, 1));
    return to_nhwc(ctx, x, 0, C, cp, dst);
  }
  Scratch nhwc;
  RTG_TRY(nhwc.alloc(ctx, static_cast<size_t>(std::max<int64_t>(1, static_cast<int64_t>(N) * H * W * cp)) * 4));
  RTG_TRY(to_nhwc(ctx, x, 0, C, cp, nhwc.f32()));
  if (kind == 3) {
    RTG_TRY(need(N, C, H, W));
    rtg_tensor o = x;
    o.data = dst;
    set_contiguous_strides(o);
    return from_nhwc(ctx, nhwc.f32(), C, cp, o, 0);
  }
  if (kind != 0) return set_error(RTG_ERR_INVALID_VALUE, "unknown patch kind");
  const int kpt = static_cast<int>(div_ceil(CG, kBK) * kBK);
  const int64_t kmat = p.kh * p.kw * kpt, M = static_cast<int64_t>(N) * OH * OW;
  RTG_TRY(need(N, static_cast<int64_t>(OH) * OW, kmat, 1));
  if (M * kmat == 0) return RTG_OK;
  Im2colGeom ig = {N, H, W, cp, group * CG, CG, static_cast<int>(p.kh), static_cast<int>(p.kw), OH, OW,
                   static_cast<int>(p.stride_h), static_cast<int>(p.stride_w), static_cast<int>(p.pads[0]),
                   static_cast<int>(p.pads[1]), static_cast<int>(p.dil_h), static_cast<int>(p.dil_w), kpt};
  const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(M * kmat, 256), static_cast<int64_t>(ctx->num_sms) * 32));
  im2col_nhwc_kernel<<<blocks, 256, 0, ctx->stream>>>(dst, nhwc.f32(), ig);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

}  // namespace rtg
