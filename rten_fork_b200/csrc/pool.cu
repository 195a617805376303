// MaxPool / AveragePool / GlobalAveragePool (reference: src/ops/pooling.rs:104-420).  HBM-bound CUDA-core kernels that
// address input and output through element strides, so NCHW-dense and NHWC-dense tensors are both handled without a
// layout change; the thread order follows whichever dimension is contiguous.
#include "ops.h"

namespace rtg {

namespace {

struct PoolGeom {
  int n, c, h, w, oh, ow;
  int kh, kw, stride_h, stride_w, pad_t, pad_l;
  int64_t is[4], os[4];   // element strides of input / output in (n, c, h, w) order
  int kind, count_include_pad;
  int chan_fastest;       // 1: thread index varies fastest over channels (NHWC), 0: over ow (NCHW)
};

__global__ void pool2d_kernel(float* __restrict__ out, const float* __restrict__ in, PoolGeom g) {
  const int64_t total = static_cast<int64_t>(g.n) * g.c * g.oh * g.ow;
  const int64_t step = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += step) {
    int n, c, oy, ox;
    int64_t r = i;
    if (g.chan_fastest) {
      c = static_cast<int>(r % g.c), r /= g.c;
      ox = static_cast<int>(r % g.ow), r /= g.ow;
      oy = static_cast<int>(r % g.oh), r /= g.oh;
      n = static_cast<int>(r);
    } else {
      ox = static_cast<int>(r % g.ow), r /= g.ow;
      oy = static_cast<int>(r % g.oh), r /= g.oh;
      c = static_cast<int>(r % g.c), r /= g.c;
      n = static_cast<int>(r);
    }
    const float* ip = in + n * g.is[0] + c * g.is[1];
    // fold in (ky, kx) order over the non-padding window, as pool_chans does (src/ops/pooling.rs:150-176)
    float acc = g.kind == RTG_POOL_MAX ? __int_as_float(0xff800000) : 0.f;
    int count = 0;
    for (int ky = 0; ky < g.kh; ++ky) {
      const int iy = oy * g.stride_h + ky - g.pad_t;
      if (iy < 0 || iy >= g.h) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int ix = ox * g.stride_w + kx - g.pad_l;
        if (ix < 0 || ix >= g.w) continue;
        const float v = ip[iy * g.is[2] + ix * g.is[3]];
        acc = g.kind == RTG_POOL_MAX ? fmaxf(acc, v) : __fadd_rn(acc, v);
        ++count;
      }
    }
    if (g.kind == RTG_POOL_AVG)
      acc = __fdiv_rn(acc, static_cast<float>(g.count_include_pad ? g.kh * g.kw : count));
    out[n * g.os[0] + c * g.os[1] + oy * g.os[2] + ox * g.os[3]] = acc;
  }
}

// NHWC-dense fast path: one thread per (pixel, 4 channels), float4 loads coalesced over channels, 32-bit index math.
// Same fold order per element as pool2d_kernel, so results are bit-identical.
__global__ void pool2d_nhwc4_kernel(float* __restrict__ out, const float* __restrict__ in, PoolGeom g) {
  const int c4 = g.c >> 2;
  const int total = g.n * g.oh * g.ow * c4;  // host guarantees < 2^31
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int r = i;
    const int cq = r % c4;
    r /= c4;
    const int ox = r % g.ow;
    r /= g.ow;
    const int oy = r % g.oh;
    const int n = r / g.oh;
    const float4* ip = reinterpret_cast<const float4*>(in + static_cast<int64_t>(n) * g.h * g.w * g.c) + cq;
    const float ninf = __int_as_float(0xff800000);
    float4 acc = g.kind == RTG_POOL_MAX ? make_float4(ninf, ninf, ninf, ninf) : make_float4(0.f, 0.f, 0.f, 0.f);
    int count = 0;
    for (int ky = 0; ky < g.kh; ++ky) {
      const int iy = oy * g.stride_h + ky - g.pad_t;
      if (iy < 0 || iy >= g.h) continue;
      for (int kx = 0; kx < g.kw; ++kx) {
        const int ix = ox * g.stride_w + kx - g.pad_l;
        if (ix < 0 || ix >= g.w) continue;
        const float4 v = __ldg(ip + (iy * g.w + ix) * c4);
        if (g.kind == RTG_POOL_MAX) {
          acc.x = fmaxf(acc.x, v.x), acc.y = fmaxf(acc.y, v.y), acc.z = fmaxf(acc.z, v.z), acc.w = fmaxf(acc.w, v.w);
        } else {
          acc.x = __fadd_rn(acc.x, v.x), acc.y = __fadd_rn(acc.y, v.y);
          acc.z = __fadd_rn(acc.z, v.z), acc.w = __fadd_rn(acc.w, v.w);
        }
        ++count;
      }
    }
    if (g.kind == RTG_POOL_AVG) {
      const float d = static_cast<float>(g.count_include_pad ? g.kh * g.kw : count);
      acc.x = __fdiv_rn(acc.x, d), acc.y = __fdiv_rn(acc.y, d), acc.z = __fdiv_rn(acc.z, d), acc.w = __fdiv_rn(acc.w, d);
    }
    reinterpret_cast<float4*>(out)[i] = acc;
  }
}

// MaxPool over NHWC with a compile-time window: the window loops of pool2d_nhwc4_kernel are data-dependent (`continue`
// on padding), so its loads leave one after the other; here all KH x KW loads of a thread are in flight together
// (out-of-image taps are replaced by -inf, which a max ignores exactly like skipping them).
template <int KH, int KW>
__global__ void __launch_bounds__(256) maxpool_nhwc4_kernel(float* __restrict__ out, const float* __restrict__ in, PoolGeom g) {
  const int c4 = g.c >> 2;
  const int total = g.n * g.oh * g.ow * c4;  // host guarantees < 2^31
  const float ninf = __int_as_float(0xff800000);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int r = i;
    const int cq = r % c4;
    r /= c4;
    const int ox = r % g.ow;
    r /= g.ow;
    const int oy = r % g.oh;
    const int n = r / g.oh;
    const float4* ip = reinterpret_cast<const float4*>(in + static_cast<int64_t>(n) * g.h * g.w * g.c) + cq;
    const int iy0 = oy * g.stride_h - g.pad_t, ix0 = ox * g.stride_w - g.pad_l;
    float4 v[KH * KW];
#pragma unroll
    for (int ky = 0; ky < KH; ++ky) {
#pragma unroll
      for (int kx = 0; kx < KW; ++kx) {
        const int iy = iy0 + ky, ix = ix0 + kx;
        const bool ok = iy >= 0 && iy < g.h && ix >= 0 && ix < g.w;
        v[ky * KW + kx] = ok ? __ldg(ip + (iy * g.w + ix) * c4) : make_float4(ninf, ninf, ninf, ninf);
      }
    }
    float4 acc = make_float4(ninf, ninf, ninf, ninf);
#pragma unroll
    for (int k = 0; k < KH * KW; ++k) {  // same fold order as pool2d_kernel: (ky, kx)
      acc.x = fmaxf(acc.x, v[k].x), acc.y = fmaxf(acc.y, v[k].y), acc.z = fmaxf(acc.z, v[k].z), acc.w = fmaxf(acc.w, v[k].w);
    }
    reinterpret_cast<float4*>(out)[i] = acc;
  }
}

// The same with two output rows (oy, oy + 1) per thread: their windows share KH - SH input rows, so a thread issues
// (KH + SH) x KW loads for two results instead of 2 x KH x KW, and the rows both windows read cross L1 / L2 once.
// Each result folds its own window in (ky, kx) order: identical bits.
template <int KH, int KW, int SH>
__global__ void __launch_bounds__(256) maxpool_nhwc4_rows2_kernel(float* __restrict__ out, const float* __restrict__ in, PoolGeom g) {
  constexpr int R = KH + SH;
  const int c4 = g.c >> 2;
  const int ohp = (g.oh + 1) >> 1;
  const int total = g.n * ohp * g.ow * c4;
  const float ninf = __int_as_float(0xff800000);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int r = i;
    const int cq = r % c4;
    r /= c4;
    const int ox = r % g.ow;
    r /= g.ow;
    const int oy = (r % ohp) * 2;
    const int n = r / ohp;
    const bool two = oy + 1 < g.oh;
    const float4* ip = reinterpret_cast<const float4*>(in + static_cast<int64_t>(n) * g.h * g.w * g.c) + cq;
    const int iy0 = oy * SH - g.pad_t, ix0 = ox * g.stride_w - g.pad_l;
    float4 v[R][KW];
#pragma unroll
    for (int ky = 0; ky < R; ++ky) {
#pragma unroll
      for (int kx = 0; kx < KW; ++kx) {
        const int iy = iy0 + ky, ix = ix0 + kx;
        const bool ok = iy >= 0 && iy < g.h && ix >= 0 && ix < g.w && (ky < KH || two);
        v[ky][kx] = ok ? __ldg(ip + (iy * g.w + ix) * c4) : make_float4(ninf, ninf, ninf, ninf);
      }
    }
    float4* op = reinterpret_cast<float4*>(out) + ((static_cast<int64_t>(n) * g.oh + oy) * g.ow + ox) * c4 + cq;
#pragma unroll
    for (int o = 0; o < 2; ++o) {
      float4 acc = make_float4(ninf, ninf, ninf, ninf);
#pragma unroll
      for (int ky = 0; ky < KH; ++ky) {
#pragma unroll
        for (int kx = 0; kx < KW; ++kx) {
          const float4 x = v[o * SH + ky][kx];
          acc.x = fmaxf(acc.x, x.x), acc.y = fmaxf(acc.y, x.y), acc.z = fmaxf(acc.z, x.z), acc.w = fmaxf(acc.w, x.w);
        }
      }
      if (o == 0 || two) op[static_cast<int64_t>(o) * g.ow * c4] = acc;
    }
  }
}

// One warp per (n, c) for NCHW (contiguous plane, shuffle reduction); for NHWC one thread per (n, c) column walks the
// pixels so that a warp reads 32 consecutive channels per pixel (coalesced).
__global__ void gap_nchw_kernel(float* __restrict__ out, const float* __restrict__ in, int64_t nc, int hw, int c,
                                int64_t sn, int64_t sc, int64_t sp, int64_t on, int64_t oc) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t i = warp; i < nc; i += nwarps) {
    const int64_t n = i / c, ch = i % c;
    const float* p = in + n * sn + ch * sc;
    float s = 0.f;
    for (int j = lane; j < hw; j += 32) s += p[j * sp];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[n * on + ch * oc] = __fdiv_rn(s, static_cast<float>(hw));
  }
}
__global__ void gap_nhwc_kernel(float* __restrict__ out, const float* __restrict__ in, int n, int hw, int c,
                                int64_t on, int64_t oc) {
  // grid: (ceil(c/32), n); block: 32 x 8 -> 8 pixel-slices per channel, reduced through shared memory
  __shared__ float part[8][33];
  const int ch = blockIdx.x * 32 + threadIdx.x;
  const int img = blockIdx.y;
  float s = 0.f;
  if (ch < c) {
    const float* p = in + static_cast<int64_t>(img) * hw * c + ch;
    for (int j = threadIdx.y; j < hw; j += 8) s += p[static_cast<int64_t>(j) * c];
  }
  part[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && ch < c) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += part[k][threadIdx.x];
    out[img * on + ch * oc] = __fdiv_rn(t, static_cast<float>(hw));
  }
}

}  // namespace

rtg_status pool_out_shape(const rtg_tensor* in, const rtg_pool_attrs* a, int64_t* shape, int64_t* pads_out) {
  if (!in) return set_error(RTG_ERR_MISSING_INPUTS, "Required inputs were missing");
  if (in->dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  if (in->ndim != 4) return set_error(RTG_ERR_INVALID_VALUE, "input must have 4 dims (NCHW)");
  const int64_t h = in->shape[2], w = in->shape[3];
  if (a->strides[0] <= 0 || a->strides[1] <= 0) return set_error(RTG_ERR_INVALID_VALUE, "Strides must be > 0");
  if (a->kernel[0] <= 0 || a->kernel[1] <= 0) return set_error(RTG_ERR_PANIC, "kernel size must be > 0");
  int64_t pads[4] = {0, 0, 0, 0}, oh, ow;
  if (a->pad_mode == RTG_PAD_SAME) {
    oh = div_ceil(h, a->strides[0]), ow = div_ceil(w, a->strides[1]);
    if (oh == 0 || ow == 0) return set_error(RTG_ERR_PANIC, "attempt to subtract with overflow");
    const int64_t th = std::max<int64_t>(0, (oh - 1) * a->strides[0] + a->kernel[0] - h);
    const int64_t tw = std::max<int64_t>(0, (ow - 1) * a->strides[1] + a->kernel[1] - w);
    pads[0] = th / 2, pads[1] = tw / 2, pads[2] = div_ceil(th, 2), pads[3] = div_ceil(tw, 2);
  } else {
    for (int i = 0; i < 4; ++i) pads[i] = a->pads[i];
    const int64_t ph = h + pads[0] + pads[2], pw = w + pads[1] + pads[3];
    if (ph < a->kernel[0] || pw < a->kernel[1]) return set_error(RTG_ERR_INVALID_VALUE, "Input too small for kernel size");
    oh = (ph - a->kernel[0]) / a->strides[0] + 1;
    ow = (pw - a->kernel[1]) / a->strides[1] + 1;
  }
  shape[0] = in->shape[0], shape[1] = in->shape[1], shape[2] = oh, shape[3] = ow;
  if (pads_out)
    for (int i = 0; i < 4; ++i) pads_out[i] = pads[i];
  return RTG_OK;
}

rtg_status pool2d_forward(Context* ctx, int kind, const rtg_tensor& in, const rtg_pool_attrs& a, const rtg_tensor& out) {
  int64_t shape[4], pads[4];
  RTG_TRY(pool_out_shape(&in, &a, shape, pads));
  const int64_t total = shape[0] * shape[1] * shape[2] * shape[3];
  if (total == 0) return RTG_OK;
  PoolGeom g = {};
  g.n = static_cast<int>(shape[0]), g.c = static_cast<int>(shape[1]), g.h = static_cast<int>(in.shape[2]);
  g.w = static_cast<int>(in.shape[3]), g.oh = static_cast<int>(shape[2]), g.ow = static_cast<int>(shape[3]);
  g.kh = static_cast<int>(a.kernel[0]), g.kw = static_cast<int>(a.kernel[1]);
  g.stride_h = static_cast<int>(a.strides[0]), g.stride_w = static_cast<int>(a.strides[1]);
  g.pad_t = static_cast<int>(pads[0]), g.pad_l = static_cast<int>(pads[1]);
  for (int i = 0; i < 4; ++i) g.is[i] = in.strides[i], g.os[i] = out.strides[i];
  g.kind = kind, g.count_include_pad = a.count_include_pad;
  g.chan_fastest = (out.strides[1] == 1 && g.c > 1) ? 1 : 0;
  if (is_nhwc_dense(in) && is_nhwc_dense(out) && g.c % 4 == 0 && total < (1ll << 31) &&
      (reinterpret_cast<uintptr_t>(in.data) & 15) == 0 && (reinterpret_cast<uintptr_t>(out.data) & 15) == 0) {
    const int blocks4 = static_cast<int>(std::min<int64_t>(div_ceil(total / 4, 256), static_cast<int64_t>(ctx->num_sms) * 32));
    float* op = static_cast<float*>(out.data);
    const float* ip = static_cast<const float*>(in.data);
    static const bool rows2 = [] { const char* e = std::getenv("RTG_POOL_ROWS2"); return !e || std::atoi(e) != 0; }();
    const int64_t pairs4 = shape[0] * ((shape[2] + 1) / 2) * shape[3] * (shape[1] / 4);
    const int blocks_r2 = static_cast<int>(std::min<int64_t>(div_ceil(pairs4, 256), static_cast<int64_t>(ctx->num_sms) * 64));
    if (rows2 && kind == RTG_POOL_MAX && g.kh == 3 && g.kw == 3 && g.stride_h == 2) {
      maxpool_nhwc4_rows2_kernel<3, 3, 2><<<blocks_r2, 256, 0, ctx->stream>>>(op, ip, g);
      ctx->last_kernel = "maxpool_nhwc4<3x3,two rows>";
    } else if (rows2 && kind == RTG_POOL_MAX && g.kh == 3 && g.kw == 3 && g.stride_h == 1) {
      maxpool_nhwc4_rows2_kernel<3, 3, 1><<<blocks_r2, 256, 0, ctx->stream>>>(op, ip, g);
      ctx->last_kernel = "maxpool_nhwc4<3x3,two rows>";
    } else if (kind == RTG_POOL_MAX && g.kh == 3 && g.kw == 3) {
      maxpool_nhwc4_kernel<3, 3><<<blocks4, 256, 0, ctx->stream>>>(op, ip, g);
      ctx->last_kernel = "maxpool_nhwc4<3x3>";
    } else if (kind == RTG_POOL_MAX && g.kh == 2 && g.kw == 2) {
      maxpool_nhwc4_kernel<2, 2><<<blocks4, 256, 0, ctx->stream>>>(op, ip, g);
      ctx->last_kernel = "maxpool_nhwc4<2x2>";
    } else {
      pool2d_nhwc4_kernel<<<blocks4, 256, 0, ctx->stream>>>(op, ip, g);
      ctx->last_kernel = "pool2d_nhwc4";
    }
    ctx->launches++;
    RTG_CUDA(cudaGetLastError());
    return RTG_OK;
  }
  const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(total, 256), static_cast<int64_t>(ctx->num_sms) * 16));
  pool2d_kernel<<<blocks, 256, 0, ctx->stream>>>(static_cast<float*>(out.data), static_cast<const float*>(in.data), g);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

rtg_status global_avg_pool_forward(Context* ctx, const rtg_tensor& in, const rtg_tensor& out) {
  if (in.dtype != RTG_F32) return set_error(RTG_ERR_INCORRECT_INPUT_TYPE, "Incorrect input type");
  if (in.ndim != 4) return set_error(RTG_ERR_INVALID_VALUE, "input must have 4 dims (NCHW)");
  const int64_t n = in.shape[0], c = in.shape[1];
  const int hw = static_cast<int>(in.shape[2] * in.shape[3]);
  if (n * c == 0) return RTG_OK;
  if (is_nhwc_dense(in) && c > 1) {
    dim3 grid(static_cast<unsigned>(div_ceil(c, 32)), static_cast<unsigned>(n));
    gap_nhwc_kernel<<<grid, dim3(32, 8), 0, ctx->stream>>>(static_cast<float*>(out.data),
                                                          static_cast<const float*>(in.data), static_cast<int>(n), hw,
                                                          static_cast<int>(c), out.strides[0], out.strides[1]);
  } else {
    if (!(in.shape[2] == 1 || in.strides[2] == in.shape[3] * in.strides[3]))
      return set_error(RTG_ERR_UNSUPPORTED_VALUE, "GlobalAveragePool input plane must be collapsible");
    const int64_t sp = (in.shape[3] == 1 && in.shape[2] > 1) ? in.strides[2] : in.strides[3];
    const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(n * c * 32, 256), static_cast<int64_t>(ctx->num_sms) * 16));
    gap_nchw_kernel<<<blocks, 256, 0, ctx->stream>>>(static_cast<float*>(out.data), static_cast<const float*>(in.data),
                                                    n * c, hw, static_cast<int>(c), in.strides[0], in.strides[1], sp,
                                                    out.strides[0], out.strides[1]);
  }
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

}  // namespace rtg
