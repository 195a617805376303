// Host side of the tcgen05 contraction engine: operand legalisation, TMA descriptor encoding, tile-shape choice,
// launch.  Reference role: GemmExecutor::{gemm_bias, prepack_*} front-end (src/gemm.rs:448-542).
#include "gemm_launch.cuh"
#include <vector>

#include <math.h>
#include <stdio.h>
#include <stdlib.h>

namespace rtg {

// ---------------------------------------------------------------- helper kernels
// Tiled strided copy: coalesced on whichever side is contiguous.  32x32 tile through shared memory.
__global__ void copy_strided_kernel(float* __restrict__ dst, int64_t dst_ld, const float* __restrict__ src,
                                    int64_t rows, int64_t cols, int64_t rs, int64_t cs) {
  __shared__ float tile[32][33];
  const int64_t r0 = static_cast<int64_t>(blockIdx.y) * 32, c0 = static_cast<int64_t>(blockIdx.x) * 32;
  const int tx = threadIdx.x, ty = threadIdx.y;  // 32 x 8
  if (cs == 1 || rs != 1) {
    // source rows contiguous (or nothing contiguous): read along columns
    for (int i = ty; i < 32; i += 8) {
      const int64_t r = r0 + i, c = c0 + tx;
      if (r < rows && c < cols) dst[r * dst_ld + c] = src[r * rs + c * cs];
    }
  } else {
    // source columns contiguous (rs == 1): read along rows, transpose through smem
    for (int i = ty; i < 32; i += 8) {
      const int64_t c = c0 + i, r = r0 + tx;
      if (r < rows && c < cols) tile[i][tx] = src[r * rs + c * cs];
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
      const int64_t r = r0 + i, c = c0 + tx;
      if (r < rows && c < cols) dst[r * dst_ld + c] = tile[tx][i];
    }
  }
}

rtg_status copy_strided_2d(Context* ctx, float* dst, int64_t dst_ld, const float* src, int64_t rows, int64_t cols,
                           int64_t rs, int64_t cs) {
  if (rows == 0 || cols == 0) return RTG_OK;
  dim3 grid(static_cast<unsigned>(div_ceil(cols, 32)), static_cast<unsigned>(div_ceil(rows, 32)));
  if (grid.y > 65535) return set_error(RTG_ERR_INTERNAL, "copy_strided_2d: too many rows");
  copy_strided_kernel<<<grid, dim3(32, 8), 0, ctx->stream>>>(dst, dst_ld, src, rows, cols, rs, cs);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

__global__ void scale_kernel(float* p, size_t n, float beta) {
  size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t step = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (; i < n; i += step) p[i] = beta == 0.f ? 0.f : beta * p[i];
}
rtg_status scale_buffer(Context* ctx, float* p, size_t n, float beta) {
  if (n == 0) return RTG_OK;
  const int blocks = static_cast<int>(std::min<size_t>((n + 255) / 256, 148 * 8));
  scale_kernel<<<blocks, 256, 0, ctx->stream>>>(p, n, beta);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

// ---------------------------------------------------------------- packed B operand of RTG_MATH_TF32_BF16X2 (BPack)
__global__ void b_pack_kernel(float* __restrict__ out_hi, uint32_t* __restrict__ out_prime, const float* __restrict__ b,
                              int64_t K, int64_t N, int64_t rs, int64_t cs, int64_t groups /* ceil(K/32)*4 */,
                              int n_fastest) {
  const int64_t total = N * groups;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    // consecutive threads walk the contiguous dimension of B
    const int64_t n = n_fastest ? i % N : i / groups, g = n_fastest ? i / N : i % groups;
    float hi[8], lo[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int64_t k = g * 8 + j;
      const float x = k < K ? b[k * rs + n * cs] : 0.f;
      hi[j] = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);  // round to nearest TF32
      lo[j] = x - hi[j];                                                        // exact in fp32
    }
    float4* oh = reinterpret_cast<float4*>(out_hi + (n * groups + g) * 8);
    oh[0] = make_float4(hi[0], hi[1], hi[2], hi[3]);
    oh[1] = make_float4(hi[4], hi[5], hi[6], hi[7]);
    uint4 hi4, lo4;
    hi4.x = ptx::pack_bf16x2(hi[0], hi[1]), hi4.y = ptx::pack_bf16x2(hi[2], hi[3]);
    hi4.z = ptx::pack_bf16x2(hi[4], hi[5]), hi4.w = ptx::pack_bf16x2(hi[6], hi[7]);
    lo4.x = ptx::pack_bf16x2(lo[0], lo[1]), lo4.y = ptx::pack_bf16x2(lo[2], lo[3]);
    lo4.z = ptx::pack_bf16x2(lo[4], lo[5]), lo4.w = ptx::pack_bf16x2(lo[6], lo[7]);
    uint4* o = reinterpret_cast<uint4*>(out_prime + (n * groups + g) * 8);
    o[0] = hi4;
    o[1] = lo4;
  }
}

rtg_status make_b_pack(Context* ctx, MatView b, float* out, int64_t plane) {
  const int64_t K = b.rows, N = b.cols;
  if (K == 0 || N == 0) return RTG_OK;
  const int64_t groups = div_ceil(K, kBK) * 4;
  const int64_t total = N * groups;
  const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(total, 256), static_cast<int64_t>(ctx->num_sms) * 16));
  b_pack_kernel<<<blocks, 256, 0, ctx->stream>>>(out, reinterpret_cast<uint32_t*>(out + plane), b.p, K, N, b.rs, b.cs,
                                                 groups, b.cs == 1 && b.rs != 1 ? 1 : 0);
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

// ---------------------------------------------------------------- tensor maps
static rtg_status encode_2d(Context* ctx, CUtensorMap* map, const float* base, uint64_t inner, uint64_t outer,
                            uint64_t outer_stride_elems, uint32_t box_inner, uint32_t box_outer,
                            CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {outer_stride_elems * sizeof(float)};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = ctx->encode_tiled(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box,
                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_errorf(RTG_ERR_CUDA,
                      "cuTensorMapEncodeTiled failed (%d): base %p dims [%llu,%llu] stride %llu box [%u,%u]", (int)r,
                      base, (unsigned long long)inner, (unsigned long long)outer,
                      (unsigned long long)strides[0], box_inner, box_outer);
  return RTG_OK;
}

// {columns, rows, planes} view of a row-major output / residual for the staged epilogue: box 32 x 32 x 1.
static rtg_status encode_out_3d(Context* ctx, CUtensorMap* map, const float* base, uint64_t cols, uint64_t rows,
                                uint64_t planes, uint64_t ld) {
  cuuint64_t dims[3] = {cols, rows, planes};
  cuuint64_t strides[2] = {ld * sizeof(float), ld * rows * sizeof(float)};
  cuuint32_t box[3] = {32, 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = ctx->encode_tiled(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box,
                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_errorf(RTG_ERR_CUDA, "cuTensorMapEncodeTiled (output) failed (%d): base %p dims [%llu,%llu,%llu] ld %llu",
                      (int)r, base, (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)planes,
                      (unsigned long long)ld);
  return RTG_OK;
}

// Packed B operand for a launch: the caller's precomputed copy, or a pre-pass into pool scratch.  Map: {K, N, 2 planes},
// box 32 x BN x 2, 128B swizzle: one load fills a stage's [B hi][B'] tiles.
static rtg_status b_pack_map(Context* ctx, const Epilogue& ep, MatView b, int bn, Scratch* tmp, CUtensorMap* map) {
  BPack bp = ep.b_pack;
  if (!bp.p) {
    bp.ld = b_pack_ld(b.rows);
    bp.plane = b.cols * bp.ld;
    RTG_TRY(tmp->alloc(ctx, static_cast<size_t>(bp.plane) * 2 * 4));
    RTG_TRY(make_b_pack(ctx, b, tmp->f32(), bp.plane));
    bp.p = tmp->f32();
  }
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(bp.ld), static_cast<cuuint64_t>(b.cols), 2};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(bp.ld) * 4, static_cast<cuuint64_t>(bp.plane) * 4};
  // (debug flag 128, wrong results, timing only) half the rows per box: what the stage would receive if a CTA pair shared
  // the B tile (cta_group::2), i.e. half the shared-memory write traffic of the B stream
  static const bool half_b = []() { const char* e = getenv("RTG_DEBUG_FLAGS"); return e && (atoi(e) & 128); }();
  cuuint32_t box[3] = {kBK, static_cast<cuuint32_t>(half_b ? bn / 2 : bn), 2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = ctx->encode_tiled(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(bp.p), dims, strides, box,
                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_errorf(RTG_ERR_CUDA, "cuTensorMapEncodeTiled (packed B) failed (%d): base %p dims [%lld,%lld,2] plane %lld",
                      (int)r, bp.p, (long long)bp.ld, (long long)b.cols, (long long)bp.plane);
  return RTG_OK;
}

// Can TMA address this operand in place?  The contiguous dimension must have stride 1, the base must be 16-byte
// aligned and the pitch of the other dimension a positive multiple of 16 bytes (irrelevant when it has one entry).
static bool direct_ok(const float* p, int64_t inner_stride, int64_t outer_stride, int64_t outer_count) {
  if ((reinterpret_cast<uintptr_t>(p) & 15) != 0 || inner_stride != 1) return false;
  return outer_count == 1 || (outer_stride > 0 && outer_stride % 4 == 0);
}

static int env_debug_flags() {
  static int f = -1;
  if (f < 0) {
    const char* e = getenv("RTG_DEBUG_FLAGS");
    f = e ? atoi(e) : 0;
  }
  return f;
}
#define debug_flags() (env_debug_flags() | (ctx->debug_flags & 0xFFFF))

// ---------------------------------------------------------------- split-K reduction
// Adds the partial-sum planes of a split-K launch in split order (fixed, so repeated runs are bit-identical) and
// applies the epilogue in the order of the kernel's own `finish`: alpha, bias, residual, activation.

__global__ void __launch_bounds__(256) splitk_reduce_kernel(const ReduceArgs a) {
  ptx::grid_dependency_wait();  // the partial sums come from the kernel launched just before
  const int nv = (a.N + 3) >> 2;
  const int64_t total = static_cast<int64_t>(a.M) * nv;
  const int64_t step = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += step) {
    const int m = static_cast<int>(i / nv);
    const int c = static_cast<int>(i - static_cast<int64_t>(m) * nv) << 2;
    const float* w = a.ws + static_cast<int64_t>(m) * a.ws_ld + c;
    const float4 x0 = *reinterpret_cast<const float4*>(w);
    float v[4] = {x0.x, x0.y, x0.z, x0.w};
    for (int s = 1; s < a.splits; ++s) {
      const float4 x = *reinterpret_cast<const float4*>(w + s * a.plane);
      v[0] = __fadd_rn(v[0], x.x), v[1] = __fadd_rn(v[1], x.y), v[2] = __fadd_rn(v[2], x.z), v[3] = __fadd_rn(v[3], x.w);
    }
    if (a.alpha_div) {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = __fdiv_rn(v[j], a.alpha);
    } else if (a.alpha != 1.f) {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] *= a.alpha;
    }
    const int ncols = min(4, a.N - c);
    if (a.bias_mode == BIAS_ROW) {
      const float b = __ldg(a.bias + m);
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] += b;
    } else if (a.bias_mode == BIAS_COL) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < ncols) v[j] += __ldg(a.bias + c + j);
    }
    if (a.R != nullptr) {
      const float* r = a.R + static_cast<int64_t>(m) * a.r_ld + c;
      if (a.vec_r) {
        const float4 r4 = *reinterpret_cast<const float4*>(r);
        v[0] += a.beta * r4.x, v[1] += a.beta * r4.y, v[2] += a.beta * r4.z, v[3] += a.beta * r4.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (j < ncols) v[j] += a.beta * r[j];
      }
    }
    apply_act_array(v, a.act, a.act_lo, a.act_hi);
    float* d = a.D + static_cast<int64_t>(m) * a.d_ld + c;
    if (a.vec_d) {
      *reinterpret_cast<float4*>(d) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < ncols) d[j] = v[j];
    }
  }
}

rtg_status launch_reduce(Context* ctx, const ReduceArgs& ra) {
  const int64_t total = static_cast<int64_t>(ra.M) * ((ra.N + 3) / 4);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(std::min<int64_t>(div_ceil(total, 256), ctx->num_sms * 8)));
  cfg.blockDim = dim3(256);
  cfg.stream = ctx->stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  RTG_CUDA(cudaLaunchKernelEx(&cfg, splitk_reduce_kernel, ra));
  ctx->launches++;
  return RTG_OK;
}

template <int A_KIND, int B_MAJOR>
static rtg_status dispatch(Context* ctx, int bn, int mode, const CUtensorMap& ma, const CUtensorMap& mb,
                           GemmArgs args, int plane_rows = 0, int64_t num_planes = 0, const Epilogue* peers = nullptr,
                           const CUtensorMap* mb2p = nullptr) {
  if (mode == RTG_MATH_TF32_BF16X2 && (!mb2p || A_KIND == 3)) mode = RTG_MATH_3XTF32;  // no B' for this launch
  args.one2 = 0x3f8000003f800000ull;
  const CUtensorMap& mb2 = mb2p ? *mb2p : mb;
  // Row-major, 16-byte aligned outputs (and residuals) move through shared memory with TMA (see GemmCfg).
  CUtensorMap md = ma, mr = ma;
  Scratch ws, peer_maps;
  const int n_peers = peers ? peers->n_peers : 0;
  if (n_peers > 0) args.splits = 1;  // the reduction pass writes the local buffer only
  ReduceArgs red = {};
  const ReduceArgs* redp = nullptr;
  auto stageable = [&](const float* p, const OutMap& om) {
    if (!p || om.hw > 0) return false;
    if ((reinterpret_cast<uintptr_t>(p + om.chan_off) & 15) != 0) return false;
    return (args.M == 1 && num_planes <= 1) || (om.ld > 0 && om.ld % 4 == 0);
  };
  // Split-K: the kernel leaves raw partial sums in a {N, M, splits} workspace through the staged (TMA store) path and
  // splitk_reduce_kernel applies the epilogue.  Row-major D / R only; plain and im2col A operands.
  if (args.splits > 1 && ((A_KIND != 0 && A_KIND != 1) || args.dmap.hw > 0 || (args.R && args.rmap.hw > 0) ||
                          (args.flags & 8) || args.num_kb < 2))
    args.splits = 1;
  if (args.splits > 1) {
    int kbs = static_cast<int>(div_ceil(args.num_kb, args.splits));
    if (mode != RTG_MATH_TF32) kbs = static_cast<int>(div_ceil(kbs, kChunkKb)) * kChunkKb;  // whole accumulation chunks
    args.kb_per_split = kbs;
    args.splits = static_cast<int>(div_ceil(args.num_kb, kbs));  // every split owns at least one k-block
  }
  if (args.splits > 1) {
    const int64_t ws_ld = div_ceil(args.N, 4) * 4, plane = ws_ld * args.M;
    RTG_TRY(ws.alloc(ctx, static_cast<size_t>(plane) * args.splits * sizeof(float)));
    red.ws = ws.f32(), red.plane = plane, red.ws_ld = ws_ld;
    red.splits = args.splits, red.M = args.M, red.N = args.N;
    red.alpha = args.alpha, red.beta = args.beta;
    red.alpha_div = (args.flags & kFlagAlphaDiv) ? 1 : 0;
    red.bias_mode = args.bias_mode, red.bias = args.bias;
    red.R = args.R, red.r_ld = args.rmap.ld;
    if (args.R) red.R += args.rmap.chan_off;
    red.D = args.D + args.dmap.chan_off, red.d_ld = args.dmap.ld;
    red.act = args.act, red.act_lo = args.act_lo, red.act_hi = args.act_hi;
    auto vec_ok = [&](const float* p, int64_t ld) {
      return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && ld % 4 == 0 && args.N % 4 == 0;
    };
    red.vec_d = vec_ok(red.D, red.d_ld);
    red.vec_r = red.R ? vec_ok(red.R, red.r_ld) : 0;
    redp = &red;
    // the kernel itself only accumulates
    args.alpha = 1.f, args.beta = 0.f, args.bias = nullptr, args.bias_mode = BIAS_NONE, args.R = nullptr;
    args.act = ACT_NONE;
    args.flags &= ~kFlagAlphaDiv;
    args.D = ws.f32();
    args.dmap = OutMap{ws_ld, 0, 0, 0};
    RTG_TRY(encode_out_3d(ctx, &md, ws.f32(), args.N, args.M, args.splits, ws_ld));
    args.flags |= kFlagStagedD;
  } else if (!(args.flags & 8) && stageable(args.D, args.dmap)) {
    // plane_rows > 0 (row-window convolution): rows are output columns of one image row, planes the image rows
    const uint64_t rows = plane_rows > 0 ? plane_rows : args.M;
    const uint64_t planes = num_planes > 0 ? num_planes : plane_rows > 0 ? args.M / plane_rows : 1;
    const uint64_t ld = (args.M == 1 && num_planes <= 1) ? static_cast<uint64_t>(div_ceil(args.N, 4) * 4) : args.dmap.ld;
    RTG_TRY(encode_out_3d(ctx, &md, args.D + args.dmap.chan_off, args.N, rows, planes, ld));
    args.flags |= kFlagStagedD;
    if (n_peers > 0) {
      CUtensorMap host_maps[7];
      for (int p = 0; p < n_peers; ++p) {
        if ((reinterpret_cast<uintptr_t>(peers->peer_D[p] + args.dmap.chan_off) & 15) != 0)
          return set_error(RTG_ERR_UNSUPPORTED_VALUE, "peer output buffers must be 16-byte aligned");
        RTG_TRY(encode_out_3d(ctx, &host_maps[p], peers->peer_D[p] + args.dmap.chan_off, args.N, rows, planes, ld));
      }
      RTG_TRY(peer_maps.alloc(ctx, sizeof(CUtensorMap) * 7));
      RTG_CUDA(cudaMemcpyAsync(peer_maps.p, host_maps, sizeof(CUtensorMap) * n_peers, cudaMemcpyHostToDevice, ctx->stream));
      args.peer_maps = static_cast<const CUtensorMap*>(peer_maps.p);
      args.n_peers = n_peers;
    }
    if (stageable(args.R, args.rmap)) {
      const uint64_t rld = (args.M == 1 && num_planes <= 1) ? static_cast<uint64_t>(div_ceil(args.N, 4) * 4) : args.rmap.ld;
      RTG_TRY(encode_out_3d(ctx, &mr, args.R + args.rmap.chan_off, args.N, rows, planes, rld));
      args.flags |= kFlagStagedR;
    }
  }
  if (n_peers > 0 && args.n_peers == 0)
    return set_error(RTG_ERR_UNSUPPORTED_VALUE, "peer stores need a row-major, 16-byte aligned output");
  return launch_variants<A_KIND, B_MAJOR>(ctx, bn, mode, ma, mb, mb2, md, mr, args, redp);
}

// Largest tile that still fills the machine; smaller tiles when the problem has too few of them.  With `splits` the
// choice is joint over (BN, split-K factor): layers whose tile count is a poor fit for 148 SMs (98 or 196 tiles)
// divide K so that the work units fill whole waves; the price is one pass over the partial sums (L2 resident).
// Rasterisation of the output tiles (GemmArgs::group_m): when the output has many tiles in both directions, walk them
// in groups of group_m m-tiles so that the tiles in flight share operand panels in L2.  A tile of A costs 16 KB per
// k-block, a tile of B (and B') 16 KB per 128 columns: the traffic of num_sms tiles in flight is minimal near
// group_m = sqrt(num_sms * b_bytes / a_bytes).
static int choose_group_m(Context* ctx, int64_t M, int64_t N, int bn) {
  static const int forced = []() {
    const char* e = getenv("RTG_GROUP_M");  // developer switch: 0 disables, n forces
    return e ? atoi(e) : -1;
  }();
  const int64_t m_tiles = div_ceil(M, kBM), n_tiles = div_ceil(N, bn);
  if (forced >= 0) return forced;
  if (n_tiles < 4 || m_tiles < 8) return 0;  // one of the operands is re-used by (almost) every tile in flight anyway
  const double ratio = bn >= 256 ? 2.0 : (bn >= 128 ? (ctx->math == RTG_MATH_TF32 ? 1.0 : 2.0) : 1.0);
  int gm = static_cast<int>(sqrt(static_cast<double>(ctx->num_sms) * ratio) + 0.5);
  gm = std::max(4, std::min<int>(gm, static_cast<int>(m_tiles)));
  return gm;
}

static int choose_bn(Context* ctx, int64_t M, int64_t N, int64_t num_kb, int64_t batch = 1, int* splits = nullptr) {
  // Cycle model fitted to the per-role counters of the PROF build (profiles/): the MMA thread needs
  // max(64, BN/2) cycles per 128 x BN x 8 instruction (64 is the dispatch floor, so BN = 64 runs the tensor pipe at
  // half rate) plus ~150 cycles of barrier traffic per k-block; the epilogue drains a tile in ~1000 + 12*BN cycles
  // with four warps (3xTF32 has eight); MMA and epilogue of consecutive tiles overlap.
  const bool x3 = ctx->math != RTG_MATH_TF32;
  const int max_bn = x3 ? 128 : 256;
  const int64_t m_tiles = div_ceil(M, kBM) * batch;
  static const int max_splits = []() {
    const char* e = getenv("RTG_SPLITK");  // developer switch: 1 disables split-K
    return e ? std::max(1, atoi(e)) : 8;
  }();
  int best = 64, best_s = 1;
  double best_cost = 1e300;
  const int cands[3] = {256, 128, 64};
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    if (bn > max_bn) continue;
    if (bn > 64 && N <= bn / 2) continue;  // mostly padding
    const int64_t tiles = m_tiles * div_ceil(N, bn);
    // per-k-block cost of a busy CTA: the fp32-accurate modes are paced by the TMA -> split -> tcgen05.st -> MMA loop
    // (1250-1350 cycles in the round-1 PROF build these constants come from; the shipped kernel measures ~720 at
    // BN = 128 and ~900 at BN = 64, and the choices below do not change anywhere between those sets: RTG_MODEL_* are
    // developer overrides used to check exactly that), TF32 by shared-memory operand reads
    static const double kb128 = []() { const char* e = getenv("RTG_MODEL_KB128"); return e ? atof(e) : 1350.0; }();
    static const double kb64 = []() { const char* e = getenv("RTG_MODEL_KB64"); return e ? atof(e) : 1250.0; }();
    static const double red_fixed = []() { const char* e = getenv("RTG_MODEL_REDUCE"); return e ? atof(e) : 5000.0; }();
    const double kb_cycles = x3 ? (bn == 128 ? kb128 : kb64) : 6.4 * std::max(64, bn / 2) + 150.0;
    const double epi_cycles = x3 ? 800.0 + 6.0 * bn : 1000.0 + 12.0 * bn;
    for (int s = 1; s <= (splits ? max_splits : 1); ++s) {
      int64_t kbs = div_ceil(num_kb, s);
      if (s > 1 && x3) kbs = div_ceil(kbs, kChunkKb) * kChunkKb;
      const int64_t s_eff = div_ceil(num_kb, std::max<int64_t>(kbs, 1));
      if (s > 1 && (s_eff != s || kbs < 4)) continue;
      const int64_t rounds = div_ceil(tiles * s, ctx->num_sms);
      const double main_cycles = static_cast<double>(kbs) * kb_cycles;
      double cost = static_cast<double>(rounds) * std::max(main_cycles, epi_cycles) + epi_cycles + 3000.0;
      // reduction pass: launch + (s partial planes read, one result written) out of L2 at ~2500 B/clk
      if (s > 1) cost += red_fixed + static_cast<double>(s + 1) * M * N * 4.0 / 2500.0;
      if (cost < best_cost * (s > 1 ? 0.9 : 1.0)) {  // a split has to pay for itself clearly
        best_cost = cost;
        best = bn;
        best_s = s;
      }
    }
  }
  if (splits) *splits = best_s;
  return best;
}

rtg_status gemm_device(Context* ctx, MatView a, MatView b, float* D, OutMap dmap, const Epilogue& ep) {
  const int64_t M = a.rows, K = a.cols, N = b.cols;
  if (M == 0 || N == 0) return RTG_OK;
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "GEMM dims exceed i32");
  // --- very few rows: memory-bound, one pass over B on the CUDA cores (src/gemm.rs:651-704 gemv)
  static const bool gemv_on = []() {
    const char* e = getenv("RTG_GEMV");  // developer switch: 0 keeps M <= 8 on the tensor path
    return !e || atoi(e) != 0;
  }();
  if (gemv_on && K > 0 && gemv_supported(a, b, dmap, ep)) return gemv_device(ctx, a, b, D, dmap, ep);

  // --- legalise A: K-major rows with 16-byte aligned base and pitch
  Scratch a_tmp, b_tmp;
  const int64_t Kp = div_ceil(K, 4) * 4;
  if (direct_ok(a.p, a.cs, a.rs, M)) {
    if (M == 1) a.rs = Kp;  // pitch of a single row is never used but must encode
  } else {
    RTG_TRY(a_tmp.alloc(ctx, static_cast<size_t>(M) * Kp * sizeof(float)));
    if (Kp != K) RTG_CUDA(cudaMemsetAsync(a_tmp.p, 0, static_cast<size_t>(M) * Kp * sizeof(float), ctx->stream));
    RTG_TRY(copy_strided_2d(ctx, a_tmp.f32(), Kp, a.p, M, K, a.rs, a.cs));
    a = MatView{a_tmp.f32(), M, K, Kp, 1};
  }
  // --- legalise B: either [N][K] K-major (b.rs == 1) or [K][N] MN-major (b.cs == 1).  RTG_MATH_TF32_BF16X2 reads B
  // only through its packed K-major copy (make_b_pack gathers any strides), so B itself needs no legal layout there.
  const bool m2 = ctx->math == RTG_MATH_TF32_BF16X2;
  int b_major = 0;
  if (m2) {
  } else if (direct_ok(b.p, b.rs, b.cs, N)) {
    b_major = 0;
    if (N == 1) b.cs = Kp;
  } else if (direct_ok(b.p, b.cs, b.rs, K)) {
    b_major = 1;
    if (K == 1) b.rs = div_ceil(N, 4) * 4;
  } else {
    // gather into [N][Kp] K-major
    RTG_TRY(b_tmp.alloc(ctx, static_cast<size_t>(N) * Kp * sizeof(float)));
    if (Kp != K) RTG_CUDA(cudaMemsetAsync(b_tmp.p, 0, static_cast<size_t>(N) * Kp * sizeof(float), ctx->stream));
    RTG_TRY(copy_strided_2d(ctx, b_tmp.f32(), Kp, b.p, N, K, b.cs, b.rs));
    b = MatView{b_tmp.f32(), K, N, 1, Kp};
    b_major = 0;
  }

  CUtensorMap ma, mb;
  int splits = 1;
  const int bn = choose_bn(ctx, M, N, div_ceil(K, kBK), 1, dmap.hw > 0 ? nullptr : &splits);
  RTG_TRY(encode_2d(ctx, &ma, a.p, K, M, a.rs, kBK, kBM));
  if (m2)
    mb = ma;  // unused by the kernel in this mode; any valid map
  else if (b_major == 0)
    RTG_TRY(encode_2d(ctx, &mb, b.p, K, N, b.cs, kBK, bn));
  else
    RTG_TRY(encode_2d(ctx, &mb, b.p, N, K, b.rs, 32, kBK, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B));

  GemmArgs g = {};
  g.M = static_cast<int>(M);
  g.N = static_cast<int>(N);
  g.K = static_cast<int>(K);
  g.num_kb = static_cast<int>(div_ceil(K, kBK));
  g.alpha = ep.alpha;
  g.beta = ep.beta;
  g.bias = ep.bias;
  g.bias_mode = ep.bias ? ep.bias_mode : BIAS_NONE;
  g.R = (ep.R && ep.beta != 0.f) ? ep.R : nullptr;
  g.D = D;
  g.dmap = dmap;
  g.rmap = ep.rmap;
  g.act = ep.act;
  g.act_lo = ep.act_lo;
  g.act_hi = ep.act_hi;
  g.flags = debug_flags() | (ep.alpha_div ? kFlagAlphaDiv : 0) | (ep.b_pack.p && ep.b_pack_const ? kFlagBConst : 0);
  g.splits = splits;
  g.group_m = choose_group_m(ctx, M, N, bn);
  ctx->pending_flops = 2.0 * g.M * g.N * g.K;
  ctx->pending_bytes = 4.0 * (static_cast<double>(g.M) * g.K + static_cast<double>(g.K) * g.N +
                              static_cast<double>(g.M) * g.N * (g.R ? 2 : 1));
  Scratch bp_tmp;
  CUtensorMap mb2;
  if (m2) RTG_TRY(b_pack_map(ctx, ep, b, bn, &bp_tmp, &mb2));
  if (b_major == 0) return dispatch<0, 0>(ctx, bn, ctx->math, ma, mb, g, 0, 0, &ep, m2 ? &mb2 : nullptr);
  return dispatch<0, 1>(ctx, bn, ctx->math, ma, mb, g, 0, 0, &ep, m2 ? &mb2 : nullptr);
}

// ---------------------------------------------------------------- batched GEMM (one launch for all batch entries)
static rtg_status encode_4d(Context* ctx, CUtensorMap* map, const float* base, const uint64_t dims_in[4],
                            const uint64_t strides_elems[3], uint32_t box_inner, uint32_t box_outer,
                            CUtensorMapSwizzle swz) {
  cuuint64_t dims[4] = {dims_in[0], dims_in[1], dims_in[2], dims_in[3]};
  cuuint64_t strides[3] = {strides_elems[0] * sizeof(float), strides_elems[1] * sizeof(float),
                           strides_elems[2] * sizeof(float)};
  cuuint32_t box[4] = {box_inner, box_outer, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = ctx->encode_tiled(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), dims, strides, box,
                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_errorf(RTG_ERR_CUDA, "cuTensorMapEncodeTiled (batched) failed (%d): base %p dims [%llu,%llu,%llu,%llu]",
                      (int)r, base, (unsigned long long)dims[0], (unsigned long long)dims[1],
                      (unsigned long long)dims[2], (unsigned long long)dims[3]);
  return RTG_OK;
}

bool gemm_batched_supported(const BatchedView& a, const BatchedView& b, int64_t n0, int64_t n1, int64_t N) {
  auto batch_ok = [&](const BatchedView& v) {
    const bool ok0 = n0 == 1 || v.bs0 == 0 || (v.bs0 > 0 && v.bs0 % 4 == 0);
    const bool ok1 = n1 == 1 || v.bs1 == 0 || (v.bs1 > 0 && v.bs1 % 4 == 0);
    return ok0 && ok1;
  };
  if (!batch_ok(a) || !batch_ok(b)) return false;
  if (n0 > INT32_MAX || n1 > INT32_MAX || n0 * n1 > INT32_MAX) return false;
  if (a.rows > INT32_MAX || a.cols > INT32_MAX || N > INT32_MAX) return false;
  if (!direct_ok(a.p, a.cs, a.rs, a.rows)) return false;
  return direct_ok(b.p, b.rs, b.cs, N) || direct_ok(b.p, b.cs, b.rs, b.rows);
}

rtg_status gemm_batched_device(Context* ctx, BatchedView a, BatchedView b, int64_t n0, int64_t n1, float* D,
                               const Epilogue& ep) {
  const int64_t M = a.rows, K = a.cols, N = b.cols, batch = n0 * n1;
  if (M == 0 || N == 0 || batch == 0) return RTG_OK;
  if (!gemm_batched_supported(a, b, n0, n1, N)) return set_error(RTG_ERR_INTERNAL, "batched GEMM operands not addressable");
  const int64_t Kp = div_ceil(K, 4) * 4, Np = div_ceil(N, 4) * 4;
  const int b_major = direct_ok(b.p, b.rs, b.cs, N) ? 0 : 1;
  const int bn = choose_bn(ctx, M, N, div_ceil(K, kBK), batch);
  // a broadcast batch dimension is encoded with extent 1 (the kernel keeps its coordinate at 0); its stride only has
  // to be a valid multiple of 16 bytes
  auto dim = [](int64_t n, int64_t bs) { return static_cast<uint64_t>(bs == 0 ? 1 : n); };
  auto str = [](int64_t bs, int64_t fallback) { return static_cast<uint64_t>(bs > 0 ? bs : fallback); };
  CUtensorMap ma, mb;
  {
    const int64_t pitch = M == 1 ? Kp : a.rs;
    const uint64_t dims[4] = {static_cast<uint64_t>(K), static_cast<uint64_t>(M), dim(n0, a.bs0), dim(n1, a.bs1)};
    const uint64_t strides[3] = {static_cast<uint64_t>(pitch), str(n0 == 1 ? 0 : a.bs0, pitch), str(n1 == 1 ? 0 : a.bs1, pitch)};
    RTG_TRY(encode_4d(ctx, &ma, a.p, dims, strides, kBK, kBM, CU_TENSOR_MAP_SWIZZLE_128B));
  }
  if (b_major == 0) {
    const int64_t pitch = N == 1 ? Kp : b.cs;
    const uint64_t dims[4] = {static_cast<uint64_t>(K), static_cast<uint64_t>(N), dim(n0, b.bs0), dim(n1, b.bs1)};
    const uint64_t strides[3] = {static_cast<uint64_t>(pitch), str(n0 == 1 ? 0 : b.bs0, pitch), str(n1 == 1 ? 0 : b.bs1, pitch)};
    RTG_TRY(encode_4d(ctx, &mb, b.p, dims, strides, kBK, bn, CU_TENSOR_MAP_SWIZZLE_128B));
  } else {
    const int64_t pitch = K == 1 ? Np : b.rs;
    const uint64_t dims[4] = {static_cast<uint64_t>(N), static_cast<uint64_t>(K), dim(n0, b.bs0), dim(n1, b.bs1)};
    const uint64_t strides[3] = {static_cast<uint64_t>(pitch), str(n0 == 1 ? 0 : b.bs0, pitch), str(n1 == 1 ? 0 : b.bs1, pitch)};
    RTG_TRY(encode_4d(ctx, &mb, b.p, dims, strides, 32, kBK, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B));
  }
  GemmArgs g = {};
  g.M = static_cast<int>(M);
  g.N = static_cast<int>(N);
  g.K = static_cast<int>(K);
  g.num_kb = static_cast<int>(div_ceil(K, kBK));
  g.alpha = ep.alpha;
  g.beta = ep.beta;
  g.bias = ep.bias;
  g.bias_mode = ep.bias ? ep.bias_mode : BIAS_NONE;
  g.R = (ep.R && ep.beta != 0.f) ? ep.R : nullptr;
  g.D = D;
  g.dmap = OutMap{N, 0, 0, 0};
  g.rmap = ep.rmap;
  g.act = ep.act;
  g.act_lo = ep.act_lo;
  g.act_hi = ep.act_hi;
  g.flags = debug_flags() | (ep.alpha_div ? kFlagAlphaDiv : 0);
  g.ic.bat_count = static_cast<int>(batch);
  g.ic.bat_n0 = static_cast<int>(n0);
  g.ic.bat_a0 = (n0 > 1 && a.bs0 != 0) ? 1 : 0;
  g.ic.bat_a1 = (n1 > 1 && a.bs1 != 0) ? 1 : 0;
  g.ic.bat_b0 = (n0 > 1 && b.bs0 != 0) ? 1 : 0;
  g.ic.bat_b1 = (n1 > 1 && b.bs1 != 0) ? 1 : 0;
  const double fb = static_cast<double>(batch);
  ctx->pending_flops = 2.0 * fb * g.M * g.N * g.K;
  ctx->pending_bytes = 4.0 * fb * (static_cast<double>(g.M) * g.K + static_cast<double>(g.K) * g.N +
                                   static_cast<double>(g.M) * g.N * (g.R ? 2 : 1));
  if (b_major == 0) return dispatch<3, 0>(ctx, bn, ctx->math, ma, mb, g, static_cast<int>(M), batch);
  return dispatch<3, 1>(ctx, bn, ctx->math, ma, mb, g, static_cast<int>(M), batch);
}

// ---------------------------------------------------------------- implicit-GEMM convolution
bool conv_im2col_supported(const ConvGeom& g) {
  // TMA im2col limits for 2 spatial dims: corners in [-128,127], tap offsets < 255, traversal stride <= 8,
  // channel pitch a multiple of 16 bytes.
  if (g.c_total % 4 != 0 || g.c_off % 4 != 0) return false;
  if (g.stride_h < 1 || g.stride_w < 1 || g.stride_h > 8 || g.stride_w > 8) return false;
  const int up_w = g.pad_r - (g.kw - 1) * g.dil_w, up_h = g.pad_b - (g.kh - 1) * g.dil_h;
  if (g.pad_l > 128 || g.pad_t > 128 || up_w < -128 || up_w > 127 || up_h < -128 || up_h > 127) return false;
  if ((g.kw - 1) * g.dil_w >= 255 || (g.kh - 1) * g.dil_h >= 255) return false;
  return true;
}

rtg_status conv_gemm_device(Context* ctx, const float* x, const ConvGeom& g, const float* wmat, int og, float* D,
                            OutMap dmap, const Epilogue& ep) {
  const int64_t M = static_cast<int64_t>(g.n) * g.oh * g.ow;
  if (M == 0 || og == 0) return RTG_OK;
  if (M > INT32_MAX) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "conv output pixels exceed i32");
  const int c_blocks = static_cast<int>(div_ceil(g.cg, kBK));
  const int64_t Kmat = static_cast<int64_t>(g.kh) * g.kw * c_blocks * kBK;

  CUtensorMap ma, mb;
  {
    // The channel extent ends with this group's slice: the 32-channel box of the last block would otherwise read the
    // next group's channels, and 0 (padded weight) x Inf / NaN there would poison outputs the reference leaves finite.
    // Past the extent TMA fills zeros.
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(std::min(g.c_total, g.c_off + g.cg)), static_cast<cuuint64_t>(g.w),
                          static_cast<cuuint64_t>(g.h), static_cast<cuuint64_t>(g.n)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(g.c_total) * 4, static_cast<cuuint64_t>(g.c_total) * g.w * 4,
                             static_cast<cuuint64_t>(g.c_total) * g.w * g.h * 4};
    int lower[2] = {-g.pad_l, -g.pad_t};
    int upper[2] = {g.pad_r - (g.kw - 1) * g.dil_w, g.pad_b - (g.kh - 1) * g.dil_h};
    cuuint32_t estr[4] = {1, static_cast<cuuint32_t>(g.stride_w), static_cast<cuuint32_t>(g.stride_h), 1};
    CUresult r = ctx->encode_im2col(&ma, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x), dims, strides,
                                    lower, upper, kBK, kBM, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
      return set_errorf(RTG_ERR_CUDA, "cuTensorMapEncodeIm2col failed (%d): NHWC [%d,%d,%d,%d] pads [%d,%d,%d,%d]",
                        (int)r, g.n, g.h, g.w, g.c_total, g.pad_t, g.pad_l, g.pad_b, g.pad_r);
    // Driver <= 13.1 mis-encodes im2col maps of tensors smaller than 128 KiB; the documented fix-up is to clear
    // bit 21 of the second descriptor word.
    const uint64_t bytes = static_cast<uint64_t>(g.n) * g.h * g.w * g.c_total * 4;
    if (ctx->driver_version <= 13010 && bytes < 131072) reinterpret_cast<uint64_t*>(&ma)[1] &= ~(1ull << 21);
  }
  int splits = 1;
  const int bn = choose_bn(ctx, M, og, Kmat / kBK, 1, dmap.hw > 0 ? nullptr : &splits);
  RTG_TRY(encode_2d(ctx, &mb, wmat, Kmat, og, Kmat, kBK, bn));

  GemmArgs a = {};
  a.M = static_cast<int>(M);
  a.N = og;
  a.K = static_cast<int>(Kmat);
  a.num_kb = static_cast<int>(Kmat / kBK);
  a.alpha = ep.alpha;
  a.beta = ep.beta;
  a.bias = ep.bias;
  a.bias_mode = ep.bias ? ep.bias_mode : BIAS_NONE;
  a.R = (ep.R && ep.beta != 0.f) ? ep.R : nullptr;
  a.D = D;
  a.dmap = dmap;
  a.rmap = ep.rmap;
  a.act = ep.act;
  a.act_lo = ep.act_lo;
  a.act_hi = ep.act_hi;
  a.flags = debug_flags() | (ep.b_pack.p && ep.b_pack_const ? kFlagBConst : 0);
  a.splits = splits;
  a.group_m = choose_group_m(ctx, M, og, bn);
  a.ic.ow = g.ow;
  a.ic.oh = g.oh;
  a.ic.stride_w = g.stride_w;
  a.ic.stride_h = g.stride_h;
  a.ic.pad_l = g.pad_l;
  a.ic.pad_t = g.pad_t;
  a.ic.dil_w = g.dil_w;
  a.ic.dil_h = g.dil_h;
  a.ic.kw = g.kw;
  a.ic.kh = g.kh;
  a.ic.c_blocks = c_blocks;
  a.ic.c_off = g.c_off;
  // algorithmic work of this launch: true K = C/g * kh * kw (not the per-tap padded K the tiles walk); bytes = the
  // input image slice once + weights + output (+ residual)
  const double k_true = static_cast<double>(g.cg) * g.kh * g.kw;
  ctx->pending_flops = 2.0 * M * og * k_true;
  ctx->pending_bytes = 4.0 * (static_cast<double>(g.n) * g.h * g.w * g.cg + og * k_true +
                              static_cast<double>(M) * og * (a.R ? 2 : 1));
  Scratch bp_tmp;
  CUtensorMap mb2;
  const bool m2 = ctx->math == RTG_MATH_TF32_BF16X2;
  if (m2) RTG_TRY(b_pack_map(ctx, ep, MatView{wmat, Kmat, og, 1, Kmat}, bn, &bp_tmp, &mb2));
  return dispatch<1, 0>(ctx, bn, ctx->math, ma, mb, a, 0, 0, nullptr, m2 ? &mb2 : nullptr);
}

// ---------------------------------------------------------------- row-window convolution
rtg_status conv_window_gemm_device(Context* ctx, const float* x_pad, const WindowGeom& g, const float* wmat, int og,
                                   float* D, OutMap dmap, const Epilogue& ep) {
  const int64_t M = static_cast<int64_t>(g.n) * g.oh * g.ow;
  if (M == 0 || og == 0) return RTG_OK;
  if (M > INT32_MAX) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "conv output pixels exceed i32");
  CUtensorMap ma, mb;
  {
    // {32-float window, output column, padded input row, image}; consecutive windows overlap (pitch stride_w * cp)
    cuuint64_t dims[4] = {32, static_cast<cuuint64_t>(g.ow), static_cast<cuuint64_t>(g.hp), static_cast<cuuint64_t>(g.n)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(g.stride_w) * g.cp * 4, static_cast<cuuint64_t>(g.wp) * g.cp * 4,
                             static_cast<cuuint64_t>(g.hp) * g.wp * g.cp * 4};
    cuuint32_t box[4] = {32, kBM, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = ctx->encode_tiled(&ma, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x_pad), dims, strides,
                                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
      return set_errorf(RTG_ERR_CUDA, "cuTensorMapEncodeTiled (row window) failed (%d): padded image [%d,%d,%d,%d]",
                        (int)r, g.n, g.hp, g.wp, g.cp);
  }
  const int64_t Kmat = static_cast<int64_t>(g.kh) * kBK;
  const int bn = choose_bn(ctx, M, og, g.kh);
  RTG_TRY(encode_2d(ctx, &mb, wmat, Kmat, og, Kmat, kBK, bn));
  GemmArgs a = {};
  a.M = static_cast<int>(M);
  a.N = og;
  a.K = static_cast<int>(Kmat);
  a.num_kb = g.kh;
  a.alpha = ep.alpha;
  a.beta = ep.beta;
  a.bias = ep.bias;
  a.bias_mode = ep.bias ? ep.bias_mode : BIAS_NONE;
  a.R = (ep.R && ep.beta != 0.f) ? ep.R : nullptr;
  a.D = D;
  a.dmap = dmap;
  a.rmap = ep.rmap;
  a.act = ep.act;
  a.act_lo = ep.act_lo;
  a.act_hi = ep.act_hi;
  a.flags = debug_flags() | (ep.b_pack.p && ep.b_pack_const ? kFlagBConst : 0);
  a.ic.ow = g.ow;
  a.ic.oh = g.oh;
  a.ic.stride_h = g.stride_h;
  a.ic.dil_h = g.dil_h;
  a.ic.win_tpr = static_cast<int>(div_ceil(g.ow, kBM));
  a.ic.win_rows = g.n * g.oh;
  const double k_true = static_cast<double>(g.c) * g.kh * g.kw;
  ctx->pending_flops = 2.0 * M * og * k_true;
  ctx->pending_bytes = 4.0 * (static_cast<double>(g.n) * g.hp * g.wp * g.cp + og * k_true +
                              static_cast<double>(M) * og * (a.R ? 2 : 1));
  Scratch bp_tmp;
  CUtensorMap mb2;
  const bool m2 = ctx->math == RTG_MATH_TF32_BF16X2;
  if (m2) RTG_TRY(b_pack_map(ctx, ep, MatView{wmat, Kmat, og, 1, Kmat}, bn, &bp_tmp, &mb2));
  return dispatch<2, 0>(ctx, bn, ctx->math, ma, mb, a, g.ow, 0, nullptr, m2 ? &mb2 : nullptr);
}

// ---------------------------------------------------------------- split-K GEMM on static scratch (no allocation)
// rtg_gemm_static_scratch (include/rten_gpu.h): the split-K form of the contraction for callers that may not allocate
// (CUDA graph capture, tight serving loops).  Its scratch — the {N, M, splits} partial-sum planes that
// splitk_reduce_kernel folds, preceded by the bf16 correction operand B' in RTG_MATH_TF32_BF16X2 — lives in this
// file-scope device array instead of the context pool.  Every GPU that loads the library holds it until exit; all calls
// share it, so calls that overlap on the GPU (two streams, two contexts of one device) race.  Nothing in it has to
// start at zero: every plane element and every B' word is written by this call before it is read (TMA stores cover the
// whole {N, M, splits} box, make_b_pack zero-fills beyond K).
static __device__ __align__(256) float g_static_scratch[RTG_STATIC_SCRATCH_BYTES / sizeof(float)];

// The array's name is a host-side stand-in in host code, not a device address: resolve it once per device, with that
// device current.
static rtg_status static_scratch_base(Context* ctx, float** out) {
  static float* base[64] = {};
  static std::mutex mu;
  if (ctx->device < 0 || ctx->device >= 64) return set_error(RTG_ERR_INVALID_VALUE, "device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!base[ctx->device]) {
    RTG_CUDA(cudaSetDevice(ctx->device));
    float* p = nullptr;
    RTG_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch));
    if (!p) return set_error(RTG_ERR_CUDA, "cudaGetSymbolAddress returned NULL for the static scratch array");
    base[ctx->device] = p;
  }
  *out = base[ctx->device];
  return RTG_OK;
}

rtg_status gemm_splitk_static(Context* ctx, const float* a, int64_t M, int64_t K, const float* b, int64_t N, float* D,
                              int64_t ldd, float alpha, float beta, const float* bias, int splits_req) {
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "GEMM dims exceed i32");
  auto aligned = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!aligned(a) || !aligned(b) || !aligned(D) || K % 4 != 0 || N % 4 != 0 || ldd % 4 != 0 || ldd < N)
    return set_error(RTG_ERR_UNSUPPORTED_VALUE,
                     "rtg_gemm_static_scratch: a, b and out must be 16-byte aligned dense row-major matrices with "
                     "k, n and out_row_stride multiples of 4 (nothing may need a gathered copy)");
  if (ctx->ktiming)
    return set_error(RTG_ERR_UNSUPPORTED_VALUE, "rtg_gemm_static_scratch: per-launch kernel timing creates events; disable it");
  const int64_t num_kb = div_ceil(K, kBK);
  if (num_kb < 2) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "rtg_gemm_static_scratch: k must span at least two 32-deep blocks");
  int mode = ctx->math;
  int splits = 1;
  const int bn = choose_bn(ctx, M, N, num_kb, 1, &splits);
  if (splits_req > 0) splits = splits_req;
  if (splits < 2) splits = 2;  // this entry point is the split-K form: at least two partial planes
  int kbs = static_cast<int>(div_ceil(num_kb, splits));
  if (mode != RTG_MATH_TF32) kbs = static_cast<int>(div_ceil(kbs, kChunkKb)) * kChunkKb;  // whole accumulation chunks
  splits = static_cast<int>(div_ceil(num_kb, kbs));
  if (splits < 2)
    return set_error(RTG_ERR_UNSUPPORTED_VALUE, "rtg_gemm_static_scratch: k is too short to split (needs more than 128 in the fp32-accurate modes)");

  // scratch layout: [B' (TF32+BF16x2 only)][splits planes of M x N partial sums]
  const int64_t kw = num_kb * kBK;
  const size_t bprime_bytes = mode == RTG_MATH_TF32_BF16X2 ? static_cast<size_t>(N) * kw * 4 * 2 : 0;  // two planes
  const size_t bprime_pad = (bprime_bytes + 255) & ~static_cast<size_t>(255);
  const int64_t plane = N * M;
  const size_t need = bprime_pad + static_cast<size_t>(plane) * splits * sizeof(float);
  if (need > RTG_STATIC_SCRATCH_BYTES)
    return set_errorf(RTG_ERR_UNSUPPORTED_VALUE,
                      "rtg_gemm_static_scratch: m=%lld n=%lld k=%lld with %d splits needs %zu bytes of scratch, the static "
                      "buffer holds %zu (RTG_STATIC_SCRATCH_BYTES); use rtg_gemm",
                      (long long)M, (long long)N, (long long)K, splits, need, static_cast<size_t>(RTG_STATIC_SCRATCH_BYTES));
  float* base = nullptr;
  RTG_TRY(static_scratch_base(ctx, &base));
  uint32_t* bprime = reinterpret_cast<uint32_t*>(base);
  float* ws = reinterpret_cast<float*>(reinterpret_cast<char*>(base) + bprime_pad);

  CUtensorMap ma, mb, mb2, md;
  RTG_TRY(encode_2d(ctx, &ma, a, K, M, K, kBK, kBM));
  RTG_TRY(encode_2d(ctx, &mb, b, N, K, N, 32, kBK, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B));  // row-major [K][N]: MN-major B
  mb2 = mb;
  if (mode == RTG_MATH_TF32_BF16X2) {
    Epilogue pk;
    pk.b_pack = BPack{reinterpret_cast<const float*>(bprime), kw, N * kw};
    RTG_TRY(make_b_pack(ctx, MatView{b, K, N, N, 1}, reinterpret_cast<float*>(bprime), N * kw));
    RTG_TRY(b_pack_map(ctx, pk, MatView{b, K, N, N, 1}, bn, nullptr, &mb2));
  }
  RTG_TRY(encode_out_3d(ctx, &md, ws, N, M, splits, N));

  GemmArgs g = {};
  g.M = static_cast<int>(M), g.N = static_cast<int>(N), g.K = static_cast<int>(K);
  g.num_kb = static_cast<int>(num_kb);
  g.alpha = 1.f, g.beta = 0.f;  // the kernel only accumulates; the reduction pass applies the epilogue
  g.bias_mode = BIAS_NONE;
  g.D = ws;
  g.dmap = OutMap{N, 0, 0, 0};
  g.act = ACT_NONE;
  g.flags = kFlagStagedD;
  g.splits = splits, g.kb_per_split = kbs;

  ReduceArgs red = {};
  red.ws = ws, red.plane = plane, red.ws_ld = N;
  red.splits = splits, red.M = g.M, red.N = g.N;
  red.alpha = alpha, red.beta = beta;
  red.bias_mode = bias ? BIAS_ROW : BIAS_NONE, red.bias = bias;
  red.R = beta != 0.f ? D : nullptr, red.r_ld = ldd;
  red.D = D, red.d_ld = ldd;
  red.act = ACT_NONE;
  red.vec_d = 1, red.vec_r = red.R ? 1 : 0;
  ctx->pending_flops = 2.0 * g.M * g.N * g.K;
  ctx->pending_bytes = 4.0 * (static_cast<double>(g.M) * g.K + static_cast<double>(g.K) * g.N +
                              static_cast<double>(g.M) * g.N * (red.R ? 2 : 1));
  if (mode == RTG_MATH_TF32_BF16X2) return launch_variants<0, 0>(ctx, bn, mode, ma, mb, mb2, md, ma, g, &red);
  return launch_variants<0, 1>(ctx, bn, mode, ma, mb, mb2, md, ma, g, &red);
}

}  // namespace rtg
