// Memory-bound GEMM for very few rows (M <= 8): the reference's gemv path (src/gemm.rs:651-704 gemv,
// src/gemm/kernels.rs:26-194 simd_gemv / simd_gemv_transposed, taken when M == 1) and the batch-1..8 heads of the
// model graphs (ResNet fc, BERT pooler).  On the tensor path such a product occupies 1/128 .. 8/128 of a 128-row tile;
// here it is what it is on any machine: one pass over B at HBM speed, fp32 FMAs on the CUDA cores.
//
//   y[m, n] = act(alpha * sum_k a[m, k] * B[k, n] + beta * r[m, n] + bias)
//
// Two kernels by the layout of B:
//   K-major  (B[k, n] at b + n*ld + k, the transposed-B / Gemm.transB case): one warp per column n; lanes stride over k
//            with 16-byte loads, the M partial sums are reduced with warp shuffles (fixed tree: reruns bit-identical).
//   MN-major (row-major B[k, n] at b + k*ld + n): a thread owns 4 adjacent columns (one float4 per k), the 8 warps of
//            a block split a K range eight ways and fold their sums through shared memory in warp order; blocks along
//            grid.y split K further, into a {splits, M, N} workspace folded by the second kernel in split order.
// Algorithmic bytes per call: 4 * (K*N + M*K + M*N (+ M*N residual)); arithmetic is exact fp32 (no TF32 anywhere), so
// every math mode gives the same result.
#include "gemm_host.h"

namespace rtg {

struct GemvArgs {
  const float* a;
  int64_t a_rs, a_cs;
  const float* b;
  int64_t b_ld;
  int M, N, K;
  float alpha, beta;
  int alpha_div;
  const float* bias;
  int bias_mode;
  const float* R;
  int64_t r_ld;
  float* D;
  int64_t d_ld;
  int act;
  float act_lo, act_hi;
  float* ws;        // MN-major with splits > 1: [splits][M][N] partial sums
  int splits, k_per_split;
};

__device__ __forceinline__ float gemv_finish(const GemvArgs& g, float v, int m, int n) {
  if (g.alpha_div) v = __fdiv_rn(v, g.alpha);
  else if (g.alpha != 1.f) v *= g.alpha;
  if (g.bias_mode == BIAS_ROW) v += __ldg(g.bias + m);
  else if (g.bias_mode == BIAS_COL) v += __ldg(g.bias + n);
  if (g.R) v += g.beta * g.R[static_cast<int64_t>(m) * g.r_ld + n];
  return apply_act(v, g.act, g.act_lo, g.act_hi);
}

// ---------------------------------------------------------------- K-major B: a warp per column
// NC columns per warp: with several rows of A, one load of an A vector serves NC columns (A is re-read per column
// group from L1 / L2: M * K * 4 bytes against K * 4 per column of B, so for M > 1 the A reads would dominate).
template <int MR, int NC, bool VEC>
__global__ void __launch_bounds__(256) gemv_kmajor_kernel(const GemvArgs g) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const int64_t groups = (g.N + NC - 1) / NC;
  for (int64_t grp = warp; grp < groups; grp += warps) {
    const int64_t n0 = grp * NC;
    const float* bcol[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) bcol[c] = g.b + min(n0 + c, static_cast<int64_t>(g.N) - 1) * g.b_ld;  // clamped: not stored
    float acc[MR][NC];
#pragma unroll
    for (int m = 0; m < MR; ++m)
#pragma unroll
      for (int c = 0; c < NC; ++c) acc[m][c] = 0.f;
    if (VEC) {  // a rows and B columns 16-byte aligned, unit stride along k
#pragma unroll 2
      for (int k = lane * 4; k < g.K; k += 128) {
        float4 bv[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) bv[c] = __ldg(reinterpret_cast<const float4*>(bcol[c] + k));
#pragma unroll
        for (int m = 0; m < MR; ++m)
          if (m < g.M) {
            const float4 av = __ldg(reinterpret_cast<const float4*>(g.a + m * g.a_rs + k));
#pragma unroll
            for (int c = 0; c < NC; ++c) {
              acc[m][c] = fmaf(av.x, bv[c].x, acc[m][c]);
              acc[m][c] = fmaf(av.y, bv[c].y, acc[m][c]);
              acc[m][c] = fmaf(av.z, bv[c].z, acc[m][c]);
              acc[m][c] = fmaf(av.w, bv[c].w, acc[m][c]);
            }
          }
      }
    } else {
      for (int k = lane; k < g.K; k += 32) {
        float bv[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) bv[c] = __ldg(bcol[c] + k);
#pragma unroll
        for (int m = 0; m < MR; ++m)
          if (m < g.M) {
            const float av = __ldg(g.a + m * g.a_rs + k * g.a_cs);
#pragma unroll
            for (int c = 0; c < NC; ++c) acc[m][c] = fmaf(av, bv[c], acc[m][c]);
          }
      }
    }
#pragma unroll
    for (int m = 0; m < MR; ++m)
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        float v = acc[m][c];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && m < g.M && n0 + c < g.N)
          g.D[static_cast<int64_t>(m) * g.d_ld + n0 + c] = gemv_finish(g, v, m, static_cast<int>(n0 + c));
      }
  }
}

// K-major B with several rows of A: the rows (M * K * 4 bytes, at most kGemvSmemBytes) are staged in shared memory once
// per CTA, so the B columns stream from HBM / L2 while A is read from shared memory instead of L2 once per column.  A warp
// per column, the CTA's eight warps walk the columns round-robin.  (The classifier head at batch 8: 1000 columns x 64 KB
// of A through L2 made the plain kernel 18 us; the weights alone are an 8 MB read.)
constexpr int kGemvSmemBytes = 96 * 1024;
template <int MR>
__global__ void __launch_bounds__(256) gemv_kmajor_smem_kernel(const GemvArgs g) {
  extern __shared__ __align__(16) float gemv_a[];  // [M][K]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int kv = g.K >> 2;
  for (int i = threadIdx.x; i < g.M * kv; i += blockDim.x) {
    const int m = i / kv, k4 = i - m * kv;
    reinterpret_cast<float4*>(gemv_a)[i] = __ldg(reinterpret_cast<const float4*>(g.a + m * g.a_rs) + k4);
  }
  __syncthreads();
  for (int n = blockIdx.x * 8 + warp; n < g.N; n += gridDim.x * 8) {
    const float4* bcol = reinterpret_cast<const float4*>(g.b + static_cast<int64_t>(n) * g.b_ld);
    float acc[MR];
#pragma unroll
    for (int m = 0; m < MR; ++m) acc[m] = 0.f;
#pragma unroll 4
    for (int k4 = lane; k4 < kv; k4 += 32) {
      const float4 bv = __ldg(bcol + k4);
#pragma unroll
      for (int m = 0; m < MR; ++m)
        if (m < g.M) {
          const float4 av = reinterpret_cast<const float4*>(gemv_a)[m * kv + k4];
          acc[m] = fmaf(av.x, bv.x, acc[m]);
          acc[m] = fmaf(av.y, bv.y, acc[m]);
          acc[m] = fmaf(av.z, bv.z, acc[m]);
          acc[m] = fmaf(av.w, bv.w, acc[m]);
        }
    }
#pragma unroll
    for (int m = 0; m < MR; ++m) {
      float v = acc[m];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && m < g.M) g.D[static_cast<int64_t>(m) * g.d_ld + n] = gemv_finish(g, v, m, n);
    }
  }
}

// ---------------------------------------------------------------- MN-major B: a thread per 4 columns
template <int MR>
__global__ void __launch_bounds__(256) gemv_mnmajor_kernel(const GemvArgs g) {
  __shared__ float4 part[8][MR][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n0 = (blockIdx.x * 32 + lane) * 4;
  const int split = blockIdx.y;
  const int k_lo = split * g.k_per_split, k_hi = min(g.K, k_lo + g.k_per_split);
  float4 acc[MR];
#pragma unroll
  for (int m = 0; m < MR; ++m) acc[m] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (n0 < g.N) {
    for (int k = k_lo + warp; k < k_hi; k += 8) {
      const float4 bv = __ldg(reinterpret_cast<const float4*>(g.b + static_cast<int64_t>(k) * g.b_ld + n0));
#pragma unroll
      for (int m = 0; m < MR; ++m)
        if (m < g.M) {
          const float av = __ldg(g.a + m * g.a_rs + k * g.a_cs);
          acc[m].x = fmaf(av, bv.x, acc[m].x);
          acc[m].y = fmaf(av, bv.y, acc[m].y);
          acc[m].z = fmaf(av, bv.z, acc[m].z);
          acc[m].w = fmaf(av, bv.w, acc[m].w);
        }
    }
  }
#pragma unroll
  for (int m = 0; m < MR; ++m) part[warp][m][lane] = acc[m];
  __syncthreads();
  if (warp == 0 && n0 < g.N) {
#pragma unroll
    for (int m = 0; m < MR; ++m) {
      if (m >= g.M) break;
      float4 v = part[0][m][lane];
      for (int w = 1; w < 8; ++w) {  // warp order: fixed
        const float4 p = part[w][m][lane];
        v.x += p.x, v.y += p.y, v.z += p.z, v.w += p.w;
      }
      if (g.splits > 1) {
        *reinterpret_cast<float4*>(g.ws + (static_cast<int64_t>(split) * g.M + m) * g.N + n0) = v;
      } else {
        float* d = g.D + static_cast<int64_t>(m) * g.d_ld + n0;
        d[0] = gemv_finish(g, v.x, m, n0), d[1] = gemv_finish(g, v.y, m, n0 + 1);
        d[2] = gemv_finish(g, v.z, m, n0 + 2), d[3] = gemv_finish(g, v.w, m, n0 + 3);
      }
    }
  }
}

__global__ void __launch_bounds__(256) gemv_reduce_kernel(const GemvArgs g) {
  const int64_t total = static_cast<int64_t>(g.M) * g.N;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int m = static_cast<int>(i / g.N), n = static_cast<int>(i - static_cast<int64_t>(m) * g.N);
    float v = g.ws[i];
    for (int s = 1; s < g.splits; ++s) v += g.ws[static_cast<int64_t>(s) * total + i];  // split order: fixed
    g.D[static_cast<int64_t>(m) * g.d_ld + n] = gemv_finish(g, v, m, n);
  }
}

// Can (and should) this product take the gemv path?  Row-major output, no peers, B addressable as K-major columns or
// as 16-byte aligned row-major rows with N % 4 == 0.
bool gemv_supported(const MatView& a, const MatView& b, const OutMap& dmap, const Epilogue& ep) {
  if (a.rows < 1 || a.rows > 8 || a.cols < 1 || b.cols < 1) return false;
  if (dmap.hw > 0 || dmap.chan_off != 0 || ep.n_peers > 0) return false;
  if (ep.R && ep.beta != 0.f && (ep.rmap.hw > 0 || ep.rmap.chan_off != 0)) return false;
  if (a.rs < 0 || a.cs < 0 || b.rs < 0 || b.cs < 0) return false;
  if (b.rs == 1 || b.rows == 1) return true;  // K-major (any column pitch)
  return b.cs == 1 && b.cols % 4 == 0 && b.rs % 4 == 0 && (reinterpret_cast<uintptr_t>(b.p) & 15) == 0;
}

rtg_status gemv_device(Context* ctx, MatView a, MatView b, float* D, OutMap dmap, const Epilogue& ep) {
  GemvArgs g = {};
  g.a = a.p, g.a_rs = a.rs, g.a_cs = a.cs;
  g.b = b.p;
  g.M = static_cast<int>(a.rows), g.K = static_cast<int>(a.cols), g.N = static_cast<int>(b.cols);
  g.alpha = ep.alpha, g.beta = ep.beta, g.alpha_div = ep.alpha_div ? 1 : 0;
  g.bias = ep.bias, g.bias_mode = ep.bias ? ep.bias_mode : BIAS_NONE;
  g.R = (ep.R && ep.beta != 0.f) ? ep.R : nullptr, g.r_ld = ep.rmap.ld;
  g.D = D, g.d_ld = dmap.ld;
  g.act = ep.act, g.act_lo = ep.act_lo, g.act_hi = ep.act_hi;
  g.splits = 1, g.k_per_split = g.K;
  ctx->pending_flops = 2.0 * g.M * g.N * g.K;
  ctx->pending_bytes = 4.0 * (static_cast<double>(g.K) * g.N + static_cast<double>(g.M) * g.K +
                              static_cast<double>(g.M) * g.N * (g.R ? 2 : 1));
  const bool kmajor = b.rs == 1 || b.rows == 1;
  Scratch ws;
  if (kmajor) {
    g.b_ld = b.cs;
    const bool vec = a.cs == 1 && g.K % 4 == 0 && (a.rows == 1 || a.rs % 4 == 0) && b.cs % 4 == 0 &&
                     (reinterpret_cast<uintptr_t>(a.p) & 15) == 0 && (reinterpret_cast<uintptr_t>(b.p) & 15) == 0;
    // Four columns per warp only when that still leaves ~16 warps per SM: the loop is a chain of dependent loads, and a
    // 1000-column classifier head with 250 busy warps ran at 45 us (batch 8) against 18 us with a warp per column.
    const int nc = (g.M > 1 && div_ceil(g.N, 4) >= static_cast<int64_t>(ctx->num_sms) * 16) ? 4 : 1;
    const int blocks = static_cast<int>(std::min<int64_t>(div_ceil(div_ceil(g.N, nc), 8), static_cast<int64_t>(ctx->num_sms) * 8));
#define RTG_GEMV_K(MR, NC)                                                                  \
  do {                                                                                      \
    if (vec) gemv_kmajor_kernel<MR, NC, true><<<blocks, 256, 0, ctx->stream>>>(g);          \
    else gemv_kmajor_kernel<MR, NC, false><<<blocks, 256, 0, ctx->stream>>>(g);             \
  } while (0)
    const size_t a_bytes = static_cast<size_t>(g.M) * g.K * 4;
    if (g.M > 1 && vec && nc == 1 && a_bytes <= kGemvSmemBytes) {
      // same summation order as gemv_kmajor_kernel<MR, 1, true> (lane-strided float4 steps, then the shuffle tree)
      const int sblocks = static_cast<int>(std::min<int64_t>(div_ceil(g.N, 8), static_cast<int64_t>(ctx->num_sms) * 2));
      auto kern = g.M <= 4 ? gemv_kmajor_smem_kernel<4> : gemv_kmajor_smem_kernel<8>;
      static bool configured[2][16] = {};
      bool& done = configured[g.M <= 4 ? 0 : 1][ctx->device & 15];
      if (!done) {
        RTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemvSmemBytes));
        done = true;
      }
      kern<<<sblocks, 256, a_bytes, ctx->stream>>>(g);
      ctx->last_kernel = "gemv<B=K,float4,A in smem>";
      ctx->launches++;
      RTG_CUDA(cudaGetLastError());
      return RTG_OK;
    }
    if (g.M <= 1) RTG_GEMV_K(1, 1);
    else if (g.M <= 4 && nc == 4) RTG_GEMV_K(4, 4);
    else if (g.M <= 4) RTG_GEMV_K(4, 1);
    else if (nc == 4) RTG_GEMV_K(8, 4);
    else RTG_GEMV_K(8, 1);
#undef RTG_GEMV_K
    ctx->last_kernel = vec ? "gemv<B=K,float4>" : "gemv<B=K>";
  } else {
    g.b_ld = b.rs;
    const int col_blocks = static_cast<int>(div_ceil(g.N, 128));
    // enough blocks to pull from every HBM channel: split K across grid.y until ~4 blocks per SM (8+ rows per warp)
    int splits = static_cast<int>(std::max<int64_t>(1, std::min<int64_t>(div_ceil(4 * ctx->num_sms, col_blocks), g.K / 64)));
    splits = std::min(splits, 64);
    g.k_per_split = static_cast<int>(div_ceil(g.K, splits));
    g.splits = static_cast<int>(div_ceil(g.K, g.k_per_split));
    if (g.splits > 1) {
      RTG_TRY(ws.alloc(ctx, static_cast<size_t>(g.splits) * g.M * g.N * 4));
      g.ws = ws.f32();
    }
    dim3 grid(col_blocks, g.splits);
    if (g.M <= 1) gemv_mnmajor_kernel<1><<<grid, 256, 0, ctx->stream>>>(g);
    else if (g.M <= 4) gemv_mnmajor_kernel<4><<<grid, 256, 0, ctx->stream>>>(g);
    else gemv_mnmajor_kernel<8><<<grid, 256, 0, ctx->stream>>>(g);
    if (g.splits > 1) {
      ctx->launches++;
      RTG_CUDA(cudaGetLastError());
      const int rb = static_cast<int>(std::min<int64_t>(div_ceil(static_cast<int64_t>(g.M) * g.N, 256), ctx->num_sms * 4));
      gemv_reduce_kernel<<<rb, 256, 0, ctx->stream>>>(g);
    }
    ctx->last_kernel = g.splits > 1 ? "gemv<B=MN,splitK>" : "gemv<B=MN>";
  }
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  return RTG_OK;
}

}  // namespace rtg
