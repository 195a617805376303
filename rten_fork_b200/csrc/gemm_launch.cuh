// Launch side of the contraction kernel: one `launch_variants<A_KIND, B_MAJOR>` per operand addressing mode, each
// explicitly instantiated in its own translation unit (gemm_inst_*.cu) so that the 36 kernel variants compile in
// parallel.  gemm_host.cu only sees the declarations.
#pragma once
#include "gemm_host.h"
#include <algorithm>
#include <vector>

#include <stdio.h>
#include <stdlib.h>

namespace rtg {

struct ReduceArgs {
  const float* ws;     // [splits][M][ws_ld]
  int64_t plane, ws_ld;
  int32_t splits, M, N;
  float alpha, beta;
  int32_t alpha_div, bias_mode;
  const float* bias;
  const float* R;
  int64_t r_ld;
  float* D;
  int64_t d_ld;
  int32_t act;
  float act_lo, act_hi;
  int32_t vec_d, vec_r;  // 16-byte accesses are legal for D / R
};

rtg_status launch_reduce(Context* ctx, const ReduceArgs& ra);

template <int A_KIND, int B_MAJOR>
rtg_status launch_variants(Context* ctx, int bn, int mode, const CUtensorMap& ma, const CUtensorMap& mb,
                           const CUtensorMap& mb2, const CUtensorMap& md, const CUtensorMap& mr, const GemmArgs& args,
                           const ReduceArgs* red);

#ifdef RTG_GEMM_INSTANTIATE
// ---------------------------------------------------------------- launch
template <int BN, int MODE, int A_KIND, int B_MAJOR, bool EPI2 = false>
static rtg_status launch(Context* ctx, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mb2,
                         const CUtensorMap& md, const CUtensorMap& mr, const GemmArgs& args,
                         const ReduceArgs* red = nullptr) {
  using Cfg = GemmCfg<BN, MODE, EPI2>;
  auto kern = gemm_tcgen05_kernel<BN, MODE, A_KIND, B_MAJOR, EPI2>;
  static bool configured[16] = {};  // per device
  if (!configured[ctx->device & 15]) {
    RTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    configured[ctx->device & 15] = true;
  }
  const int m_tiles_1 = (args.M + kBM - 1) / kBM, n_tiles = (args.N + BN - 1) / BN;
  const int64_t m_tiles = A_KIND == 2 ? static_cast<int64_t>(args.ic.win_rows) * args.ic.win_tpr
                                      : A_KIND == 3 ? static_cast<int64_t>(args.ic.bat_count) * m_tiles_1 : m_tiles_1;
  if (m_tiles * n_tiles > INT32_MAX) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "too many output tiles");
  const int64_t units = m_tiles * n_tiles * (args.splits > 1 ? args.splits : 1);
  if (units > INT32_MAX) return set_error(RTG_ERR_UNSUPPORTED_VALUE, "too many work units");
  const int grid = static_cast<int>(std::min<int64_t>(units, ctx->num_sms));
#ifdef RTG_PROF
  static unsigned long long* prof_buf = nullptr;
  GemmArgs pargs = args;
  if (getenv("RTG_PROF_PRINT")) {
    if (!prof_buf) RTG_CUDA(cudaMalloc(&prof_buf, 148 * 32 * 8));
    RTG_CUDA(cudaMemsetAsync(prof_buf, 0, 148 * 32 * 8, ctx->stream));
    pargs.prof = prof_buf;
  }
#define args pargs
#endif
  Context::KernelRec rec{};
  if (ctx->ktiming) {
    RTG_CUDA(cudaEventCreate(&rec.a));
    RTG_CUDA(cudaEventCreate(&rec.b));
    rec.flops = ctx->pending_flops, rec.bytes = ctx->pending_bytes;
    RTG_CUDA(cudaEventRecord(rec.a, ctx->stream));
  }
  {
    // Programmatic stream serialisation: the kernel's prologue (barrier init, TMEM allocation, tensor-map prefetch)
    // overlaps the tail of the previous kernel; it synchronises itself with griddepcontrol.wait.
    static const bool pdl = []() {
      const char* e = getenv("RTG_PDL");
      return !e || atoi(e) != 0;
    }();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(Cfg::kThreads);
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    RTG_CUDA(cudaLaunchKernelEx(&cfg, kern, ma, mb, mb2, md, mr, args));
  }
  ctx->launches++;
  RTG_CUDA(cudaGetLastError());
  if (red) RTG_TRY(launch_reduce(ctx, *red));  // part of the same timing record: the contraction is not done before
  if (ctx->ktiming) {
    RTG_CUDA(cudaEventRecord(rec.b, ctx->stream));
    ctx->krecs.push_back(rec);
  }
#ifdef RTG_PROF
#undef args
  if (pargs.prof) {
    std::vector<unsigned long long> h(148 * 32);
    RTG_CUDA(cudaStreamSynchronize(ctx->stream));
    RTG_CUDA(cudaMemcpy(h.data(), prof_buf, h.size() * 8, cudaMemcpyDeviceToHost));
    static const char* roles[4] = {"tma", "mma", "epi", "split"};
    static const char* slots[4][4] = {{"wait_empty", "issue", "-", "-"},
                                      {"wait_tempty", "wait_full", "issue", "-"},
                                      {"waits(tfull,resid,stread)", "drain(ldtm)", "math(bias,res,act)", "stage+store+setup"},
                                      {"wait_afull", "wait_empty", "st_wait_and_arrive", "lds_and_split"}};
    fprintf(stderr, "[prof] BN=%d mode=%d akind=%d grid=%d M=%d N=%d kb=%d:", BN, MODE, A_KIND, grid, args.M, args.N,
            args.num_kb);
    for (int r = 0; r < 4; ++r) {
      double acc[5] = {0, 0, 0, 0, 0};
      for (int b = 0; b < grid; ++b)
        for (int i = 0; i < 5; ++i) acc[i] += static_cast<double>(h[(b * 4 + r) * 8 + i]) / grid;
      if (acc[4] == 0) continue;
      fprintf(stderr, " | %s total %.0f", roles[r], acc[4]);
      for (int i = 0; i < 4; ++i)
        if (slots[r][i][0] != '-') fprintf(stderr, " %s %.0f", slots[r][i], acc[i]);
      double m5 = 0, m6 = 0, m7 = 0;
      for (int b = 0; b < grid; ++b) {
        m5 += static_cast<double>(h[(b * 4 + r) * 8 + 5]) / grid;
        m6 += static_cast<double>(h[(b * 4 + r) * 8 + 6]) / grid;
        m7 += static_cast<double>(h[(b * 4 + r) * 8 + 7]) / grid;
      }
      if (r == 0) fprintf(stderr, " first_tma@%.0f", m5);
      if (r == 1) fprintf(stderr, " first_mma@%.0f last_commit@%.0f", m5, m6);
      if (r == 2) fprintf(stderr, " setup_done@%.0f epi_drained@%.0f cta_end@%.0f", m7, m5, m6);
    }
    unsigned long long t0 = ~0ull, t0max = 0, t1 = 0, t1min = ~0ull;
    for (int b = 0; b < grid; ++b) {
      const unsigned long long a = h[(b * 4 + 3) * 8 + 5], e = h[(b * 4 + 3) * 8 + 6];
      t0 = std::min(t0, a), t0max = std::max(t0max, a), t1 = std::max(t1, e), t1min = std::min(t1min, e);
    }
    fprintf(stderr, " | wall: first CTA start -> last CTA end %.1f us, start skew %.1f us, end skew %.1f us",
            (t1 - t0) * 1e-3, (t0max - t0) * 1e-3, (t1 - t1min) * 1e-3);
    fprintf(stderr, "\n");
  }
#endif
  char name[96];
  char sk[24] = "";
  if (args.splits > 1) snprintf(sk, sizeof(sk), ",splitK=%d", args.splits);
  snprintf(name, sizeof(name), "gemm_tcgen05<BN=%d,%s,%s,B=%s%s%s>", BN, MODE == 2 ? "TF32+BF16x2" : MODE ? "3xTF32" : "TF32",
           A_KIND == 3 ? "batched" : A_KIND == 2 ? "window" : A_KIND ? "im2col" : "tiled", B_MAJOR ? "MN" : "K", sk,
           EPI2 ? ",epi2" : "");
  ctx->last_kernel = name;
  return RTG_OK;
}


template <int A_KIND, int B_MAJOR>
rtg_status launch_variants(Context* ctx, int bn, int mode, const CUtensorMap& ma, const CUtensorMap& mb,
                           const CUtensorMap& mb2, const CUtensorMap& md, const CUtensorMap& mr, const GemmArgs& args,
                           const ReduceArgs* redp) {
  if (mode == RTG_MATH_TF32) {
    if (bn == 64) return launch<64, 0, A_KIND, B_MAJOR>(ctx, ma, mb, mb2, md, mr, args, redp);
    if (bn == 128) return launch<128, 0, A_KIND, B_MAJOR>(ctx, ma, mb, mb2, md, mr, args, redp);
    if (bn == 256) return launch<256, 0, A_KIND, B_MAJOR>(ctx, ma, mb, mb2, md, mr, args, redp);
  } else if (mode == RTG_MATH_TF32_BF16X2 && A_KIND != 3) {  // batched products keep the three-MMA split
    // this mode reads B through its packed K-major copy: there is no MN-major variant
    if constexpr (B_MAJOR == 0) {
      // tiles with few k-blocks are paced by their epilogue: two staging buffers per warp (GemmCfg EPI2)
      static const int epi2_kb = []() {
        const char* e = getenv("RTG_EPI2_KB");  // developer switch: 0 disables
        // measured on the ResNet-50 layer table: 4 beats 0, 8 and 16; never above kChunkKb (the variant accumulates a
        // unit in one chunk, which is the fp32-accurate modes' summation only up to that depth)
        return std::min(e ? atoi(e) : 4, kChunkKb);
      }();
      const int kb_per_unit = args.splits > 1 ? args.kb_per_split : args.num_kb;
      if (kb_per_unit <= epi2_kb && (args.flags & kFlagStagedD)) {
        if (bn == 64) return launch<64, 2, A_KIND, 0, true>(ctx, ma, mb, mb2, md, mr, args, redp);
        if (bn == 128) return launch<128, 2, A_KIND, 0, true>(ctx, ma, mb, mb2, md, mr, args, redp);
      }
      if (bn == 64) return launch<64, 2, A_KIND, 0>(ctx, ma, mb, mb2, md, mr, args, redp);
      if (bn == 128) return launch<128, 2, A_KIND, 0>(ctx, ma, mb, mb2, md, mr, args, redp);
    } else {
      return set_error(RTG_ERR_INTERNAL, "TF32+BF16x2 launches take the packed (K-major) B operand");
    }
  } else {
    if (bn == 64) return launch<64, 1, A_KIND, B_MAJOR>(ctx, ma, mb, mb2, md, mr, args, redp);
    if (bn == 128) return launch<128, 1, A_KIND, B_MAJOR>(ctx, ma, mb, mb2, md, mr, args, redp);
  }
  return set_error(RTG_ERR_INTERNAL, "bad BN");
}
#else
extern template rtg_status launch_variants<0, 0>(Context*, int, int, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const GemmArgs&, const ReduceArgs*);
extern template rtg_status launch_variants<0, 1>(Context*, int, int, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const GemmArgs&, const ReduceArgs*);
extern template rtg_status launch_variants<1, 0>(Context*, int, int, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const GemmArgs&, const ReduceArgs*);
extern template rtg_status launch_variants<2, 0>(Context*, int, int, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const GemmArgs&, const ReduceArgs*);
extern template rtg_status launch_variants<3, 0>(Context*, int, int, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const GemmArgs&, const ReduceArgs*);
extern template rtg_status launch_variants<3, 1>(Context*, int, int, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const GemmArgs&, const ReduceArgs*);
#endif

}  // namespace rtg
