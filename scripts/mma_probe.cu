// Developer probe (not part of the library): issue / pipe rates of tcgen05.mma on this GPU, the numbers the contraction
// kernel's main loop is designed around.  One CTA (or one 2-CTA cluster) per launch, one issuing thread, operands are
// whatever is in shared / tensor memory.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I rten_fork_b200/csrc scripts/mma_probe.cu -o scripts/mma_probe
#include "sm100_ptx.cuh"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

using namespace rtg;

#define CK(x)                                                                        \
  do {                                                                               \
    cudaError_t e_ = (x);                                                            \
    if (e_ != cudaSuccess) {                                                         \
      printf("%s failed: %s\n", #x, cudaGetErrorString(e_));                         \
      exit(1);                                                                       \
    }                                                                                \
  } while (0)

enum Form { SS_TF32 = 0, TS_TF32 = 1, TS_F16 = 2, ALT_TS = 3, GRP_TS = 4, SS_F16 = 5, DUAL_TS_TF32 = 6, DUAL_ALT = 7, DUAL_SS_TF32 = 8, QUAD_TS_TF32 = 9 };

__device__ __forceinline__ void mma2_tf32_ts(uint32_t d, uint32_t a_t, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a_t), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void mma2_f16_ts(uint32_t d, uint32_t a_t, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a_t), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void mma2_tf32_ss(uint32_t d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void mma_f16_ss(uint32_t d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void commit2(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

template <int form>
__device__ __forceinline__ void issue1(int i, uint32_t d, uint32_t a_t, uint64_t adesc, uint64_t bdesc,
                                       uint32_t id_tf32, uint32_t id_bf16, uint32_t dn = 128) {
  const uint32_t bl = static_cast<uint32_t>(bdesc), bh = static_cast<uint32_t>(bdesc >> 32);
  const uint32_t al = static_cast<uint32_t>(adesc), ah = static_cast<uint32_t>(adesc >> 32);
  const uint32_t ks = (i & 3) * 2;  // walk the four k-steps of a 128-byte row
  switch (form) {
    case SS_TF32: ptx::mma_tf32_ss_w(d, al + ks, ah, bl + ks, bh, id_tf32, 1); break;
    case TS_TF32: ptx::mma_tf32_ts_w(d, a_t + (i & 3) * 8, bl + ks, bh, id_tf32, 1); break;
    case TS_F16: ptx::mma_f16_ts_w(d, a_t + (i & 3) * 8, bl + ks, bh, id_bf16, 1); break;
    case ALT_TS:
      if (i & 1) ptx::mma_tf32_ts_w(d, a_t + (i & 3) * 8, bl + ks, bh, id_tf32, 1);
      else ptx::mma_f16_ts_w(d, a_t + 32 + (i & 3) * 8, bl + 1024 + ks, bh, id_bf16, 1);
      break;
    case GRP_TS:
      if (i & 4) ptx::mma_tf32_ts_w(d, a_t + (i & 3) * 8, bl + ks, bh, id_tf32, 1);
      else ptx::mma_f16_ts_w(d, a_t + 32 + (i & 3) * 8, bl + 1024 + ks, bh, id_bf16, 1);
      break;
    case SS_F16: mma_f16_ss(d, adesc + ks, bdesc + ks, id_bf16, 1); break;
    // independent accumulators, interleaved (dn = columns per accumulator)
    case DUAL_TS_TF32: ptx::mma_tf32_ts_w(d + (i & 1) * dn, a_t + ((i >> 1) & 3) * 8, bl + ks, bh, id_tf32, 1); break;
    case DUAL_SS_TF32: ptx::mma_tf32_ss_w(d + (i & 1) * dn, al + ks, ah, bl + ks, bh, id_tf32, 1); break;
    case QUAD_TS_TF32: ptx::mma_tf32_ts_w(d + (i & 3) * dn, a_t + ((i >> 2) & 3) * 8, bl + ks, bh, id_tf32, 1); break;
    case DUAL_ALT:  // TF32 hi*hi into accumulator 0, bf16 corrections into accumulator 1
      if (i & 1) ptx::mma_tf32_ts_w(d, a_t + ((i >> 1) & 3) * 8, bl + ks, bh, id_tf32, 1);
      else ptx::mma_f16_ts_w(d + dn, a_t + 32 + ((i >> 1) & 3) * 8, bl + 1024 + ks, bh, id_bf16, 1);
      break;
  }
}

// mode 0: throughput of a chain of `reps` MMAs; mode 1: clock after every issue (queue depth); mode 2: pipeline skeleton
// (a producer warp arrives on `full`, the MMA thread waits `nwait` barriers, issues `per_round` MMAs, commits `empty`)
template <int N, int form>
__global__ void __launch_bounds__(128, 1) probe1(int mode, int reps, int per_round, int nwait, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = base + 160 * 1024;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) ptx::sts_f32(base + i * 4, 1.0f);
  const int S = (mode == 6 && nwait == 0) ? 8 : 4;
  constexpr int SMAX = 8;
  if (threadIdx.x == 0) {
    for (int s = 0; s < 3 * SMAX + 2; ++s) ptx::mbar_init(bars + 8 * s, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tm = tmem_slot;
  const uint32_t id_tf32 = ptx::make_idesc_tf32(128, N, 0, 0), id_bf16 = ptx::make_idesc_bf16(128, N);
  const uint64_t adesc = ptx::make_smem_desc(base, 16, 1024);
  const uint64_t bdesc = ptx::make_smem_desc(base + 32 * 1024, 16, 1024);
  const uint32_t done = bars;  // barrier 0
  auto full = [&](int s) { return bars + 8 * (1 + s); };
  auto empty = [&](int s) { return bars + 8 * (1 + SMAX + s); };
  auto full2 = [&](int s) { return bars + 8 * (2 * SMAX + 2 + s); };
  if ((mode == 2 || mode == 3 || mode == 4 || mode == 6) && warp == 1) {  // producer skeleton
    int s = 0;
    uint32_t ph = 0;
    const int rounds = reps / per_round;
    for (int r = 0; r < rounds; ++r) {
      ptx::mbar_wait(empty(s), ph ^ 1u);
      if (lane == 0) {
        ptx::mbar_arrive(full(s));
        ptx::mbar_arrive(full2(s));
      }
      __syncwarp();
      if (++s == S) s = 0, ph ^= 1u;
    }
  }
  if (warp == 0) {
    const uint32_t d = tm, a_t = tm + 256;
    if (mode == 0) {
      long long t0 = 0, t1 = 0;
      for (int pass = 0; pass < 2; ++pass) {  // pass 0 warms up
        __syncwarp();
        t0 = clock64();
        if (ptx::elect_one()) {
          _Pragma("unroll 8") for (int i = 0; i < reps; ++i) issue1<form>(i, d, a_t, adesc, bdesc, id_tf32, id_bf16, N);
          ptx::mma_commit(done);
        }
        __syncwarp();
        ptx::mbar_wait(done, pass & 1);
        t1 = clock64();
      }
      if (lane == 0) out[0] = t1 - t0;
    } else if (mode == 1) {
      if (lane == 0) {
        const long long t0 = clock64();
        for (int i = 0; i < reps && i < 64; ++i) {
          issue1<form>(i, d, a_t, adesc, bdesc, id_tf32, id_bf16, N);
          out[1 + i] = clock64() - t0;
        }
        ptx::mma_commit(done);
        ptx::mbar_wait(done, 0);
        out[0] = clock64() - t0;
      }
    } else if (mode == 3) {  // skeleton with per-phase clocks: wait / fence / issue / commit
      int s = 0;
      uint32_t ph = 0;
      const int rounds = reps / per_round;
      long long tw = 0, tf = 0, ti = 0, tc = 0;
      __syncwarp();
      const long long t0 = clock64();
      long long t = t0;
      for (int r = 0; r < rounds; ++r) {
        ptx::mbar_wait(full(s), ph);
        if (nwait > 1) ptx::mbar_wait(full2(s), ph);
        long long u = clock64(); tw += u - t; t = u;
        ptx::tc_fence_after();
        u = clock64(); tf += u - t; t = u;
        if (ptx::elect_one()) {
          _Pragma("unroll 8") for (int i = 0; i < per_round; ++i) issue1<form>(i, d, a_t, adesc, bdesc, id_tf32, id_bf16, N);
        }
        u = clock64(); ti += u - t; t = u;
        if (ptx::elect_one()) ptx::mma_commit(empty(s));
        __syncwarp();
        u = clock64(); tc += u - t; t = u;
        if (++s == S) s = 0, ph ^= 1u;
      }
      if (ptx::elect_one()) ptx::mma_commit(done);
      __syncwarp();
      ptx::mbar_wait(done, 0);
      if (lane == 0) out[0] = clock64() - t0, out[1] = tw, out[2] = tf, out[3] = ti, out[4] = tc;
    } else if (mode == 6) {  // variants: nwait 0: 8 stages; 1: a shared-memory load between commit and wait; 2: plain arrive, no commit
      int s = 0;
      uint32_t ph = 0;
      const int rounds = reps / per_round;
      long long tw = 0, tf = 0, ti = 0, tc = 0, tl = 0;
      float sink = 0.f;
      __syncwarp();
      const long long t0 = clock64();
      long long t = t0;
      for (int r = 0; r < rounds; ++r) {
        long long u;
        if (nwait == 1) {
          sink += ptx::lds_f32(base + lane * 4);
          if (sink == 12345.f) printf("x");
          u = clock64(); tl += u - t; t = u;
        }
        ptx::mbar_wait(full(s), ph);
        u = clock64(); tw += u - t; t = u;
        ptx::tc_fence_after();
        u = clock64(); tf += u - t; t = u;
        if (ptx::elect_one()) {
          _Pragma("unroll 8") for (int i = 0; i < per_round; ++i) issue1<form>(i, d, a_t, adesc, bdesc, id_tf32, id_bf16, N);
        }
        u = clock64(); ti += u - t; t = u;
        if (nwait == 2) {
          if (ptx::elect_one()) ptx::mbar_arrive(empty(s));
        } else {
          if (ptx::elect_one()) ptx::mma_commit(empty(s));
        }
        __syncwarp();
        u = clock64(); tc += u - t; t = u;
        if (++s == S) s = 0, ph ^= 1u;
      }
      if (ptx::elect_one()) ptx::mma_commit(done);
      __syncwarp();
      ptx::mbar_wait(done, 0);
      if (lane == 0) out[0] = clock64() - t0, out[1] = tw, out[2] = tf, out[3] = ti, out[4] = tc, out[5] = tl;
    } else if (mode == 4) {  // one thread runs the whole loop (no elect / syncwarp per round)
      const int rounds = reps / per_round;
      __syncwarp();
      const long long t0 = clock64();
      if (lane == 0) {
        int s = 0;
        uint32_t ph = 0;
        for (int r = 0; r < rounds; ++r) {
          ptx::mbar_wait(full(s), ph);
          if (nwait > 1) ptx::mbar_wait(full2(s), ph);
          ptx::tc_fence_after();
          _Pragma("unroll 8") for (int i = 0; i < per_round; ++i) issue1<form>(i, d, a_t, adesc, bdesc, id_tf32, id_bf16, N);
          ptx::mma_commit(empty(s));
          if (++s == S) s = 0, ph ^= 1u;
        }
        ptx::mma_commit(done);
      }
      __syncwarp();
      ptx::mbar_wait(done, 0);
      if (lane == 0) out[0] = clock64() - t0;
    } else if (mode == 5) {  // no producer round trip: issue + commit only (what the commits cost by themselves)
      const int rounds = reps / per_round;
      __syncwarp();
      const long long t0 = clock64();
      int s = 0;
      for (int r = 0; r < rounds; ++r) {
        if (ptx::elect_one()) {
          _Pragma("unroll 8") for (int i = 0; i < per_round; ++i) issue1<form>(i, d, a_t, adesc, bdesc, id_tf32, id_bf16, N);
          if (nwait > 0) ptx::mma_commit(full2(s));  // nobody waits on it
        }
        __syncwarp();
        if (++s == S) s = 0;
      }
      if (ptx::elect_one()) ptx::mma_commit(done);
      __syncwarp();
      ptx::mbar_wait(done, 0);
      if (lane == 0) out[0] = clock64() - t0;
    } else {
      int s = 0;
      uint32_t ph = 0;
      const int rounds = reps / per_round;
      __syncwarp();
      const long long t0 = clock64();
      for (int r = 0; r < rounds; ++r) {
        ptx::mbar_wait(full(s), ph);
        if (nwait > 1) ptx::mbar_wait(full2(s), ph);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          _Pragma("unroll 8") for (int i = 0; i < per_round; ++i) issue1<form>(i, d, a_t, adesc, bdesc, id_tf32, id_bf16, N);
          ptx::mma_commit(empty(s));
        }
        __syncwarp();
        if (++s == S) s = 0, ph ^= 1u;
      }
      if (ptx::elect_one()) ptx::mma_commit(done);
      __syncwarp();
      ptx::mbar_wait(done, 0);
      if (lane == 0) out[0] = clock64() - t0;
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tm, 512);
  }
}

// 2-CTA pair: M = 256 (128 rows per CTA), each CTA holds N/2 columns of B.  The leader CTA's thread issues.
template <int N, int form>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) probe2(int reps, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = base + 160 * 1024;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) ptx::sts_f32(base + i * 4, 1.0f);
  if (threadIdx.x == 0) {
    ptx::mbar_init(bars, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32(&tmem_slot)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  ptx::tc_fence_after();
  const uint32_t tm = tmem_slot;
  const uint32_t id_tf32 = ptx::make_idesc_tf32(256, N, 0, 0), id_bf16 = ptx::make_idesc_bf16(256, N);
  const uint64_t adesc = ptx::make_smem_desc(base, 16, 1024);
  const uint64_t bdesc = ptx::make_smem_desc(base + 32 * 1024, 16, 1024);
  if (rank == 0 && warp == 0) {
    const uint32_t d = tm, a_t = tm + 256;
    long long t0 = 0, t1 = 0;
    for (int pass = 0; pass < 2; ++pass) {
      __syncwarp();
      t0 = clock64();
      if (ptx::elect_one()) {
_Pragma("unroll 8")
        for (int i = 0; i < reps; ++i) {
          const uint32_t ks = (i & 3) * 2;
          if (form == TS_TF32) mma2_tf32_ts(d, a_t + (i & 3) * 8, bdesc + ks, id_tf32, 1);
          else if (form == TS_F16) mma2_f16_ts(d, a_t + (i & 3) * 8, bdesc + ks, id_bf16, 1);
          else if (form == ALT_TS) {
            if (i & 1) mma2_tf32_ts(d, a_t + (i & 3) * 8, bdesc + ks, id_tf32, 1);
            else mma2_f16_ts(d, a_t + 32 + (i & 3) * 8, bdesc + 1024 + ks, id_bf16, 1);
          } else mma2_tf32_ss(d, adesc + ks, bdesc + ks, id_tf32, 1);
        }
        commit2(bars);
      }
      __syncwarp();
      ptx::mbar_wait(bars, pass & 1);
      t1 = clock64();
    }
    if (lane == 0) out[0] = t1 - t0;
  }
  ptx::tc_fence_before();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 0) {
    ptx::tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512) : "memory");
  }
}

static const char* kFormName[] = {"SS tf32", "TS tf32", "TS bf16", "alt TS bf16/tf32", "4 bf16 + 4 tf32 TS", "SS bf16", "TS tf32, 2 accs", "bf16->acc1 tf32->acc0", "SS tf32, 2 accs", "TS tf32, 4 accs"};
constexpr int kSmem = 160 * 1024 + 1024 + 256;

template <int N, int form>
static void run1f(long long* dout, int mode, int reps, int per_round, int nwait) {
  CK(cudaFuncSetAttribute(probe1<N, form>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  CK(cudaMemset(dout, 0, 80 * 8));
  probe1<N, form><<<1, 128, kSmem>>>(mode, reps, per_round, nwait, dout);
  CK(cudaDeviceSynchronize());
  long long h[80];
  CK(cudaMemcpy(h, dout, sizeof(h), cudaMemcpyDeviceToHost));
  if (mode == 0) {
    printf("N=%3d %-20s chain of %d: %6.1f cycles / MMA\n", N, kFormName[form], reps, double(h[0]) / reps);
  } else if (mode == 1) {
    printf("N=%3d %-20s issue clocks:", N, kFormName[form]);
    for (int i = 0; i < 24; ++i) printf(" %lld", h[1 + i]);
    printf(" ... total %lld\n", h[0]);
  } else if (mode == 6) {
    const double r = reps / per_round;
    printf("N=%3d %-20s variant %d (%s): %2d MMAs / round: %7.1f cycles / round = lds %.1f + wait %.1f + fence %.1f + issue %.1f + commit %.1f\n",
           N, kFormName[form], nwait, nwait == 0 ? "8 stages" : nwait == 1 ? "lds before wait" : "plain arrive, no commit", per_round,
           h[0] / r, h[5] / r, h[1] / r, h[2] / r, h[3] / r, h[4] / r);
  } else if (mode == 3) {
    const double r = reps / per_round;
    printf("N=%3d %-20s skeleton phases: %2d MMAs / round, %d wait(s): %7.1f cycles / round = wait %.1f + fence %.1f + issue %.1f + commit %.1f\n",
           N, kFormName[form], per_round, nwait, h[0] / r, h[1] / r, h[2] / r, h[3] / r, h[4] / r);
  } else {
    printf("N=%3d %-20s %s: %2d MMAs / round, %d wait(s): %7.1f cycles / round (%5.1f / MMA)\n", N, kFormName[form],
           mode == 4 ? "one-thread skeleton" : mode == 5 ? "issue + commit only" : "pipeline skeleton", per_round, nwait,
           double(h[0]) / (reps / per_round), double(h[0]) / reps);
  }
}


// TMEM store throughput: the warps in `mask` each issue `reps` tcgen05.st.32x32b.x16 (2 KB each) into their lane quarter,
// optionally while one thread keeps the tensor pipe busy with TS-form MMAs (mma != 0).
__global__ void __launch_bounds__(320, 1) probe_st(int mask, int reps, int mma, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = base + 160 * 1024;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) ptx::sts_f32(base + i * 4, 1.0f);
  if (threadIdx.x == 0) {
    ptx::mbar_init(bars, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tm = tmem_slot;
  if (warp < 8 && ((mask >> warp) & 1)) {
    uint32_t r[16];
    for (int j = 0; j < 16; ++j) r[j] = lane * 16 + j;
    const uint32_t dst = tm + (static_cast<uint32_t>((warp & 3) * 32) << 16) + 256 + (warp >> 2) * 64;
    long long t0 = 0, t1 = 0;
    for (int pass = 0; pass < 2; ++pass) {
      __syncwarp();
      t0 = clock64();
#pragma unroll 4
      for (int i = 0; i < reps; ++i) ptx::tmem_st_32x16(dst + (i & 3) * 16, r);
      t1 = clock64();
      ptx::tmem_st_wait();
    }
    const long long t2 = clock64();
    if (lane == 0) out[warp * 2] = t1 - t0, out[warp * 2 + 1] = t2 - t0;
  } else if (warp == 8 && mma) {
    // mma > 0: N = 128 TS-form chain, mma < 0: N = 64; its own duration goes to out[20] (does TMEM store / load traffic
    // of the other warps lengthen the tensor pipe's time per instruction?)
    const int n_mma = mma > 0 ? mma : -mma;
    const uint32_t id_tf32 = mma > 0 ? ptx::make_idesc_tf32(128, 128, 0, 0) : ptx::make_idesc_tf32(128, 64, 0, 0);
    const uint64_t bdesc = ptx::make_smem_desc(base + 32 * 1024, 16, 1024);
    const long long t0 = clock64();
    if (ptx::elect_one()) {
      for (int i = 0; i < n_mma; ++i)
        ptx::mma_tf32_ts_w(tm, tm + 448 + (i & 3) * 8, static_cast<uint32_t>(bdesc) + (i & 3) * 2, static_cast<uint32_t>(bdesc >> 32), id_tf32, 1);
      ptx::mma_commit(bars);
    }
    __syncwarp();
    ptx::mbar_wait(bars, 0);
    if (lane == 0) out[20] = clock64() - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tm, 512);
  }
}

static void run_st(long long* dout, int mask, int reps, int mma) {
  CK(cudaFuncSetAttribute(probe_st, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  CK(cudaMemset(dout, 0, 80 * 8));
  probe_st<<<1, 320, kSmem>>>(mask, reps, mma, dout);
  CK(cudaDeviceSynchronize());
  long long h[24];
  CK(cudaMemcpy(h, dout, sizeof(h), cudaMemcpyDeviceToHost));
  if (mma) printf("[MMA chain N=%d: %.1f cycles / MMA] ", mma > 0 ? 128 : 64, double(h[20]) / (mma > 0 ? mma : -mma));
  printf("tcgen05.st.32x32b.x16 warps 0x%02x%s:", mask, mma ? " + MMA chain" : "");
  for (int w = 0; w < 8; ++w)
    if ((mask >> w) & 1) printf("  w%d %.1f issue / %.1f done", w, double(h[2 * w]) / reps, double(h[2 * w + 1]) / reps);
  printf("  cycles per store\n");
}

template <int N>
static void run1(long long* dout, int form, int mode, int reps, int per_round, int nwait) {
  switch (form) {
    case 0: run1f<N, 0>(dout, mode, reps, per_round, nwait); break;
    case 1: run1f<N, 1>(dout, mode, reps, per_round, nwait); break;
    case 2: run1f<N, 2>(dout, mode, reps, per_round, nwait); break;
    case 3: run1f<N, 3>(dout, mode, reps, per_round, nwait); break;
    case 4: run1f<N, 4>(dout, mode, reps, per_round, nwait); break;
    case 5: run1f<N, 5>(dout, mode, reps, per_round, nwait); break;
    case 6: run1f<N, 6>(dout, mode, reps, per_round, nwait); break;
    case 7: run1f<N, 7>(dout, mode, reps, per_round, nwait); break;
    case 8: run1f<N, 8>(dout, mode, reps, per_round, nwait); break;
    case 9: run1f<N, 9>(dout, mode, reps, per_round, nwait); break;
  }
}

template <int N>
static void run2(long long* dout, int form, int reps) {
  CK(cudaMemset(dout, 0, 80 * 8));
  switch (form) {
    case 0: CK(cudaFuncSetAttribute(probe2<N, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem)); probe2<N, 0><<<2, 128, kSmem>>>(reps, dout); break;
    case 1: CK(cudaFuncSetAttribute(probe2<N, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem)); probe2<N, 1><<<2, 128, kSmem>>>(reps, dout); break;
    case 2: CK(cudaFuncSetAttribute(probe2<N, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem)); probe2<N, 2><<<2, 128, kSmem>>>(reps, dout); break;
    case 3: CK(cudaFuncSetAttribute(probe2<N, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem)); probe2<N, 3><<<2, 128, kSmem>>>(reps, dout); break;
  }
  CK(cudaDeviceSynchronize());
  long long h[1];
  CK(cudaMemcpy(h, dout, sizeof(h), cudaMemcpyDeviceToHost));
  printf("2-CTA M=256 N=%3d %-20s chain of %d: %6.1f cycles / MMA\n", N, kFormName[form], reps, double(h[0]) / reps);
}

int main(int argc, char** argv) {
  long long* dout;
  CK(cudaMalloc(&dout, 80 * 8));
  const int reps = 512;
  if (argc > 1 && !strcmp(argv[1], "st")) {
    for (int mma : {0, 4096})
      for (int mask : {0x01, 0x11, 0x03, 0x0f, 0xff}) run_st(dout, mask, 256, mma);
    // the other direction: a long store stream (8 warps x 16384 stores) under a chain of 2048 MMAs
    for (int mma : {2048, -2048})
      for (int mask : {0x00, 0x0f, 0xff}) run_st(dout, mask, 16384, mma);
    printf("probe done\n");
    return 0;
  }
  if (argc > 1 && !strcmp(argv[1], "round")) {  // where a round's fixed cost goes
    for (int per_round : {8, 16}) {
      for (int nwait : {1, 2}) {
        run1<64>(dout, ALT_TS, 2, 1024, per_round, nwait);
        run1<64>(dout, ALT_TS, 3, 1024, per_round, nwait);
        run1<64>(dout, ALT_TS, 4, 1024, per_round, nwait);
        run1<128>(dout, ALT_TS, 2, 1024, per_round, nwait);
        run1<128>(dout, ALT_TS, 3, 1024, per_round, nwait);
        run1<128>(dout, ALT_TS, 4, 1024, per_round, nwait);
      }
      run1<64>(dout, ALT_TS, 5, 1024, per_round, 0);
      run1<64>(dout, ALT_TS, 5, 1024, per_round, 1);
      run1<128>(dout, ALT_TS, 5, 1024, per_round, 0);
      run1<128>(dout, ALT_TS, 5, 1024, per_round, 1);
    }
    for (int v : {0, 1, 2}) {
      run1<64>(dout, ALT_TS, 6, 1024, 8, v);
      run1<128>(dout, ALT_TS, 6, 1024, 8, v);
    }
    printf("probe done\n");
    return 0;
  }
  for (int form : {SS_TF32, TS_TF32, TS_F16, ALT_TS, GRP_TS, SS_F16}) {
    run1<64>(dout, form, 0, reps, 0, 0);
    run1<128>(dout, form, 0, reps, 0, 0);
    if (form != ALT_TS && form != GRP_TS) run1<256>(dout, form, 0, reps, 0, 0);
  }
  for (int form : {DUAL_TS_TF32, DUAL_ALT, DUAL_SS_TF32}) {
    run1<64>(dout, form, 0, reps, 0, 0);
    run1<128>(dout, form, 0, reps, 0, 0);
  }
  run1<64>(dout, QUAD_TS_TF32, 0, reps, 0, 0);
  run1<32>(dout, QUAD_TS_TF32, 0, reps, 0, 0);
  run1<32>(dout, TS_TF32, 0, reps, 0, 0);
  run1<192>(dout, TS_TF32, 0, reps, 0, 0);
  for (int per_round : {4, 8, 16}) run1<128>(dout, DUAL_ALT, 2, 1024, per_round, 1);
  for (int form : {SS_TF32, TS_TF32, ALT_TS, DUAL_ALT}) run1<128>(dout, form, 1, 64, 0, 0);
  for (int form : {SS_TF32, ALT_TS, GRP_TS})
    for (int per_round : {4, 8, 16, 32})
      for (int nwait : {1, 2}) run1<128>(dout, form, 2, 1024, per_round, nwait);
  for (int form : {TS_TF32, ALT_TS}) run1<64>(dout, form, 1, 64, 0, 0);
  for (int per_round : {4, 8, 16, 32})
    for (int nwait : {1, 2}) run1<64>(dout, ALT_TS, 2, 1024, per_round, nwait);
  run1<256>(dout, SS_TF32, 2, 1024, 4, 1);
  run1<256>(dout, SS_TF32, 2, 1024, 8, 1);
  for (int form : {SS_TF32, TS_TF32, TS_F16, ALT_TS}) {
    run2<64>(dout, form, reps);
    run2<128>(dout, form, reps);
    run2<256>(dout, form, reps);
  }
  printf("probe done\n");
  return 0;
}
