"""First-light GPU check of the tcgen05 GEMM (run under gpurun)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import rten_fork_b200 as R

ctx = R.Context(0)
ex = R.GemmExecutor(ctx)
rng = np.random.default_rng(0)

def run(m, n, k, mode, b_kmajor=False, bias=False, alpha=1.0, beta=0.0, signed=False):
    ctx.set_math(mode)
    a = rng.random((m, k), dtype=np.float32)
    if b_kmajor:
        b = rng.random((n, k), dtype=np.float32).T
    else:
        b = rng.random((k, n), dtype=np.float32)
    if signed:
        a = a - 0.5; b = b - 0.5
    bv = rng.random(m, dtype=np.float32) if bias else None
    out = np.full((m, n), np.nan, dtype=np.float32) if beta == 0 else rng.random((m, n), dtype=np.float32)
    ref = alpha * (a.astype(np.float64) @ b.astype(np.float64)) + (beta * out.astype(np.float64) if beta else 0) + (bv[:, None] if bias else 0)
    ex.gemm_bias(out.reshape(-1), n, a, b, alpha, beta, bv)
    err = np.abs(out - ref)
    rel = (err / (np.abs(ref) + 1e-30)).max() if not signed else err.max() / np.abs(ref).max()
    print(f"m={m} n={n} k={k} mode={mode} bK={b_kmajor} bias={bias} a={alpha} b={beta} signed={signed} -> maxrel {rel:.3e} nan={np.isnan(out).sum()} [{ctx.last_kernel}]", flush=True)
    return rel

if os.environ.get("RTG_DEBUG_FLAGS"):
    for k in (32, 256, 4096):
        run(256, 128, k, R.MATH_3XTF32, True)
    sys.exit(0)
for k in (128, 512, 2048, 4608, 16384):
    run(512, 256, k, R.MATH_3XTF32, True)
for mode in (R.MATH_TF32, R.MATH_3XTF32):
    for bk in (True, False):
        run(128, 64, 32, mode, bk)
        run(128, 256, 64, mode, bk)
        run(256, 512, 1024, mode, bk)
        run(1000, 1000, 1000, mode, bk, bias=True, alpha=0.5, beta=2.0)
        run(77, 50, 19, mode, bk)
        run(1024, 1024, 4096, mode, bk, signed=True)
    run(2048, 2048, 16384, mode)

# timing, device resident
for mode, name in ((R.MATH_TF32, "tf32"), (R.MATH_3XTF32, "3xtf32")):
    ctx.set_math(mode)
    for n in (1024, 4096, 8192):
        a = torch.rand(n, n, device="cuda"); b = torch.rand(n, n, device="cuda"); bt = b.t().contiguous().t()
        o = torch.empty(n, n, device="cuda")
        torch.cuda.synchronize()
        for lab, bb in (("B[K,N]", b), ("B[N,K]^T", bt)):
            for _ in range(3): ex.gemm(o.view(-1), n, a, bb)
            ctx.sync(); t = time.perf_counter(); it = 20
            for _ in range(it): ex.gemm(o.view(-1), n, a, bb)
            ctx.sync(); dt = (time.perf_counter() - t) / it
            ref = a.double() @ b.double()
            rel = ((o.double() - ref).abs() / ref.abs()).max().item()
            print(f"{name} n={n} {lab}: {dt*1e3:.3f} ms  {2*n**3/dt/1e12:.1f} TFLOP/s  maxrel {rel:.2e} [{ctx.last_kernel}]", flush=True)
# torch reference timing
for n in (4096, 8192):
    a = torch.rand(n, n, device="cuda"); b = torch.rand(n, n, device="cuda")
    for tf in (False, True):
        torch.backends.cuda.matmul.allow_tf32 = tf
        for _ in range(3): a @ b
        torch.cuda.synchronize(); t = time.perf_counter()
        for _ in range(20): a @ b
        torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 20
        print(f"torch matmul n={n} tf32={tf}: {2*n**3/dt/1e12:.1f} TFLOP/s")
