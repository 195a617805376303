"""Accuracy / speed probe of RTG_MATH_TF32_BF16X2 against float64 (run under gpurun)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import rten_fork_b200 as R
from rten_fork_b200 import ops
ctx = R.default_context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
g = torch.Generator(device="cuda").manual_seed(1)
for mode, name in ((R.MATH_3XTF32, "3xtf32"), (R.MATH_TF32_BF16X2, "tf32+bf16x2")):
    ctx.set_math(mode)
    for (m, n, k) in ((128, 128, 32), (256, 128, 256), (512, 512, 4096), (2048, 2048, 2048)):
        for signed in (False, True):
            a = torch.rand((m, k), device="cuda", generator=g) - (0.5 if signed else 0)
            b = torch.rand((k, n), device="cuda", generator=g) - (0.5 if signed else 0)
            ref = a.double() @ b.double()
            scale = (a.double().abs() @ b.double().abs())
            for lab, bb in (("B[K,N]", b), ("B[N,K]^T", b.t().contiguous().t())):
                out = torch.empty((m, n), device="cuda")
                ops.matmul(a, bb, out=out, ctx=ctx)
                torch.cuda.synchronize()
                err = (out.double() - ref).abs()
                print(f"{name:12s} {m}x{n}x{k} signed={signed} {lab:9s} max|err|/sum|ab| {float((err / scale).max()):.2e}  "
                      f"max rel {float((err / ref.abs().clamp_min(1e-30)).max()):.2e}  [{ctx.last_kernel}]", flush=True)
    for n in (1024, 4096, 8192):
        a = torch.rand((n, n), device="cuda", generator=g); b = torch.rand((n, n), device="cuda", generator=g)
        out = torch.empty((n, n), device="cuda")
        for _ in range(3): ops.matmul(a, b, out=out, ctx=ctx)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        it = 20 if n <= 4096 else 5
        e0.record()
        for _ in range(it): ops.matmul(a, b, out=out, ctx=ctx)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / it
        print(f"{name:12s} GEMM {n}^3: {ms*1e3:.1f} us  {2*n**3/ms/1e9:.1f} TFLOP/s [{ctx.last_kernel}]", flush=True)
