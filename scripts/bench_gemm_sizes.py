"""Square GEMM timing per math mode (device-resident): python scripts/bench_gemm_sizes.py [sizes...]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import ops

sizes = [int(x) for x in sys.argv[1:]] or [2048, 4096, 8192, 16384]
ctx = R.default_context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
for n in sizes:
    a, b = torch.rand((n, n), device="cuda"), torch.rand((n, n), device="cuda")
    c = torch.empty((n, n), device="cuda")
    row = []
    for name, m in (("tf32_bf16x2", R.MATH_TF32_BF16X2), ("3xtf32", R.MATH_3XTF32), ("tf32", R.MATH_TF32)):
        ctx.set_math(m)
        for _ in range(2):
            ops.matmul(a, b, out=c, ctx=ctx)
        torch.cuda.synchronize()
        it = 3 if n >= 8192 else 10
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(it):
            ops.matmul(a, b, out=c, ctx=ctx)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / it * 1e3
        row.append(f"{name} {us:9.1f} us {2.0 * n ** 3 / us / 1e6:6.1f} TF/s")
    print(n, " | ".join(row), ctx.last_kernel, flush=True)
