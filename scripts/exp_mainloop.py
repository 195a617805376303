"""Steady-state cost of one k-block of the contraction main loop: one 128 x BN tile per CTA (M = 128 * 148), K = 16384
(512 k-blocks per tile), so fill, drain and the epilogue are negligible.  Usage: python scripts/exp_mainloop.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import ops

ctx = R.default_context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
M, K = 128 * 148, 16384
for mode, name in ((R.MATH_TF32_BF16X2, "tf32_bf16x2"), (R.MATH_3XTF32, "3xtf32"), (R.MATH_TF32, "tf32")):
    ctx.set_math(mode)
    for N in (64, 128, 256):
        a = torch.rand((M, K), device="cuda") - 0.5
        b = (torch.rand((N, K), device="cuda") - 0.5).t()   # K-major B
        out = torch.empty((M, N), device="cuda")
        for _ in range(2):
            ops.matmul(a, b, out=out, ctx=ctx)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            ops.matmul(a, b, out=out, ctx=ctx)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 5 * 1e3
        import subprocess
        mhz = float(subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True).stdout.split()[0])
        kbs = K // 32
        print(f"{name:12s} N={N:4d} {ctx.last_kernel:50s} {us:9.1f} us  {2.0 * M * N * K / us / 1e6:7.1f} TF/s  "
              f"{us * 1e-6 * 1.965e9 / kbs:7.1f} cycles / k-block at 1965 MHz (idle clock now {mhz:.0f})")
        del a, b, out
