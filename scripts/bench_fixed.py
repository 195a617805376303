"""Fixed-cost probe: one tile per CTA (M = 128 * 148), sweep K.  Kernel-only time from ctx kernel timing."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import ops

ctx = R.default_context(0)
mode = sys.argv[1] if len(sys.argv) > 1 else "tf32"
ctx.set_math(R.MATH_TF32 if mode == "tf32" else R.MATH_3XTF32)
for tiles_per_cta in (1, 4):
    for N in (64, 128, 256):
        for K in (32, 64, 128, 256, 512, 1024, 2048, 4096):
            M = 128 * 148 * tiles_per_cta
            a = torch.rand((M, K), device="cuda")
            b = torch.rand((K, N), device="cuda").t().contiguous().t()   # K-major B ([N][K] storage)
            out = torch.empty((M, N), device="cuda")
            for _ in range(2):
                ops.matmul(a, b, out=out, ctx=ctx)
            ctx.set_kernel_timing(True)
            for _ in range(5):
                ops.matmul(a, b, out=out, ctx=ctx)
            kt = ctx.kernel_timing()
            ctx.set_kernel_timing(False)
            us = kt["ms"] / kt["launches"] * 1e3
            print(f"tiles/cta {tiles_per_cta} N {N:4d} K {K:5d}: {us:8.1f} us  {us / (K / 32) / tiles_per_cta:6.3f} us/kblock  "
                  f"{kt['flops'] / kt['launches'] / us / 1e6:7.1f} TF/s {ctx.last_kernel}")
