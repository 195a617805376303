"""Developer aid: per-role cycle counters (make PROF=1 build) of one square GEMM.
Usage: RTG_PROF_LIB=1 python scripts/prof_gemm.py [n] [mode]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import ops

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
mode = sys.argv[2] if len(sys.argv) > 2 else "tf32_bf16x2"
ctx = R.default_context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
ctx.set_math({"3xtf32": R.MATH_3XTF32, "tf32": R.MATH_TF32, "tf32_bf16x2": R.MATH_TF32_BF16X2}[mode])
a, b = torch.rand((n, n), device="cuda"), torch.rand((n, n), device="cuda")
if "bt" in sys.argv:
    b = b.t()  # K-major B: one TMA box per k-block instead of BN/32
out = torch.empty((n, n), device="cuda")
for _ in range(3):
    ops.matmul(a, b, out=out, ctx=ctx)
torch.cuda.synchronize()
os.environ["RTG_PROF_PRINT"] = "1"
ops.matmul(a, b, out=out, ctx=ctx)
torch.cuda.synchronize()
