"""Developer experiment: BERT's 768 -> 3072 layer as a pointwise convolution with each fused activation.
Usage: python scripts/exp_ffn.py [cin cout]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import ops

cin = int(sys.argv[1]) if len(sys.argv) > 1 else 768
cout = int(sys.argv[2]) if len(sys.argv) > 2 else 3072
ctx = R.default_context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
ctx.set_math(R.MATH_TF32_BF16X2)
x = (torch.rand((64, cin, 128, 1), device="cuda") - 0.5).contiguous(memory_format=torch.channels_last)
w = (torch.rand((cout, cin, 1, 1), device="cuda") - 0.5) * 0.1
b = torch.rand((cout,), device="cuda")
out = torch.empty((64, cout, 128, 1), device="cuda").contiguous(memory_format=torch.channels_last)
for act in ("none", "relu", "gelu", "sigmoid"):
    kw = dict(padding=(0, 0, 0, 0), activation=act, out=out, ctx=ctx)
    for _ in range(3):
        ops.conv(x, w, b, **kw)
    ctx.set_kernel_timing(True)
    for _ in range(10):
        ops.conv(x, w, b, **kw)
    kt = ctx.kernel_timing()
    ctx.set_kernel_timing(False)
    us = kt["ms"] / kt["launches"] * 1e3
    print(f"{act:8s} {us:8.1f} us  {2.0 * 8192 * cin * cout / us / 1e6:7.1f} TF/s  {ctx.last_kernel}", flush=True)
    if os.environ.get("RTG_PROF_LIB") == "1":
        os.environ["RTG_PROF_PRINT"] = "1"
        ops.conv(x, w, b, **kw)
        torch.cuda.synchronize()
        del os.environ["RTG_PROF_PRINT"]
