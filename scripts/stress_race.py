"""Developer aid: the contraction kernel is deterministic, so any run-to-run difference is a pipeline race.  Repeats a
few split-K / multi-tile convolutions and GEMMs, interleaved with unrelated launches (programmatic dependent launch
overlaps their tails), and compares every result bitwise with the first.  Usage: python scripts/stress_race.py [iters]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import ops

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
ctx = R.default_context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
torch.manual_seed(0)
cl = torch.channels_last
cases = []
for (B, C, O, H, k, s) in [(4, 512, 512, 7, 3, 1), (4, 256, 256, 14, 3, 1), (4, 1024, 256, 14, 1, 1), (32, 256, 128, 56, 1, 1),
                           (32, 128, 128, 28, 3, 1)]:
    p = k // 2
    x = (torch.rand((B, C, H, H), device="cuda") - 0.4).contiguous(memory_format=cl)
    w = torch.rand((O, C, k, k), device="cuda") - 0.5
    b = torch.rand((O,), device="cuda")
    oh = (H + 2 * p - k) // s + 1
    r = torch.rand((B, O, oh, oh), device="cuda").contiguous(memory_format=cl)
    cases.append((f"conv {C}->{O} {H}x{H} k{k}", lambda x=x, w=w, b=b, p=p, s=s, r=r: ops.conv(
        x, w, b, padding=(p, p, p, p), groups=1, strides=(s, s), residual=r, activation="relu", out_nhwc=True, ctx=ctx)))
for n, kk in [(1024, 1024), (257, 1000), (2048, 512)]:
    a, bm = torch.rand((n, kk), device="cuda") - 0.5, torch.rand((kk, n), device="cuda") - 0.5
    cases.append((f"gemm {n}x{n}x{kk}", lambda a=a, bm=bm: ops.matmul(a, bm, ctx=ctx)))
# the TMA-staged depthwise kernel (row ring, warp-granular slot release) and the gemv kernels
for (B, C, H, s) in [(8, 96, 56, 2), (8, 144, 28, 1), (4, 960, 7, 1)]:
    x = (torch.rand((B, C, H, H), device="cuda") - 0.4).contiguous(memory_format=cl)
    w = torch.rand((C, 1, 3, 3), device="cuda") - 0.5
    b = torch.rand((C,), device="cuda")
    cases.append((f"depthwise {C}ch {H}x{H} s{s}", lambda x=x, w=w, b=b, C=C, s=s: ops.conv(
        x, w, b, padding=(1, 1, 1, 1), groups=C, strides=(s, s), activation="clip", clip=(0.0, 6.0), out_nhwc=True, ctx=ctx)))
for (m, n, kk, tb) in [(1, 4096, 512, False), (8, 1000, 2048, True), (3, 1000, 2048, True)]:
    a = torch.rand((m, kk), device="cuda") - 0.5
    bm = (torch.rand((n, kk), device="cuda") - 0.5).t() if tb else torch.rand((kk, n), device="cuda") - 0.5
    cases.append((f"gemv {m}x{n}x{kk}", lambda a=a, bm=bm: ops.matmul(a, bm, ctx=ctx)))
sa, sb = torch.rand((64, 64), device="cuda"), torch.rand((64, 64), device="cuda")
bad_total = 0
for mode in (sys.argv[2].split(",") if len(sys.argv) > 2 else ("3xtf32", "tf32_bf16x2", "tf32")):
    ctx.set_math({"3xtf32": R.MATH_3XTF32, "tf32": R.MATH_TF32, "tf32_bf16x2": R.MATH_TF32_BF16X2}[mode])
    for name, fn in cases:
        ref = fn().clone()
        bad, worst = 0, 0.0
        for i in range(iters):
            if i % 3 == 0:
                ops.matmul(sa, sb, ctx=ctx)
            if i % 7 == 0:
                torch.cuda.synchronize()
            out = fn()
            if not torch.equal(out, ref):
                bad += 1
                worst = max(worst, float((out - ref).abs().max()))
        print(f"{mode:12s} {name:28s} {bad}/{iters} runs differ  max |diff| {worst:.3e}  [{ctx.last_kernel}]", flush=True)
        bad_total += bad
print("TOTAL", bad_total)
