"""Per-layer timing of the contraction kernel on ResNet-50's distinct convolution shapes (batch 32), device-resident
NHWC activations.  Prints kernel-only time (CUDA events around the launch, ctx kernel timing), achieved algorithmic
TFLOP/s and GB/s.  Usage: python scripts/bench_layers.py [tf32|3xtf32] [filter]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import ops

# (name, cin, cout, k, stride, pad, in_hw, residual)
LAYERS = [
    ("conv1 7x7/2", 3, 64, 7, 2, 3, 224, False),
    ("l1 1x1 64->64", 64, 64, 1, 1, 0, 56, False),
    ("l1 3x3 64->64", 64, 64, 3, 1, 1, 56, False),
    ("l1 1x1 64->256 +res", 64, 256, 1, 1, 0, 56, True),
    ("l1 1x1 256->64", 256, 64, 1, 1, 0, 56, False),
    ("l2 1x1 256->128", 256, 128, 1, 1, 0, 56, False),
    ("l2 3x3/2 128->128", 128, 128, 3, 2, 1, 56, False),
    ("l2 1x1 128->512 +res", 128, 512, 1, 1, 0, 28, True),
    ("l2 1x1/2 256->512", 256, 512, 1, 2, 0, 56, False),
    ("l2 1x1 512->128", 512, 128, 1, 1, 0, 28, False),
    ("l2 3x3 128->128", 128, 128, 3, 1, 1, 28, False),
    ("l3 3x3 256->256", 256, 256, 3, 1, 1, 14, False),
    ("l3 1x1 256->1024 +res", 256, 1024, 1, 1, 0, 14, True),
    ("l3 1x1 1024->256", 1024, 256, 1, 1, 0, 14, False),
    ("l4 3x3 512->512", 512, 512, 3, 1, 1, 7, False),
    ("l4 1x1 512->2048 +res", 512, 2048, 1, 1, 0, 7, True),
    ("l4 1x1 2048->512", 2048, 512, 1, 1, 0, 7, False),
]


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "3xtf32"
    filt = sys.argv[2] if len(sys.argv) > 2 else ""
    ctx = R.default_context(0)
    ctx.set_math(R.MATH_3XTF32 if mode == "3xtf32" else R.MATH_TF32)
    B = 32
    tot = 0.0
    for name, cin, cout, k, s, p, hw, res in LAYERS:
        if filt and filt not in name:
            continue
        x = torch.rand((B, cin, hw, hw), device="cuda").contiguous(memory_format=torch.channels_last)
        w = torch.rand((cout, cin, k, k), device="cuda") - 0.5
        b = torch.rand((cout,), device="cuda")
        ohw = (hw + 2 * p - k) // s + 1
        out = torch.empty((B, cout, ohw, ohw), device="cuda").contiguous(memory_format=torch.channels_last)
        r = torch.rand((B, cout, ohw, ohw), device="cuda").contiguous(memory_format=torch.channels_last) if res else None
        kw = dict(padding=(p, p, p, p), strides=(s, s), residual=r, activation="relu", out=out, ctx=ctx)
        for _ in range(3):
            ops.conv(x, w, b, **kw)
        ctx.set_kernel_timing(True)
        n = 10
        for _ in range(n):
            ops.conv(x, w, b, **kw)
        kt = ctx.kernel_timing()
        ctx.set_kernel_timing(False)
        us = kt["ms"] / kt["launches"] * 1e3
        tot += us
        print(f"{name:24s} {us:8.1f} us  {kt['flops'] / kt['launches'] / us / 1e6:7.1f} TF/s  "
              f"{kt['bytes'] / kt['launches'] / us / 1e3:7.1f} GB/s  {ctx.last_kernel}")
    print(f"sum {tot:.1f} us")


if __name__ == "__main__":
    main()
