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


# MobileNetV2 (torchvision widths), batch 64: (name, cin, cout, k, stride, pad, in_hw, residual, groups, act)
MBV2 = [
    ("conv1 3x3/2 3->32", 3, 32, 3, 2, 1, 224, False, 1, "clip"),
    ("dw32 @112", 32, 32, 3, 1, 1, 112, False, 32, "clip"),
    ("pw 32->16 @112", 32, 16, 1, 1, 0, 112, False, 1, "none"),
    ("pw 16->96 @112", 16, 96, 1, 1, 0, 112, False, 1, "clip"),
    ("dw96 /2 @112", 96, 96, 3, 2, 1, 112, False, 96, "clip"),
    ("pw 96->24 @56", 96, 24, 1, 1, 0, 56, False, 1, "none"),
    ("pw 24->144 @56", 24, 144, 1, 1, 0, 56, False, 1, "clip"),
    ("dw144 @56", 144, 144, 3, 1, 1, 56, False, 144, "clip"),
    ("pw 144->24 @56 +res", 144, 24, 1, 1, 0, 56, True, 1, "none"),
    ("dw144 /2 @56", 144, 144, 3, 2, 1, 56, False, 144, "clip"),
    ("pw 144->32 @28", 144, 32, 1, 1, 0, 28, False, 1, "none"),
    ("pw 32->192 @28", 32, 192, 1, 1, 0, 28, False, 1, "clip"),
    ("dw192 @28", 192, 192, 3, 1, 1, 28, False, 192, "clip"),
    ("pw 192->32 @28 +res", 192, 32, 1, 1, 0, 28, True, 1, "none"),
    ("pw 64->384 @14", 64, 384, 1, 1, 0, 14, False, 1, "clip"),
    ("dw384 @14", 384, 384, 3, 1, 1, 14, False, 384, "clip"),
    ("pw 384->64 @14 +res", 384, 64, 1, 1, 0, 14, True, 1, "none"),
    ("pw 96->576 @14", 96, 576, 1, 1, 0, 14, False, 1, "clip"),
    ("dw576 @14", 576, 576, 3, 1, 1, 14, False, 576, "clip"),
    ("pw 576->96 @14 +res", 576, 96, 1, 1, 0, 14, True, 1, "none"),
    ("pw 160->960 @7", 160, 960, 1, 1, 0, 7, False, 1, "clip"),
    ("dw960 @7", 960, 960, 3, 1, 1, 7, False, 960, "clip"),
    ("pw 960->160 @7 +res", 960, 160, 1, 1, 0, 7, True, 1, "none"),
    ("pw 320->1280 @7", 320, 1280, 1, 1, 0, 7, False, 1, "clip"),
]


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "3xtf32"
    filt = sys.argv[2] if len(sys.argv) > 2 else ""
    if "--mbv2" in sys.argv:
        return run_table(mode, "", MBV2, 64)
    return run_table(mode, filt, [l + (1, "relu") for l in LAYERS], 32)


def run_table(mode, filt, layers, B):
    ctx = R.default_context(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ctx.set_math({"3xtf32": R.MATH_3XTF32, "tf32": R.MATH_TF32, "tf32_bf16x2": R.MATH_TF32_BF16X2}[mode])
    tot = 0.0
    tot_chain = 0.0
    for name, cin, cout, k, s, p, hw, res, groups, act in layers:
        if filt and filt not in name:
            continue
        x = torch.rand((B, cin, hw, hw), device="cuda").contiguous(memory_format=torch.channels_last)
        w = torch.rand((cout, cin // groups, k, k), device="cuda") - 0.5
        b = torch.rand((cout,), device="cuda")
        ohw = (hw + 2 * p - k) // s + 1
        out = torch.empty((B, cout, ohw, ohw), device="cuda").contiguous(memory_format=torch.channels_last)
        r = torch.rand((B, cout, ohw, ohw), device="cuda").contiguous(memory_format=torch.channels_last) if res else None
        kw = dict(padding=(p, p, p, p), groups=groups, strides=(s, s), residual=r, activation=act, clip=(0.0, 6.0),
                  out=out, ctx=ctx)
        for _ in range(3):
            ops.conv(x, w, b, **kw)
        ctx.set_kernel_timing(True)
        n = 10
        for _ in range(n):
            ops.conv(x, w, b, **kw)
        kt = ctx.kernel_timing()
        ctx.set_kernel_timing(False)
        if not kt["launches"]:  # not the contraction kernel (depthwise): events around back-to-back calls
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                ops.conv(x, w, b, **kw)
            e1.record()
            torch.cuda.synchronize()
            nbytes = 4.0 * (x.numel() + out.numel() * (2 if res else 1) + w.numel())
            kt = {"ms": e0.elapsed_time(e1), "launches": n, "flops": 2.0 * out.numel() * (cin // groups) * k * k * n,
                  "bytes": nbytes * n}
        us = kt["ms"] / kt["launches"] * 1e3
        tot += us
        # steady state inside a stream: 20 launches back to back between ONE pair of events (what Graph::run sees:
        # programmatic dependent launch overlaps each prologue with the previous tail, no event between launches)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            ops.conv(x, w, b, **kw)
        e1.record()
        torch.cuda.synchronize()
        chain = e0.elapsed_time(e1) / 20 * 1e3
        tot_chain += chain
        print(f"{name:24s} {us:8.1f} us  {kt['flops'] / kt['launches'] / us / 1e6:7.1f} TF/s  "
              f"{kt['bytes'] / kt['launches'] / us / 1e3:7.1f} GB/s  chained {chain:6.1f} us  {ctx.last_kernel}")
    print(f"sum {tot:.1f} us, chained {tot_chain:.1f} us")


if __name__ == "__main__":
    main()
