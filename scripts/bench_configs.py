"""Timings of the other BASELINE.json configs on one GPU (device-resident inputs, CUDA events, L2 flushed per step):
C1 GEMM 1024^3 through the GemmExecutor path, C3 BERT-base encoder b64 s128, C4 MobileNetV2 b64, plus the per-operator
timing report (the RTEN_TIMING equivalent).  Usage: python scripts/bench_configs.py [c1] [c3] [c4] [--math tf32|3xtf32]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import ops, synth


def timed(fn, steps=10, warmup=3, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(steps):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / steps


def main():
    argv = sys.argv[1:]
    math = "tf32_bf16x2"
    if "--math" in argv:
        i = argv.index("--math")
        math = argv[i + 1]
        del argv[i:i + 2]
    args = [a for a in argv if not a.startswith("--")]
    which = args or ["c1", "c3", "c4"]
    ctx = R.default_context(0)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ctx.set_math({"3xtf32": R.MATH_3XTF32, "tf32": R.MATH_TF32, "tf32_bf16x2": R.MATH_TF32_BF16X2}[math])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    if "c1" in which:
        for n in (512, 1024, 2048, 4096, 8192):
            a, b = torch.rand((n, n), device="cuda"), torch.rand((n, n), device="cuda")
            out = torch.empty((n, n), device="cuda")
            ms = timed(lambda: ops.matmul(a, b, out=out, ctx=ctx), steps=10, flush=flush if n <= 2048 else None)
            # the reference's bench_gemm rule (src/gemm.rs:1780-1903): iters = min(1000, 512^3 * 1000 / (m n k))
            # back-to-back calls, GFLOPS = 2 m n k iters / t  (operands stay L2-resident below ~2048, as they stay
            # cache-resident in the reference's loop)
            iters = max(1, min(1000, 512 ** 3 * 1000 // n ** 3))
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                ops.matmul(a, b, out=out, ctx=ctx)
            e1.record()
            torch.cuda.synchronize()
            loop_ms = e0.elapsed_time(e1) / iters
            print(f"C1/C5 GEMM {n}^3 {math}: {ms * 1e3:9.1f} us  {2 * n ** 3 / ms / 1e9:8.1f} TFLOP/s (single call, L2 "
                  f"flushed) | bench_gemm rule x{iters}: {loop_ms * 1e3:9.1f} us  {2 * n ** 3 / loop_ms / 1e9:8.1f} TFLOP/s  "
                  f"[{ctx.last_kernel}]")
    for tag, build, shape_in in (("c3", lambda: synth.bert_encoder(batch=64, seq=128), None),
                                 ("c4", lambda: synth.mobilenet_v2(), (64, 3, 224, 224))):
        if tag not in which:
            continue
        t0 = time.time()
        b = build()
        model = R.Model.load(b.build(), ctx=ctx)
        print(f"{tag}: built + loaded in {time.time() - t0:.1f} s, {model.num_ops} operators")
        if tag == "c3":
            ins = [torch.zeros((64, 128), dtype=torch.int32, device="cuda"),
                   torch.ones((64, 128), dtype=torch.int32, device="cuda")]
            flops = 1.430e12
        else:
            ins = [torch.rand(shape_in, device="cuda")]
            flops = 2 * 0.3e9 * 64
        outs = None

        def step():
            nonlocal outs
            outs = model.run(ins, out=outs)

        # first run allocates outputs as numpy; keep device outputs instead
        r = model.run(ins)
        outs = [torch.empty(tuple(o.shape), dtype=torch.float32, device="cuda") for o in r]
        l0 = ctx.launch_count
        ms = timed(step, steps=5, warmup=2, flush=flush)
        launches = (ctx.launch_count - l0) / 7
        unit = "sequences" if tag == "c3" else "images"
        print(f"{tag} {math}: {ms:8.3f} ms/step  {64 / ms * 1e3:9.1f} {unit}/s  {flops / ms / 1e9:7.1f} TFLOP/s (algorithmic)  "
              f"{launches:.0f} launches/step")
        model.set_timing(True)
        model.run(ins, out=outs)
        print(model.timing_report())
        model.set_timing(False)


if __name__ == "__main__":
    main()
