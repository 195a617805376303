"""Developer experiment: the depthwise 3x3 cases of bench.py's `hbm_kernels` under RTG_DW_ROWS / RTG_DW_PRED overrides
(the switches are read once per process, so every setting runs in its own process).
Usage: python scripts/exp_depthwise.py            # sweep
       python scripts/exp_depthwise.py one        # one process: prints the two cases with the current environment"""
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one():
    import torch
    import rten_fork_b200 as R
    from rten_fork_b200 import ops
    ctx = R.default_context(0)
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev).manual_seed(99)
    res = {}
    with torch.cuda.stream(stream):
        for name, (n, c, hw, s) in {"s1_32ch_112": (64, 32, 112, 1), "s2_96ch_112": (64, 96, 112, 2),
                                    "s1_144ch_56": (64, 144, 56, 1), "s1_384ch_14": (64, 384, 14, 1),
                                    "s1_960ch_7": (64, 960, 7, 1)}.items():
            x = torch.rand((n, c, hw, hw), device=dev, generator=gen).contiguous(memory_format=torch.channels_last)
            w = torch.rand((c, 1, 3, 3), device=dev, generator=gen) - 0.5
            b = torch.rand((c,), device=dev, generator=gen)
            ohw = (hw + 2 - 3) // s + 1
            y = torch.empty((n, c, ohw, ohw), device=dev).contiguous(memory_format=torch.channels_last)
            fn = lambda: ops.conv(x, w, b, padding=(1, 1, 1, 1), groups=c, strides=(s, s), activation="clip",
                                  clip=(0.0, 6.0), out=y, ctx=ctx)
            for _ in range(3):
                fn()
            stream.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(20):
                fn()
            e1.record(stream)
            stream.synchronize()
            us = e0.elapsed_time(e1) / 20 * 1e3
            nbytes = 4.0 * (x.numel() + y.numel() + w.numel())
            res[name] = (round(us, 1), round(nbytes / us / 1e3 / 6550.1, 3))
        x = torch.rand((32, 64, 112, 112), device=dev, generator=gen).contiguous(memory_format=torch.channels_last)
        y = torch.empty((32, 64, 56, 56), device=dev).contiguous(memory_format=torch.channels_last)
        fn = lambda: ops.max_pool(x, (3, 3), (2, 2), (1, 1, 1, 1), ctx=ctx, out=y)
        for _ in range(3):
            fn()
        stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(20):
            fn()
        e1.record(stream)
        stream.synchronize()
        us = e0.elapsed_time(e1) / 20 * 1e3
        res["maxpool"] = (round(us, 1), round(4.0 * (x.numel() + y.numel()) / us / 1e3 / 6550.1, 3))
    print(json.dumps(res))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "one":
        one()
    else:
        settings = [{}] + [{"RTG_DW_PRED": "1", "RTG_POOL_ROWS2": "0"}] + [{"RTG_DW_ROWS": str(r)} for r in (4, 6, 7, 8, 10, 12, 13, 14, 16, 19, 23, 28, 38, 56)]
        for s in settings:
            env = dict(os.environ, **s)
            r = subprocess.run([sys.executable, __file__, "one"], env=env, capture_output=True, text=True, timeout=300)
            print(s, r.stdout.strip() or r.stderr[-300:], flush=True)
