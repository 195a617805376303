"""Experiment: one batch of 32 split over S concurrent streams (S contexts, each running ResNet-50 at batch 32/S from its
own host thread) against one stream at batch 32.  Usage: python scripts/exp_streams.py [math]"""
import os
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import synth

math = sys.argv[1] if len(sys.argv) > 1 else "tf32_bf16x2"
modes = {"tf32_bf16x2": R.MATH_TF32_BF16X2, "3xtf32": R.MATH_3XTF32, "tf32": R.MATH_TF32}
blob = synth.resnet50(image=224, classes=1000, seed=0).build()
STEPS = 20
for S in (1, 2, 4):
    per = 32 // S
    ctxs, models, xs, outs, streams = [], [], [], [], []
    for s in range(S):
        c = R.Context(0)
        st = torch.cuda.Stream()
        c.set_stream(st.cuda_stream)
        c.set_math(modes[math])
        ctxs.append(c), streams.append(st)
        models.append(R.Model.load(blob, ctx=c))
        xs.append(torch.rand((per, 3, 224, 224), device="cuda"))
        outs.append(torch.empty((per, 1000), device="cuda"))

    def work(s, n):
        for _ in range(n):
            models[s].run([xs[s]], out=[outs[s]])

    for s in range(S):
        work(s, 3)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        th = [threading.Thread(target=work, args=(s, STEPS)) for s in range(S)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    print(f"S={S} (batch {per} per stream): {best / STEPS * 1e3:.3f} ms per 32 images, {32 * STEPS / best:.0f} img/s", flush=True)
    for m in models:
        m.close()
    for c in ctxs:
        c.close()
