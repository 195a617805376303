"""Timing of rtg_sharded_model_run over different device lists (developer aid)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import synth
blob = synth.resnet50(image=224, classes=1000, seed=0).build()
n = torch.cuda.device_count()
xp = torch.rand((32, 3, 224, 224)).pin_memory()
op = torch.empty((32, 1000)).pin_memory()
for devs in [[0], [0, 0]] + ([[0, 1]] if n > 1 else []) + ([[0, 1, 2, 3]] if n > 3 else []) + ([list(range(8))] if n > 7 else []):
    sm = R.ShardedModel(blob, devs, math=R.MATH_TF32_BF16X2)
    for name, x, o in (("pinned", xp.numpy(), op.numpy()), ("pageable", xp.numpy().copy(), np.empty((32, 1000), np.float32))):
        for _ in range(4):
            sm.run([x], [o])
        t0 = time.perf_counter()
        for _ in range(20):
            sm.run([x], [o])
        dt = (time.perf_counter() - t0) / 20
        print(f"devices {devs} {name}: {dt * 1e3:.3f} ms per batch-32 call, {32 / dt:.0f} img/s", flush=True)
    sm.close()
