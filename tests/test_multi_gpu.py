"""Two-GPU tests (skipped on a one-GPU box): N-split GEMM with the NCCL all-gather, and the batch-sharded graph run."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_WORKER = r"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "oracle"))
rank, world = int(sys.argv[3]), int(sys.argv[4])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", init_method="tcp://127.0.0.1:" + sys.argv[2], rank=rank, world_size=world,
                        device_id=torch.device("cuda", rank))
import rten_fork_b200 as R
from rten_fork_b200 import synth
from rten_fork_b200.parallel import gemm_nsplit, gemm_nsplit_fused, SymmetricMatrix, run_batch_sharded, gather_outputs
import oracle as o, graph_oracle
ctx = R.default_context(rank)
ctx.set_math(R.MATH_3XTF32)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)   # kernels and NCCL are ordered on torch's stream
rng = o.XorShiftRng(1234)
for (m, k, n) in ((200, 300, 1000), (64, 128, 512)):
    a, b = rng.fill((m, k)), rng.fill((k, n))
    out = gemm_nsplit(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()).cpu().numpy()
    want = o.gemm(a, b)
    assert np.all(np.abs(out - want) <= 1e-8 + 1e-5 * np.abs(want)), (m, k, n)
# the same product with the gather fused into the GEMM epilogue (TMA stores into every rank's buffer over NVLink)
for (m, k, n) in ((200, 300, 1000), (64, 128, 512), (384, 96, 260)):
    a, b = rng.fill((m, k)), rng.fill((k, n))
    sym = SymmetricMatrix(m, n, ctx=ctx)
    sym.tensor.fill_(float("nan"))
    dist.barrier()
    out = gemm_nsplit_fused(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), sym)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    want = o.gemm(a, b)
    assert np.all(np.abs(got - want) <= 1e-8 + 1e-5 * np.abs(want)), ("fused", m, k, n)
    sym.close()
bld = synth.resnet50(image=32, width=8, blocks=(1, 1, 1, 1), classes=10, seed=0)
x = rng.fill((5, 3, 32, 32))
m = R.Model.load(bld.build(), ctx=ctx)
outs, (lo, hi) = run_batch_sharded(m, [x], world, rank)
full = gather_outputs(torch.from_numpy(outs[0]).cuda(), 5).cpu().numpy()
want = graph_oracle.run_graph(bld, [x])[0]
scale = float(np.abs(want).max())
assert np.all(np.abs(full - want) <= 2e-5 * scale + 1e-5 * np.abs(want))
dist.destroy_process_group()
print("ok", rank)
"""


def _launch(tmp_path, world):
    import time

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = str(s.getsockname()[1])
    s.close()
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    logs = [open(tmp_path / f"rank{r}.log", "w+") for r in range(world)]
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r), str(world)], stdout=logs[r],
                              stderr=subprocess.STDOUT) for r in range(world)]
    deadline = time.time() + 240
    # a rank that dies leaves its peers blocked inside the collective: stop everything as soon as one fails
    while time.time() < deadline and any(p.poll() is None for p in procs):
        if any(p.poll() not in (None, 0) for p in procs):
            break
        time.sleep(0.5)
    for p in procs:
        if p.poll() is None:
            p.kill()
    outs = []
    for f in logs:
        f.seek(0)
        outs.append(f.read())
        f.close()
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"ok {r}" in out, "\n".join(f"--- rank {i}:\n{o[-2500:]}" for i, o in enumerate(outs))


def test_nsplit_and_batch_shard_world_1(tmp_path):
    """The same worker with a one-rank NCCL group: every code path except the cross-GPU transfer."""
    _launch(tmp_path, 1)


def test_two_gpu_nsplit_and_batch_shard(tmp_path):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _launch(tmp_path, 2)
