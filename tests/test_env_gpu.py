"""Environment switches of the runtime (reference: RTEN_TIMING src/model.rs:124-147,587-590; RTEN_USE_POOL
src/graph.rs:871; RTEN_NUM_THREADS src/threading.rs:34-60 -> RTEN_GPU_SMS here; RTEN_GPU_MATH is new).  They are read
once per process / context, so every case runs in a fresh interpreter."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import json, sys
import numpy as np
sys.path.insert(0, %r)
import rten_fork_b200 as R
from rten_fork_b200.rten_builder import ModelBuilder
ctx = R.Context(0)
b = ModelBuilder()
x = b.add_value("x", ["n", 8, 6, 6])
b.inputs = [x]
rng = np.random.default_rng(0)
w = b.add_constant("w", rng.random((8, 8, 3, 3), dtype=np.float32) - 0.5)
y = b.op("Conv", [x, w], ModelBuilder.conv_attrs((1, 1, 1, 1)))
y = b.op("Relu", [y])
b.outputs = [b.op("Add", [y, x])]
m = R.Model.load(b.build(), ctx=ctx)
xin = rng.random((2, 8, 6, 6), dtype=np.float32)
outs = [m.run([xin])[0] for _ in range(3)]
print("RESULT " + json.dumps({"math": ctx.math, "stats": ctx.pool_stats(), "replays": m.graph_replays,
                              "sum": float(outs[-1].sum()), "same": bool((outs[0] == outs[-1]).all())}))
""" % ROOT


def _run(env):
    e = dict(os.environ)
    e.update(env)
    p = subprocess.run([sys.executable, "-c", SCRIPT], env=e, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[7:]), p.stdout, p.stderr


def test_defaults_and_math_env(rtg):
    base, out, _ = _run({})
    assert base["math"] == rtg.MATH_3XTF32 and base["same"] and "Graph run of" not in out
    assert base["replays"] >= 1                       # third run replays the captured graph
    tf32, _, _ = _run({"RTEN_GPU_MATH": "tf32"})
    assert tf32["math"] == rtg.MATH_TF32
    assert abs(tf32["sum"] - base["sum"]) <= 2e-3 * abs(base["sum"]) + 1e-3
    e = dict(os.environ, RTEN_GPU_MATH="fp64")
    p = subprocess.run([sys.executable, "-c", SCRIPT], env=e, capture_output=True, text=True, timeout=300)
    assert p.returncode != 0 and "RTEN_GPU_MATH must be" in p.stderr


def test_timing_env_prints_the_table():
    res, out, _ = _run({"RTEN_TIMING": "sort=name"})
    assert out.count("Graph run of") == 3 and res["replays"] == 0     # timed runs are eager
    rows = [l.split()[0] for l in out.split("Graph run of")[1].splitlines()[2:] if "ms (" in l]
    assert rows == sorted(rows) and len(rows) >= 1
    _, _, err = _run({"RTEN_TIMING": "colour=red sort=sideways"})
    assert 'Unrecognized timing option "colour"' in err and 'Unrecognized sort order "sideways"' in err


def test_use_pool_off_and_sm_limit():
    base, _, _ = _run({})
    off, _, _ = _run({"RTEN_USE_POOL": "0"})
    assert off["stats"]["hit_count"] == 0 and off["stats"]["len"] == 0 and off["replays"] == 0
    assert off["sum"] == base["sum"]
    few, _, _ = _run({"RTEN_GPU_SMS": "8"})
    assert few["sum"] == base["sum"] and few["same"]
    e = dict(os.environ, RTEN_GPU_SMS="none")
    p = subprocess.run([sys.executable, "-c", SCRIPT], env=e, capture_output=True, text=True, timeout=300)
    assert p.returncode != 0 and "RTEN_GPU_SMS must be" in p.stderr
