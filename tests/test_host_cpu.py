"""CPU tests of the host side: the C-ABI library exports what include/rten_gpu.h declares, the package refuses to run
without its CUDA library / a device, the `.rten` writer produces the container layout the loader reads, and the
multi-GPU partition logic (world_size 2, gloo)."""
import ctypes
import os
import re
import struct
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "rten_fork_b200", "librten_gpu.so")


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "rten_gpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rtg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(LIB):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "rten_fork_b200", "csrc"), "-j8"])
    lib = ctypes.CDLL(LIB)  # loads without a GPU: no CUDA call happens at load time
    names = _declared_symbols()
    assert len(names) >= 40
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/rten_gpu.h but not exported: {missing}"


def test_status_codes_and_error_string_without_gpu():
    lib = ctypes.CDLL(LIB)
    lib.rtg_last_error.restype = ctypes.c_char_p
    # NULL arguments are rejected before any CUDA work
    assert lib.rtg_model_num_inputs(None, None) != 0
    assert b"NULL" in lib.rtg_last_error()


def test_no_cpu_fallback():
    """Creating a context without a CUDA device fails loudly (there is no CPU path behind the API)."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import rten_fork_b200 as R

    with pytest.raises(R.RtgError):
        R.Context(0)
    src = open(os.path.join(ROOT, "rten_fork_b200", "_native.py")).read()
    assert "oracle" not in src and "liboracle" not in src


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "rten_fork_b200")
    for dirpath, _dirs, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "graph_oracle" not in text and "liboracle" not in text, f


def test_rten_container_layout():
    """V2: 'RTEN' magic, version 2, model offset / length and tensor-data offset (src/header.rs:39-134); V1: a bare
    FlatBuffer whose file identifier is 'RTEN' (src/model.rs:280-300)."""
    from rten_fork_b200.rten_builder import ModelBuilder

    b = ModelBuilder()
    x = b.add_value("x", ["batch", 8])
    b.inputs = [x]
    w = b.add_constant("w", np.arange(8 * 200, dtype=np.float32).reshape(8, 200))
    b.outputs = [b.op("MatMul", [x, w])]
    v2 = b.build(version=2)
    assert v2[:4] == b"RTEN"
    version, model_off, model_len, data_off = struct.unpack_from("<IQQQ", v2, 4)  # 32-byte header
    data_len = len(v2) - data_off
    assert version == 2 and model_off == 32
    assert model_off + model_len <= data_off <= len(v2)
    assert v2[model_off + 4:model_off + 8] == b"RTEN"          # FlatBuffers file identifier
    assert data_len >= 8 * 200 * 4                              # the large constant lives in the tensor-data region
    raw = np.frombuffer(v2, dtype=np.float32, count=8 * 200, offset=data_off)
    assert (raw == np.arange(8 * 200, dtype=np.float32)).all()
    assert data_off % 16 == 0
    v1 = b.build(version=1)
    assert v1[4:8] == b"RTEN" and v1[:4] != b"RTEN"


def test_shard_ranges():
    from rten_fork_b200.parallel import column_blocks, shard_range

    for total in (0, 1, 7, 32, 33):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)
    for n in (1, 100, 128, 129, 1000, 16384):
        for world in (1, 2, 4, 8):
            blocks = column_blocks(n, world)
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
            assert all(lo % 128 == 0 for lo, hi in blocks if hi > lo)


_WORKER = r"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from rten_fork_b200.parallel import gemm_nsplit, gather_outputs, run_batch_sharded, shard_range
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:" + sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
g = torch.Generator().manual_seed(5)
for (m, k, n) in ((5, 16, 300), (4, 8, 100), (3, 8, 512)):   # ragged, one empty rank, equal blocks
    a, b = torch.rand((m, k), generator=g), torch.rand((k, n), generator=g)
    out = gemm_nsplit(a, b, local_gemm=lambda x, y: x @ y)
    assert out.shape == (m, n) and torch.allclose(out, a @ b, atol=1e-6), (m, k, n)
class FakeModel:                                               # Model.run stand-in: y = 2 x over the local rows
    def run(self, inputs):
        return [inputs[0] * 2.0]
x = np.arange(7 * 3, dtype=np.float32).reshape(7, 3)
outs, (lo, hi) = run_batch_sharded(FakeModel(), [x], 2, rank)
assert (lo, hi) == shard_range(7, 2, rank) and outs[0].shape == (hi - lo, 3)
full = gather_outputs(torch.from_numpy(outs[0]), 7)
assert torch.equal(full, torch.from_numpy(x * 2.0))
dist.destroy_process_group()
print("ok", rank)
"""


def test_nsplit_gemm_and_batch_sharding_two_ranks(tmp_path):
    """world_size 2 over gloo: column partition + all-gather of the N-split GEMM, batch-sharded run + gather."""
    import socket

    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = str(s.getsockname()[1])
    s.close()
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r} failed:\n{o}"
        assert f"ok {r}" in o


def test_bench_graph_flops():
    """bench.py's algorithmic flop count for ResNet-50 matches BASELINE.md's 8.178 GFLOP per image."""
    sys.path.insert(0, ROOT)
    import bench
    from rten_fork_b200 import synth

    flops = bench.graph_flops(synth.resnet50(), 32)
    assert abs(flops / 32 / 1e9 - 8.178) < 0.01


def test_cli_dim_specs_and_input_heuristics():
    """rten-cli's input generation rules (rten-cli/src/main.rs:60-94, 249-267), exercised without a device."""
    from rten_fork_b200 import cli

    dims = cli._parse_dims(["dim0=4,input_ids.dim1=7", "dim1=9"])
    assert dims == [(None, "dim0", 4), ("input_ids", "dim1", 7), (None, "dim1", 9)]
    assert cli._resolve("input_ids", ["dim0", "dim1"], dims, quiet=True) == [4, 7]
    assert cli._resolve("image", ["dim0", 3, "dim2"], dims, quiet=True) == [4, 3, 1]
    with pytest.raises(SystemExit):
        cli._parse_dims(["dim0=abc"])

    class FakeModel:
        inputs = [{"name": "input_ids", "shape": [None, 8]}, {"name": "attention_mask", "shape": [None, 8]},
                  {"name": "pixels", "shape": [2, 3]}]

    ids, mask, px = cli.generate_inputs(FakeModel(), dims, quiet=True, seed=0)
    assert ids.dtype == np.int32 and ids.shape == (4, 8) and not ids.any()
    assert mask.dtype == np.int32 and mask.shape == (4, 8) and (mask == 1).all()
    assert px.dtype == np.float32 and px.shape == (2, 3) and (px >= 0).all() and (px < 1).all()
