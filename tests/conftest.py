import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with `pytest -m gpu` under gpurun)")


def _has_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    import oracle as o

    o.build()
    return o


@pytest.fixture(scope="session")
def rtg():
    import rten_fork_b200 as R

    return R


@pytest.fixture(scope="session")
def ctx(rtg):
    return rtg.default_context(0)


@pytest.fixture(params=["3xtf32", "tf32", "tf32_bf16x2"])
def math_mode(request, rtg, ctx):
    """The contraction modes with the tolerance BASELINE.json's north_star states for each: the two fp32-accurate
    splits (3xTF32, and TF32 + one bf16 correction MMA) under the reference's 1e-5 rule, plain TF32 under 2e-3."""
    mode = {"3xtf32": rtg.MATH_3XTF32, "tf32": rtg.MATH_TF32, "tf32_bf16x2": rtg.MATH_TF32_BF16X2}[request.param]
    ctx.set_math(mode)
    yield {"name": request.param, "rtol": 2e-3 if request.param == "tf32" else 1e-5}
    ctx.set_math(rtg.MATH_3XTF32)


def expect_equal(x, y, rtol=1e-5, atol=1e-8):
    """rten-tensor/src/test_util.rs:44-61: |x - y| <= atol + rtol * |y| for every element."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    assert x.shape == y.shape, f"shape {x.shape} != {y.shape}"
    bad = np.abs(x - y) > atol + rtol * np.abs(y)
    if bad.any():
        i = np.argwhere(bad)[0]
        raise AssertionError(
            f"{bad.sum()} of {bad.size} elements differ; first at {tuple(i)}: {x[tuple(i)]} vs {y[tuple(i)]} "
            f"(max rel {np.max(np.abs(x - y) / (np.abs(y) + 1e-30)):.3e})")
