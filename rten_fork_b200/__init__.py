"""rten_fork_b200 — B200 (sm_100a) implementation of RTen's dense-compute hot path behind the reference's operator API.

The compute lives in ``librten_gpu.so`` (hand-written CUDA, C ABI in ``include/rten_gpu.h``); this package is the
host-side mirror of the reference interface for that path.  There is no CPU fallback.
"""
from ._native import MATH_3XTF32, MATH_TF32, MATH_TF32_BF16X2, RtgError  # noqa: F401
from .context import Context, default_context  # noqa: F401
from .gemm import GemmExecutor, gemm  # noqa: F401
from . import ops, synth  # noqa: F401
from .model import Model, ShardedModel  # noqa: F401
from .rten_builder import ModelBuilder  # noqa: F401
