"""The HBM-bound kernels of bench.py's `hbm_kernels` key by themselves (for ncu captures and quick checks).
Usage: python scripts/bench_hbm.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import rten_fork_b200 as R
from rten_fork_b200 import ops

ctx = R.default_context(0)
stream = torch.cuda.Stream()
ctx.set_stream(stream.cuda_stream)
ctx.set_math(R.MATH_TF32_BF16X2)
print(json.dumps(bench.hbm_section(torch, R, ops, ctx, stream), indent=1))
