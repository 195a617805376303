"""C5: square f32 GEMM n x n x n, N-split over the ranks of a torchrun job (BASELINE.json configs[4]).

Three timings per size, CUDA events on the launching stream, max over ranks:
  local   the rank's column block only (A . B[:, block]) -- what perfect overlap would cost
  nccl    parallel.gemm_nsplit: local GEMM, one NCCL all-gather, strided copy into [n, n]   (the baseline)
  fused   parallel.gemm_nsplit_fused: every finished tile is TMA-stored into all ranks' buffers from the GEMM epilogue,
          then a 4-byte all-reduce orders the ranks
Usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P \
       scripts/bench_nsplit.py [--math tf32|3xtf32] [sizes...]"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rten_fork_b200 as R
from rten_fork_b200 import ops
from rten_fork_b200.parallel import SymmetricMatrix, column_blocks, gemm_nsplit, gemm_nsplit_fused


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local_rank)
    if world > 1 or "RANK" in os.environ:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    else:
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29655", rank=0, world_size=1,
                                device_id=torch.device("cuda", 0))
    argv = sys.argv[1:]
    math = "3xtf32"
    if "--math" in argv:
        i = argv.index("--math")
        math = argv[i + 1]
        del argv[i:i + 2]
    sizes = [int(x) for x in argv] or [2048, 4096, 8192, 16384]
    ctx = R.default_context(local_rank)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ctx.set_math({"3xtf32": R.MATH_3XTF32, "tf32": R.MATH_TF32, "tf32_bf16x2": R.MATH_TF32_BF16X2}[math])

    def timed(fn, iters):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for n in sizes:
        g = torch.Generator(device="cuda").manual_seed(1234)
        a = torch.rand((n, n), device="cuda", generator=g)
        b = torch.rand((n, n), device="cuda", generator=g)
        lo, hi = column_blocks(n, world)[rank]
        iters = 20 if n <= 4096 else 5
        blk = torch.empty((n, max(hi - lo, 1)), device="cuda")
        t_local = timed(lambda: ops.matmul(a, b[:, lo:hi], out=blk[:, : hi - lo], ctx=ctx) if hi > lo else None, iters)
        t_nccl = timed(lambda: gemm_nsplit(a, b), iters)
        sym = SymmetricMatrix(n, n, ctx=ctx)
        t_fused = timed(lambda: gemm_nsplit_fused(a, b, sym), iters)
        ref = gemm_nsplit(a, b)
        torch.cuda.synchronize()
        same = bool(torch.equal(ref, sym.tensor))
        sym.close()
        if rank == 0:
            fl = 2.0 * n ** 3
            print(json.dumps({"n": n, "gpus": world, "math": math, "local_ms": round(t_local, 4),
                              "nccl_ms": round(t_nccl, 4), "fused_ms": round(t_fused, 4),
                              "nccl_tflops": round(fl / t_nccl / 1e9, 1), "fused_tflops": round(fl / t_fused / 1e9, 1),
                              "gather_bytes_per_gpu": (world - 1) * 4 * n * n // world,
                              "fused_equals_nccl_bitwise": same}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
