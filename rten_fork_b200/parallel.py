"""Multi-GPU partitioning of the hot path (one process per GPU, ``torch.distributed``).

Two partitionings, both from BASELINE.json's north_star:

* **Batch-sharded Graph::run** — every rank owns a contiguous slice of the batch and runs the whole graph on it.  The
  data path has no exchange step, so there is no collective; ``gather_outputs`` exists only for callers that want the
  full result on every rank (it is not part of the timed path in ``bench.py``).
* **N-split GEMM** — a single large ``A[M,K] @ B[K,N]`` is split by output columns: rank r multiplies A by its column
  block of B and the blocks are gathered with one all-gather (NCCL over NVLink on GPUs; gloo in the CPU tests).

The reference is single-process (src/threading.rs: one rayon pool); these helpers are the multi-GPU counterpart of its
per-thread column partition ``nc = clamp(N / threads, 128..1024)`` (src/gemm.rs:546-571): one column block per GPU.
"""
from __future__ import annotations

import numpy as np


def shard_range(total: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous [start, stop) of ``total`` items for ``rank``; the first ``total % world`` ranks get one extra."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} for world size {world}")
    base, extra = divmod(total, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def column_blocks(n: int, world: int, align: int = 128) -> list[tuple[int, int]]:
    """Column ranges of an N-split GEMM: blocks are multiples of ``align`` (one tcgen05 tile column, so that no rank
    runs a partial tile except the last) and differ by at most one tile; trailing ranks may be empty when N is small."""
    tiles = -(-n // align)
    out = []
    for r in range(world):
        t0, t1 = shard_range(tiles, world, r)
        out.append((min(t0 * align, n), min(t1 * align, n)))
    return out


def run_batch_sharded(model, inputs, world: int, rank: int, batch_axis: int = 0):
    """Model::run on this rank's slice of the batch (no communication).  Returns (outputs, (start, stop))."""
    n = inputs[0].shape[batch_axis]
    lo, hi = shard_range(n, world, rank)
    local = [x[lo:hi] if batch_axis == 0 else np.take(x, range(lo, hi), axis=batch_axis) for x in inputs]
    if hi == lo:
        return [], (lo, hi)
    return model.run(local), (lo, hi)


def gather_outputs(local, total_rows: int, group=None):
    """All-gather of ragged per-rank row blocks (torch tensors on the backend's device).  Not on the timed path."""
    import torch
    import torch.distributed as dist

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    rows = [shard_range(total_rows, world, r) for r in range(world)]
    width = max(hi - lo for lo, hi in rows)
    pad = torch.zeros((width,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, rows)], dim=0)


def gemm_nsplit(a, b, group=None, local_gemm=None, align: int = 128, ctx=None):
    """``A @ B`` with B's columns split over the ranks of ``group`` and the result gathered on every rank.

    ``a``: [M, K], ``b``: [K, N] torch tensors on this rank's device.  ``local_gemm(a, b_block) -> [M, n_r]`` defaults
    to the tcgen05 GEMM of this package on context ``ctx`` (CUDA tensors); the CPU tests pass a stand-in so that only the partition /
    gather logic is exercised.  The gather is ONE all-gather of equal-sized padded blocks (NCCL needs equal sizes);
    the padding columns are dropped when the blocks are laid side by side.
    """
    import torch
    import torch.distributed as dist

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    m, n = a.shape[0], b.shape[1]
    blocks = column_blocks(n, world, align)
    lo, hi = blocks[rank]
    if local_gemm is None:
        from . import ops
        from .context import default_context

        # the contraction must be ordered with the torch work around it (the copy into `mine`, the all-gather): run it
        # on torch's current stream, not on the context's own
        ctx = ctx or default_context(a.device.index)
        ctx.set_stream(torch.cuda.current_stream(a.device).cuda_stream)

        def local_gemm(x, y):
            out = torch.empty((x.shape[0], y.shape[1]), dtype=torch.float32, device=x.device)
            return ops.matmul(x, y, out=out, ctx=ctx)

    width = max(h - l for l, h in blocks)
    mine = torch.zeros((m, width), dtype=torch.float32, device=a.device)
    if hi > lo:
        mine[:, : hi - lo] = local_gemm(a, b[:, lo:hi])
    flat = torch.empty((world * m, width), dtype=torch.float32, device=a.device)
    dist.all_gather_into_tensor(flat, mine, group=group)  # rank r's block lands in rows [r*m, (r+1)*m)
    gathered = flat.view(world, m, width)
    if all(h - l == width for l, h in blocks):
        return gathered.permute(1, 0, 2).reshape(m, world * width)  # one strided copy
    return torch.cat([gathered[r, :, : h - l] for r, (l, h) in enumerate(blocks)], dim=1)


class _DevicePtrView:
    """Minimal ``__cuda_array_interface__`` carrier so that torch can wrap a raw device pointer without a copy."""

    def __init__(self, ptr: int, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class SymmetricMatrix:
    """A row-major ``[m, n]`` f32 result buffer on every rank of ``group``, each mapped into every other rank's address
    space (``rtg_symm_alloc`` / ``rtg_symm_open``: CUDA IPC handles exchanged with ``all_gather_object``).  ``ptrs[r]``
    is rank r's buffer as seen from this process; ``tensor`` is the local buffer as a torch tensor."""

    def __init__(self, m: int, n: int, ctx=None, group=None):
        import ctypes as C

        import torch
        import torch.distributed as dist

        from . import _native as N
        from .context import default_context

        self.ctx = ctx or default_context(torch.cuda.current_device())
        self.group, self.m, self.n = group, m, n
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if n % 4:
            raise N.RtgError(6, "n must be a multiple of 4")
        ptr = C.c_void_p()
        handle = (C.c_uint8 * 64)()
        N.check(N.lib().rtg_symm_alloc(self.ctx.handle, C.c_size_t(m * n * 4), C.byref(ptr), handle))
        self._local = ptr.value
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=group)
        self.ptrs, self._opened = [], []
        for r, h in enumerate(handles):
            if r == self.rank:
                self.ptrs.append(self._local)
                continue
            peer = C.c_void_p()
            buf = (C.c_uint8 * 64).from_buffer_copy(h)
            N.check(N.lib().rtg_symm_open(self.ctx.handle, buf, C.byref(peer)))
            self.ptrs.append(peer.value)
            self._opened.append(peer.value)
        self.tensor = torch.as_tensor(_DevicePtrView(self._local, (m, n)), device=torch.device("cuda", self.ctx.device))
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.tensor.device)
        self._written = False   # a fused product has been stored into the buffers since the last leading fence

    def bind_stream(self):
        """The context launches on ITS stream; the fences (NCCL all-reduces) and the caller's reads of ``tensor`` run on
        torch's current stream.  Binding the context to that stream orders all three."""
        import torch

        self.ctx.set_stream(torch.cuda.current_stream(self.tensor.device).cuda_stream)

    def fence(self):
        """Stream-ordered cross-rank barrier: once it has run on this rank's stream, every rank's GEMM (and with it
        every peer store into this rank's buffer) has completed.  One 4-byte all-reduce; no host synchronisation.
        Only orders work on torch's current stream: ``gemm_nsplit_fused`` binds the context to it first."""
        import torch.distributed as dist

        dist.all_reduce(self._flag, group=self.group)

    def close(self):
        import ctypes as C

        import torch.distributed as dist

        from . import _native as N

        dist.barrier(self.group)  # no rank unmaps or frees while a peer may still be writing
        for p in self._opened:
            N.check(N.lib().rtg_symm_close(self.ctx.handle, C.c_void_p(p)))
        self._opened = []
        dist.barrier(self.group)
        if self._local:
            N.check(N.lib().rtg_symm_free(self.ctx.handle, C.c_void_p(self._local)))
            self._local = None
        self.tensor = None


def gemm_nsplit_fused(a, b, out: SymmetricMatrix, align: int = 128, fence: bool = True):
    """``A @ B`` with B's columns split over the ranks, the gather fused into the GEMM epilogue: rank r multiplies A by
    its column block and every finished tile is stored by TMA into columns ``[lo, hi)`` of ALL ranks' ``out`` buffers
    (its own through HBM, the peers' over NVLink), so no all-gather runs afterwards — only ``out.fence()``, a 4-byte
    all-reduce that orders the ranks (and a second one in front when the same buffers are stored into again).  The
    context is bound to torch's current stream, so the product, the fences and the caller's reads are stream-ordered.  ``a``: [M, K], ``b``: [K, N] CUDA tensors (any strides).  Returns the local
    ``[M, N]`` tensor of ``out`` (complete once the fence has run on the stream).

    ``gemm_nsplit`` (local GEMM, then one NCCL all-gather and a strided copy) is the baseline this replaces."""
    import ctypes as C

    from . import _native as N

    m, k = a.shape
    n = b.shape[1]
    if (m, n) != (out.m, out.n):
        raise N.RtgError(3, "output buffer shape does not match the product")
    out.bind_stream()
    if out._written:
        # Re-use of the buffers: a fast rank's peer stores of this product must not land in a buffer a slower rank is
        # still reading from the previous one.  A leading fence makes every rank's stream reach this point (all of its
        # earlier reads are ordered before it) before any rank stores again.
        out.fence()
    out._written = True
    lo, hi = column_blocks(n, out.world, align)[out.rank]
    if hi > lo:
        bb = b[:, lo:hi]
        order = [out.rank] + [r for r in range(out.world) if r != out.rank]
        dst = (C.c_void_p * len(order))(*[out.ptrs[r] for r in order])
        N.check(N.lib().rtg_gemm_scatter(
            out.ctx.handle, C.c_void_p(a.data_ptr()), C.c_size_t(m), C.c_size_t(k),
            C.c_ssize_t(a.stride(0)), C.c_ssize_t(a.stride(1)), C.c_void_p(bb.data_ptr()), C.c_size_t(hi - lo),
            C.c_ssize_t(bb.stride(0)), C.c_ssize_t(bb.stride(1)), C.c_size_t(lo), C.c_size_t(n), dst,
            C.c_int32(len(order))))
    if fence:
        out.fence()
    return out.tensor
