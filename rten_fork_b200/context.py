"""Per-GPU context: stream, device TensorPool and weight cache (``rtg_ctx`` in include/rten_gpu.h)."""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import _native as N


FORCE_IM2COL_FALLBACK = 1 << 16


class Context:
    """Owns one ``rtg_ctx``.  Mirrors what a run of the reference owns implicitly: a TensorPool
    (src/tensor_pool.rs:103-199) plus, new here, the CUDA stream and device weight cache."""

    def __init__(self, device: int = 0, math: int | None = None):
        self._h = C.c_void_p()
        N.check(N.lib().rtg_ctx_create(C.c_int32(device), C.byref(self._h)))
        self.device = device
        if math is not None:  # None: the library's default (3xTF32, or what RTEN_GPU_MATH names)
            self.set_math(math)

    # -- lifecycle
    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            N.lib().rtg_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    # -- configuration
    def set_math(self, mode: int):
        N.check(N.lib().rtg_ctx_set_math(self._h, C.c_int32(mode)))

    @property
    def math(self) -> int:
        m = C.c_int32()
        N.check(N.lib().rtg_ctx_get_math(self._h, C.byref(m)))
        return m.value

    def set_stream(self, cuda_stream: int):
        N.check(N.lib().rtg_ctx_set_stream(self._h, C.c_void_p(cuda_stream)))

    def set_debug_flags(self, flags: int):
        """Developer / test switches (``rtg_ctx_set_debug_flags``), e.g. ``FORCE_IM2COL_FALLBACK``."""
        N.check(N.lib().rtg_ctx_set_debug_flags(self._h, C.c_int32(flags)))

    def sync(self):
        N.check(N.lib().rtg_ctx_sync(self._h))

    @property
    def launch_count(self) -> int:
        c = C.c_uint64()
        N.check(N.lib().rtg_ctx_launch_count(self._h, C.byref(c)))
        return c.value

    @property
    def last_kernel(self) -> str:
        return N.lib().rtg_ctx_last_kernel(self._h).decode()

    def set_kernel_timing(self, enabled: bool):
        N.check(N.lib().rtg_ctx_set_kernel_timing(self._h, C.c_int32(bool(enabled))))

    def kernel_timing(self):
        """Launch count, summed device ms, algorithmic flops and bytes of the contraction-kernel launches recorded
        since the last call (synchronises the stream)."""
        n, ms, fl, by = C.c_uint64(), C.c_double(), C.c_double(), C.c_double()
        N.check(N.lib().rtg_ctx_kernel_timing(self._h, C.byref(n), C.byref(ms), C.byref(fl), C.byref(by)))
        return {"launches": n.value, "ms": ms.value, "flops": fl.value, "bytes": by.value}

    # -- TensorPool surface (alloc / add / stats)
    def pool_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        N.check(N.lib().rtg_pool_alloc(self._h, C.c_size_t(nbytes), C.byref(p)))
        return p.value

    def pool_free(self, ptr: int):
        N.check(N.lib().rtg_pool_free(self._h, C.c_void_p(ptr)))

    def pool_stats(self):
        a, h, l = C.c_uint64(), C.c_uint64(), C.c_uint64()
        N.check(N.lib().rtg_pool_stats(self._h, C.byref(a), C.byref(h), C.byref(l)))
        return {"alloc_count": a.value, "hit_count": h.value, "len": l.value}

    def upload(self, dst_ptr: int, src: np.ndarray):
        src = np.ascontiguousarray(src)
        N.check(N.lib().rtg_upload(self._h, C.c_void_p(dst_ptr), src.ctypes.data_as(C.c_void_p), C.c_size_t(src.nbytes)))

    def download(self, dst: np.ndarray, src_ptr: int):
        assert dst.flags.c_contiguous
        N.check(N.lib().rtg_download(self._h, dst.ctypes.data_as(C.c_void_p), C.c_void_p(src_ptr), C.c_size_t(dst.nbytes)))


_default = {}
_default_lock = threading.Lock()


def default_context(device: int = 0) -> Context:
    """Process-wide context per device (created on first use)."""
    with _default_lock:
        ctx = _default.get(device)
        if ctx is None:
            ctx = _default[device] = Context(device)
        return ctx


def context_for(x) -> Context:
    """Default context of the device ``x`` lives on (torch CUDA tensor), else of the current / first device."""
    if not isinstance(x, np.ndarray) and getattr(x, "is_cuda", False):
        return default_context(x.device.index)
    try:
        import sys

        torch = sys.modules.get("torch")
        if torch is not None and torch.cuda.is_available():
            return default_context(torch.cuda.current_device())
    except Exception:
        pass
    return default_context(0)


def as_tensor(x, device=None) -> N.Tensor:
    """Describe a numpy array (host) or a torch CUDA tensor (device) as an ``rtg_tensor`` view.  The caller keeps the
    object alive for the duration of the call."""
    t = N.Tensor()
    if isinstance(x, np.ndarray):
        if x.dtype == np.float32:
            t.dtype = N.RTG_F32
        elif x.dtype == np.int32:
            t.dtype = N.RTG_I32
        else:
            raise N.RtgError(1, f"unsupported dtype {x.dtype}")
        t.data = x.ctypes.data
        t.ndim = x.ndim
        for i in range(x.ndim):
            t.shape[i] = x.shape[i]
            if x.strides[i] % x.itemsize or x.strides[i] < 0:
                raise N.RtgError(6, "strides must be non-negative multiples of the element size")
            t.strides[i] = x.strides[i] // x.itemsize
        t.device = N.RTG_HOST
        return t
    # torch tensor (duck-typed so that torch is not imported unless the caller already uses it)
    import torch

    if not isinstance(x, torch.Tensor):
        raise TypeError(f"expected numpy.ndarray or torch.Tensor, got {type(x)}")
    if x.dtype == torch.float32:
        t.dtype = N.RTG_F32
    elif x.dtype == torch.int32:
        t.dtype = N.RTG_I32
    else:
        raise N.RtgError(1, f"unsupported dtype {x.dtype}")
    t.data = x.data_ptr()
    t.ndim = x.dim()
    for i in range(x.dim()):
        t.shape[i] = x.shape[i]
        t.strides[i] = x.stride(i)
    t.device = x.device.index if x.is_cuda else N.RTG_HOST
    return t
