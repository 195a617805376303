"""Host-side mirror of ``GemmExecutor`` (reference: src/gemm.rs:233-543) over the C-ABI ``rtg_gemm``."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N
from .context import Context, default_context

_f32p = C.POINTER(C.c_float)


def _mat(x):
    """(pointer, rows, cols, row_stride, col_stride, device, keepalive) of a 2-D f32 matrix view."""
    if isinstance(x, np.ndarray):
        if x.dtype != np.float32 or x.ndim != 2:
            raise N.RtgError(1, "expected a 2-D float32 matrix")
        return x.ctypes.data, x.shape[0], x.shape[1], x.strides[0] // 4, x.strides[1] // 4, N.RTG_HOST
    import torch

    assert isinstance(x, torch.Tensor) and x.dtype == torch.float32 and x.dim() == 2
    dev = x.device.index if x.is_cuda else N.RTG_HOST
    return x.data_ptr(), x.shape[0], x.shape[1], x.stride(0), x.stride(1), dev


class GemmExecutor:
    """``out = alpha * a @ b + beta * out (+ bias[row])`` on the GPU.

    Same call surface as the reference executor: ``gemm`` / ``gemm_bias`` take the output as a flat buffer plus a row
    stride, ``a`` and ``b`` as arbitrarily strided matrix views.  ``prepack_a`` / ``prepack_b`` exist for API parity;
    TMA reads operands in place, so they return their argument.
    """

    def __init__(self, ctx: Context | None = None):
        self.ctx = ctx or default_context()  # host-slice API: device 0 unless a context is given

    # reference: GemmExecutor::kernel_type / b_panel_width are CPU-kernel properties; the GPU analogue is the tile.
    def kernel_name(self) -> str:
        return self.ctx.last_kernel

    def prepack_a(self, a):
        return a

    def prepack_b(self, b):
        return b

    def gemm_bias(self, out, out_row_stride, a, b, alpha=1.0, beta=0.0, bias=None):
        pa, m, k, ars, acs, da = _mat(a)
        pb, kb, n, brs, bcs, db = _mat(b)
        if isinstance(out, np.ndarray):
            if out.dtype != np.float32 or not out.flags.c_contiguous:
                raise N.RtgError(2, "output must be a contiguous float32 buffer")
            po, out_len, do = out.ctypes.data, out.size, N.RTG_HOST
        else:
            po, out_len, do = out.data_ptr(), out.numel(), (out.device.index if out.is_cuda else N.RTG_HOST)
        if not (da == db == do):
            raise N.RtgError(5, "a, b and out must live on the same device")
        pbias, blen = None, 0
        keep = None
        if bias is not None:
            if isinstance(bias, np.ndarray):
                keep = np.ascontiguousarray(bias, dtype=np.float32)
                pbias, blen = keep.ctypes.data, keep.size
            else:
                keep = bias.contiguous()
                pbias, blen = keep.data_ptr(), keep.numel()
        N.check(N.lib().rtg_gemm(
            self.ctx.handle, C.c_int32(da), C.c_void_p(po), C.c_size_t(out_len), C.c_size_t(out_row_stride),
            C.c_void_p(pa), C.c_size_t(m), C.c_size_t(k), C.c_ssize_t(ars), C.c_ssize_t(acs),
            C.c_void_p(pb), C.c_size_t(kb), C.c_size_t(n), C.c_ssize_t(brs), C.c_ssize_t(bcs),
            C.c_float(alpha), C.c_float(beta), C.c_void_p(pbias), C.c_size_t(blen)))
        return out

    def gemm_static_scratch(self, out, out_row_stride, a, b, alpha=1.0, beta=0.0, bias=None, splits=0):
        """``rtg_gemm_static_scratch``: split-K GEMM over device tensors whose scratch is the library's static device
        array (no allocation of any kind).  ``a`` / ``b`` dense row-major CUDA tensors, ``out`` a flat CUDA buffer."""
        for t in (a, b, out) + ((bias,) if bias is not None else ()):
            if not (hasattr(t, "is_cuda") and t.is_cuda and t.is_contiguous()):
                raise N.RtgError(5, "gemm_static_scratch takes contiguous CUDA tensors")
        m, k = a.shape
        kb, n = b.shape
        if k != kb:
            raise N.RtgError(50, "Columns of matrix `a` must match rows of matrix `b`")
        N.check(N.lib().rtg_gemm_static_scratch(
            self.ctx.handle, C.c_void_p(out.data_ptr()), C.c_size_t(out.numel()), C.c_size_t(out_row_stride),
            C.c_void_p(a.data_ptr()), C.c_size_t(m), C.c_size_t(k), C.c_void_p(b.data_ptr()), C.c_size_t(n),
            C.c_float(alpha), C.c_float(beta), C.c_void_p(bias.data_ptr() if bias is not None else None),
            C.c_int32(splits)))
        return out

    def gemm(self, out, out_row_stride, a, b, alpha=1.0, beta=0.0):
        return self.gemm_bias(out, out_row_stride, a, b, alpha, beta, None)

    # uninit variants: beta is 0 and the output is never read
    def gemm_uninit(self, out, out_row_stride, a, b, alpha=1.0):
        return self.gemm_bias(out, out_row_stride, a, b, alpha, 0.0, None)

    def gemm_uninit_bias(self, out, out_row_stride, a, b, alpha=1.0, bias=None):
        return self.gemm_bias(out, out_row_stride, a, b, alpha, 0.0, bias)


def gemm(a, b, alpha=1.0, beta=0.0, bias=None, out=None, ctx: Context | None = None):
    """Convenience wrapper for host arrays: returns a new ``[m, n]`` array."""
    m, n = a.shape[0], b.shape[1]
    if out is None:
        out = np.zeros((m, n), dtype=np.float32)
    GemmExecutor(ctx).gemm_bias(out.reshape(-1), n, a, b, alpha, beta, bias)
    return out
