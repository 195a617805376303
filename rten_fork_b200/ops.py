"""Host-side mirror of the reference operator functions for the dense hot path.

Same names, argument meaning and error behaviour as the reference's free functions / operator structs:

* ``conv``           src/ops/conv.rs:86-280    (``Conv`` struct :282-311)
* ``matmul``         src/ops/matmul.rs:119-239 (``MatMul`` :241-254)
* ``gemm_op``        src/ops/matmul.rs:27-81   (``Gemm`` :13-19)
* ``relu`` ...       src/ops/unary_elementwise.rs, ``add`` ... src/ops/binary_elementwise.rs
* ``max_pool`` ...   src/ops/pooling.rs

Inputs are numpy arrays (host; staged inside the call) or torch CUDA tensors (used in place).  Outputs have the type
of the first input.  Errors raise :class:`RtgError` whose ``kind`` / ``msg`` equal the reference's ``OpError``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N
from .context import context_for, Context, as_tensor, default_context

ACT = {"none": 0, "relu": 1, "clip": 2, "gelu": 3, "sigmoid": 4, "silu": 5, "erf": 6, "tanh": 7, "sqrt": 8}
BIN = {"add": 0, "sub": 1, "mul": 2, "div": 3, "pow": 4}


def _is_torch(x):
    return not isinstance(x, np.ndarray)


def _empty_like_kind(x, shape, nhwc=False):
    """Allocate the output next to ``x``: numpy for host inputs, torch CUDA for device inputs."""
    shape = tuple(int(s) for s in shape)
    if not _is_torch(x):
        return np.empty(shape, dtype=np.float32)
    import torch

    out = torch.empty(shape, dtype=torch.float32, device=x.device)
    if nhwc and len(shape) == 4:
        out = torch.empty((shape[0], shape[2], shape[3], shape[1]), dtype=torch.float32, device=x.device).permute(0, 3, 1, 2)
    return out


def _conv_attrs(padding, groups, strides, dilations):
    a = N.ConvAttrs()
    a.groups = int(groups)
    if isinstance(padding, str):
        if padding.lower() != "same":
            raise N.RtgError(5, "unknown padding mode")
        a.pad_mode, a.n_pads = 1, 0
    else:
        pads = list(padding)
        if len(pads) > 4:
            raise N.RtgError(5, "Expected 4 padding values")
        a.pad_mode, a.n_pads = 0, len(pads)
        for i, p in enumerate(pads):
            a.pads[i] = int(p)
    strides, dilations = list(strides), list(dilations)
    a.n_strides, a.n_dilations = min(len(strides), 3), min(len(dilations), 3)
    for i, s in enumerate(strides[:2]):
        a.strides[i] = int(s)
    for i, d in enumerate(dilations[:2]):
        a.dilations[i] = int(d)
    # lengths other than the expected ones are reported by the library with the reference's messages
    a.n_strides, a.n_dilations = len(strides), len(dilations)
    return a


def conv_output_shape(input, kernel, bias=None, padding=(0, 0, 0, 0), groups=1, strides=(1, 1), dilations=(1, 1)):
    a = _conv_attrs(padding, groups, strides, dilations)
    ti, tw = as_tensor(input), as_tensor(kernel)
    tb = as_tensor(bias) if bias is not None else None
    shape = (C.c_int64 * 4)()
    nd = C.c_int32()
    N.check(N.lib().rtg_conv_out_shape(C.byref(ti), C.byref(tw), C.byref(tb) if tb is not None else None, C.byref(a),
                                       shape, C.byref(nd)))
    return tuple(shape[: nd.value])


def conv(input, kernel, bias=None, padding=(0, 0, 0, 0), groups=1, strides=(1, 1), dilations=(1, 1), *,
         residual=None, activation="none", clip=(0.0, 0.0), out=None, out_nhwc=False, ctx: Context | None = None):
    """``conv(pool, input, kernel, bias, padding, groups, strides, dilations)``.

    ``padding`` is ``"same"`` (ONNX SAME_UPPER) or ``[top, left, bottom, right]`` (``[left, right]`` for 1-D input).
    ``residual`` / ``activation`` select the fused epilogue ``act(conv + bias + residual)``; the defaults give the
    reference operator.
    """
    ctx = ctx or context_for(input)
    a = _conv_attrs(padding, groups, strides, dilations)
    ti, tw = as_tensor(input), as_tensor(kernel)
    tb = as_tensor(bias) if bias is not None else None
    tr = as_tensor(residual) if residual is not None else None
    shape = (C.c_int64 * 4)()
    nd = C.c_int32()
    N.check(N.lib().rtg_conv_out_shape(C.byref(ti), C.byref(tw), C.byref(tb) if tb is not None else None, C.byref(a),
                                       shape, C.byref(nd)))
    if out is None:
        out = _empty_like_kind(input, shape[: nd.value], nhwc=out_nhwc)
    to = as_tensor(out)
    N.check(N.lib().rtg_conv(ctx.handle, C.byref(ti), C.byref(tw), C.byref(tb) if tb is not None else None,
                             C.byref(a), C.byref(tr) if tr is not None else None, C.c_int32(ACT[activation]),
                             C.c_float(clip[0]), C.c_float(clip[1]), C.byref(to)))
    return out


PATCH_KINDS = {"im2col": 0, "window": 1, "nhwc": 2, "nchw_round_trip": 3}


def conv_patches(kind, input, kernel, padding=(0, 0, 0, 0), groups=1, strides=(1, 1), dilations=(1, 1), group=0,
                 ctx: Context | None = None):
    """``rtg_conv_patches``: the index-driven layouts the convolution path builds (the software counterpart of
    ``VirtualIm2Col``, src/ops/conv/im2col.rs:75-260), as a host array, for bit-exact comparison with the oracle."""
    ctx = ctx or context_for(input)
    a = _conv_attrs(padding, groups, strides, dilations)
    ti, tw = as_tensor(input), as_tensor(kernel)
    shape = (C.c_int64 * 4)()
    nd = C.c_int32()
    N.check(N.lib().rtg_conv_out_shape(C.byref(ti), C.byref(tw), None, C.byref(a), shape, C.byref(nd)))
    n, c = input.shape[0], input.shape[1]
    spatial = 1
    for d in input.shape[2:]:
        spatial *= d
    out_pix = 1
    for d in list(shape)[2: nd.value]:
        out_pix *= d
    kh, kw = (kernel.shape[2], kernel.shape[3]) if kernel.ndim == 4 else (1, kernel.shape[2])
    cap = max(n * out_pix * kh * kw * ((c // groups + 31) // 32 * 32),          # patch matrix
              n * ((c + 3) // 4 * 4) * (spatial + 64) * 4 + 4096, 16)            # padded image (generous)
    pw = 0
    if kind == "window":
        pads = [0, 0, 0, 0] if isinstance(padding, str) else list(padding)
        cap = max(cap, n * ((c + 3) // 4 * 4) * (input.shape[2] + 64) * (input.shape[3] * 2 + 64 + 8 * 32))
    buf = np.empty(cap, dtype=np.float32)
    dims = (C.c_int64 * 4)()
    N.check(N.lib().rtg_conv_patches(ctx.handle, C.c_int32(PATCH_KINDS[kind]), C.byref(ti), C.byref(tw), C.byref(a),
                                     C.c_int32(group), buf.ctypes.data_as(C.c_void_p), C.c_size_t(cap), dims))
    shp = tuple(int(d) for d in dims)
    cnt = int(np.prod(shp))
    return buf[:cnt].reshape(shp).copy()


class Conv:
    """``struct Conv {groups, dilations, padding, strides}`` + ``Operator::run`` (src/ops/conv.rs:282-311)."""

    name = "Conv"

    def __init__(self, groups=1, dilations=(1, 1), padding=(0, 0, 0, 0), strides=(1, 1)):
        self.groups, self.dilations, self.padding, self.strides = groups, dilations, padding, strides

    def run(self, inputs, ctx=None):
        if len(inputs) < 2 or inputs[0] is None or inputs[1] is None:
            raise N.RtgError(4, "Required inputs were missing")
        bias = inputs[2] if len(inputs) > 2 else None
        return [conv(inputs[0], inputs[1], bias, self.padding, self.groups, self.strides, self.dilations, ctx=ctx)]


def _conv_transpose_attrs(padding, strides):
    a = N.ConvTransposeAttrs()
    if isinstance(padding, str):
        if padding.lower() != "same":
            raise N.RtgError(5, "unknown padding mode")
        a.pad_mode, a.n_pads = 1, 0
    else:
        pads = list(padding)
        a.pad_mode, a.n_pads = 0, min(len(pads), 5)
        for i, p in enumerate(pads[:4]):
            a.pads[i] = int(p)
    strides = list(strides)
    a.n_strides = min(len(strides), 3)
    for i, s in enumerate(strides[:2]):
        a.strides[i] = int(s)
    return a


def conv_transpose(input, kernel, bias=None, padding=(0, 0, 0, 0), strides=(1, 1), *, out=None,
                   ctx: Context | None = None):
    """``conv_transpose(pool, input, kernel, bias, padding, strides)`` (src/ops/conv.rs:441-545): input NCHW / NCW,
    kernel ``[C, O, Kh, Kw]`` / ``[C, O, Kw]``.  One tcgen05 GEMM over the whole batch, then the col2im gather."""
    ctx = ctx or context_for(input)
    a = _conv_transpose_attrs(padding, strides)
    ti, tw = as_tensor(input), as_tensor(kernel)
    tb = as_tensor(bias) if bias is not None else None
    shape = (C.c_int64 * 4)()
    nd = C.c_int32()
    N.check(N.lib().rtg_conv_transpose_out_shape(C.byref(ti), C.byref(tw), C.byref(tb) if tb is not None else None,
                                                 C.byref(a), shape, C.byref(nd)))
    if out is None:
        out = _empty_like_kind(input, shape[: nd.value])
    to = as_tensor(out)
    N.check(N.lib().rtg_conv_transpose(ctx.handle, C.byref(ti), C.byref(tw), C.byref(tb) if tb is not None else None,
                                       C.byref(a), C.byref(to)))
    return out


def conv_transpose_output_shape(input, kernel, bias=None, padding=(0, 0, 0, 0), strides=(1, 1)):
    a = _conv_transpose_attrs(padding, strides)
    ti, tw = as_tensor(input), as_tensor(kernel)
    tb = as_tensor(bias) if bias is not None else None
    shape = (C.c_int64 * 4)()
    nd = C.c_int32()
    N.check(N.lib().rtg_conv_transpose_out_shape(C.byref(ti), C.byref(tw), C.byref(tb) if tb is not None else None,
                                                 C.byref(a), shape, C.byref(nd)))
    return tuple(shape[: nd.value])


def matmul(a, b, *, out=None, ctx: Context | None = None):
    ctx = ctx or context_for(a)
    ta, tb = as_tensor(a), as_tensor(b)
    shape = (C.c_int64 * N.RTG_MAX_DIMS)()
    nd = C.c_int32()
    N.check(N.lib().rtg_matmul_out_shape(C.byref(ta), C.byref(tb), shape, C.byref(nd)))
    if out is None:
        out = _empty_like_kind(a, shape[: nd.value])
    to = as_tensor(out)
    N.check(N.lib().rtg_matmul(ctx.handle, C.byref(ta), C.byref(tb), C.byref(to)))
    return out


class MatMul:
    name = "MatMul"

    def run(self, inputs, ctx=None):
        if len(inputs) < 2 or inputs[0] is None or inputs[1] is None:
            raise N.RtgError(4, "Required inputs were missing")
        return [matmul(inputs[0], inputs[1], ctx=ctx)]


def gemm_op(a, b, c=None, alpha=1.0, beta=1.0, transpose_a=False, transpose_b=False, *, ctx: Context | None = None):
    ctx = ctx or context_for(a)
    ta, tb = as_tensor(a), as_tensor(b)
    tc = as_tensor(c) if c is not None else None
    if ta.ndim != 2:
        raise N.RtgError(5, "a must have 2 dims")
    if tb.ndim != 2:
        raise N.RtgError(5, "b must have 2 dims")
    m = a.shape[1] if transpose_a else a.shape[0]
    n = b.shape[0] if transpose_b else b.shape[1]
    out = _empty_like_kind(a, (m, n))
    to = as_tensor(out)
    N.check(N.lib().rtg_gemm_op(ctx.handle, C.byref(ta), C.byref(tb), C.byref(tc) if tc is not None else None,
                                C.c_float(alpha), C.c_float(beta), C.c_int32(bool(transpose_a)),
                                C.c_int32(bool(transpose_b)), C.byref(to)))
    return out


class Gemm:
    name = "Gemm"

    def __init__(self, alpha=1.0, beta=1.0, transpose_a=False, transpose_b=False):
        self.alpha, self.beta, self.transpose_a, self.transpose_b = alpha, beta, transpose_a, transpose_b

    def run(self, inputs, ctx=None):
        if len(inputs) < 2 or inputs[0] is None or inputs[1] is None:
            raise N.RtgError(4, "Required inputs were missing")
        c = inputs[2] if len(inputs) > 2 else None
        return [gemm_op(inputs[0], inputs[1], c, self.alpha, self.beta, self.transpose_a, self.transpose_b, ctx=ctx)]


def _unary(kind, x, lo=0.0, hi=0.0, in_place=False, ctx=None):
    ctx = ctx or context_for(x)
    ti = as_tensor(x)
    out = x if in_place else _empty_like_kind(x, x.shape)
    if in_place and not _is_torch(x) and not x.flags.c_contiguous:
        raise N.RtgError(6, "in-place host tensors must be contiguous")
    to = as_tensor(out)
    N.check(N.lib().rtg_unary(ctx.handle, C.c_int32(ACT[kind]), C.byref(ti), C.c_float(lo), C.c_float(hi), C.byref(to)))
    return out


def relu(x, **kw):
    return _unary("relu", x, **kw)


def gelu(x, **kw):
    return _unary("gelu", x, **kw)


def sigmoid(x, **kw):
    return _unary("sigmoid", x, **kw)


def silu(x, **kw):
    return _unary("silu", x, **kw)


def erf(x, **kw):
    return _unary("erf", x, **kw)


def tanh(x, **kw):
    return _unary("tanh", x, **kw)


def sqrt(x, **kw):
    return _unary("sqrt", x, **kw)


def clip(x, min_val=None, max_val=None, **kw):
    lo = float(np.finfo(np.float32).min) if min_val is None else float(min_val)
    hi = float(np.finfo(np.float32).max) if max_val is None else float(max_val)
    return _unary("clip", x, lo, hi, **kw)


def _binary(kind, a, b, out=None, ctx=None):
    # numpy hands back scalars (np.float32) for 0-d results; describe them as 0-d arrays kept alive for the call
    if isinstance(a, np.generic):
        a = np.asarray(a)
    if isinstance(b, np.generic):
        b = np.asarray(b)
    ctx = ctx or context_for(a)
    ta, tb = as_tensor(a), as_tensor(b)
    shape = (C.c_int64 * N.RTG_MAX_DIMS)()
    nd = C.c_int32()
    N.check(N.lib().rtg_binary_out_shape(C.byref(ta), C.byref(tb), shape, C.byref(nd)))
    if out is None:
        out = _empty_like_kind(a, shape[: nd.value])
    to = as_tensor(out)
    N.check(N.lib().rtg_binary(ctx.handle, C.c_int32(BIN[kind]), C.byref(ta), C.byref(tb), C.byref(to)))
    return out


def add(a, b, **kw):
    return _binary("add", a, b, **kw)


def sub(a, b, **kw):
    return _binary("sub", a, b, **kw)


def mul(a, b, **kw):
    return _binary("mul", a, b, **kw)


def div(a, b, **kw):
    return _binary("div", a, b, **kw)


def pow(a, b, **kw):  # noqa: A001 - mirrors the reference operator name
    return _binary("pow", a, b, **kw)


def _pool_attrs(kernel_size, strides, padding, count_include_pad=False):
    a = N.PoolAttrs()
    a.kernel[0], a.kernel[1] = int(kernel_size[0]), int(kernel_size[1])
    a.strides[0], a.strides[1] = int(strides[0]), int(strides[1])
    if isinstance(padding, str):
        a.pad_mode = 1
    else:
        if len(padding) != 4:
            raise N.RtgError(5, "Expected 4 padding values")
        a.pad_mode = 0
        for i, p in enumerate(padding):
            a.pads[i] = int(p)
    a.count_include_pad = int(bool(count_include_pad))
    return a


def _pool(kind, x, kernel_size, strides, padding, count_include_pad=False, ctx=None, out=None):
    ctx = ctx or context_for(x)
    a = _pool_attrs(kernel_size, strides, padding, count_include_pad)
    ti = as_tensor(x)
    shape = (C.c_int64 * 4)()
    N.check(N.lib().rtg_pool2d_out_shape(C.byref(ti), C.byref(a), shape))
    if out is None:
        out = _empty_like_kind(x, shape[:4])
    elif tuple(out.shape) != tuple(shape[:4]):
        raise N.RtgError(3, "output has the wrong shape")
    to = as_tensor(out)
    N.check(N.lib().rtg_pool2d(ctx.handle, C.c_int32(kind), C.byref(ti), C.byref(a), C.byref(to)))
    return out


def max_pool(x, kernel_size, strides, padding=(0, 0, 0, 0), **kw):
    return _pool(0, x, kernel_size, strides, padding, **kw)


def average_pool(x, kernel_size, strides, padding=(0, 0, 0, 0), count_include_pad=False, **kw):
    return _pool(1, x, kernel_size, strides, padding, count_include_pad, **kw)


def global_average_pool(x, *, ctx=None):
    ctx = ctx or context_for(x)
    ti = as_tensor(x)
    if ti.ndim != 4:
        raise N.RtgError(5, "input must have 4 dims (NCHW)")
    out = _empty_like_kind(x, (x.shape[0], x.shape[1], 1, 1))
    to = as_tensor(out)
    N.check(N.lib().rtg_global_average_pool(ctx.handle, C.byref(ti), C.byref(to)))
    return out
