"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes front-end for ``oracle/liboracle.so``, the C++ CPU restatement of the RTen
0.12.0 dense hot path (GemmExecutor, Conv, MatMul/Gemm, elementwise epilogues).
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module; the product package
``rten_fork_b200`` never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

KERNEL_AUTO, KERNEL_BASE, KERNEL_FMA, KERNEL_AVX512 = -1, 0, 1, 2
UNARY = {"relu": 0, "gelu": 1, "sigmoid": 2, "erf": 3, "silu": 4, "clip": 5}
BINARY = {"add": 0, "sub": 1, "mul": 2, "div": 3}

_f32p = C.POINTER(C.c_float)
_szp = C.POINTER(C.c_size_t)
_i32p = C.POINTER(C.c_int32)
_pdp = C.POINTER(C.c_ssize_t)


def build(force: bool = False) -> str:
    """Compile liboracle.so with the committed Makefile (gcc only, no reference sources)."""
    if force or not os.path.exists(_LIB_PATH):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-j8"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_last_error.restype = C.c_char_p
        _lib.orc_vm_scalar.restype = C.c_float
        _lib.orc_vm_scalar.argtypes = [C.c_int, C.c_float]
    return _lib


class OracleError(Exception):
    def __init__(self, code, msg):
        super().__init__(f"[{code}] {msg}")
        self.code, self.msg = code, msg


def _check(rc):
    if rc != 0:
        raise OracleError(rc, lib().orc_last_error().decode() if rc > 0 else f"panic({rc})")


def _fp(a):
    return a.ctypes.data_as(_f32p) if a is not None else None


def _sz(vals):
    return (C.c_size_t * len(vals))(*[int(v) for v in vals])


def _pd(vals):
    return (C.c_ssize_t * len(vals))(*[int(v) for v in vals])


def _estrides(a):
    return [s // a.itemsize for s in a.strides]


def set_num_threads(n: int):
    lib().orc_set_num_threads(int(n))


def num_threads() -> int:
    return lib().orc_num_threads()


class XorShiftRng:
    """rten-tensor/src/rng.rs:6-33."""

    def __init__(self, seed: int):
        self.state = C.c_uint64(seed)

    def fill(self, shape) -> np.ndarray:
        out = np.empty(shape, dtype=np.float32)
        lib().orc_rng_fill(C.byref(self.state), _fp(out), C.c_size_t(out.size))
        return out


def kernel_info(kernel=KERNEL_AUTO):
    mr, nr = C.c_size_t(), C.c_size_t()
    t = lib().orc_kernel_info(kernel, C.byref(mr), C.byref(nr))
    return None if t < 0 else (t, mr.value, nr.value)


def gemm(a, b, out=None, alpha=1.0, beta=0.0, bias=None, kernel=KERNEL_AUTO, prepack=0, out_row_stride=None):
    """GemmExecutor::gemm_bias.  a [M,K], b [K,N] may be arbitrary strided f32 views."""
    assert a.dtype == np.float32 and b.dtype == np.float32
    m, k = a.shape
    k2, n = b.shape
    rs = n if out_row_stride is None else out_row_stride
    if out is None:
        out = np.zeros(max(0, (m - 1) * rs + n) if m else 0, dtype=np.float32)
    flat = out.reshape(-1)
    ars, acs = _estrides(a)
    brs, bcs = _estrides(b)
    bias_c = np.ascontiguousarray(bias, dtype=np.float32) if bias is not None else None
    rc = lib().orc_gemm(C.c_int(kernel), _fp(flat), C.c_size_t(flat.size), C.c_size_t(rs), _fp(a), C.c_size_t(m),
                        C.c_size_t(k), C.c_ssize_t(ars), C.c_ssize_t(acs), _fp(b), C.c_size_t(k2), C.c_size_t(n),
                        C.c_ssize_t(brs), C.c_ssize_t(bcs), C.c_float(alpha), C.c_float(beta), _fp(bias_c),
                        C.c_size_t(0 if bias_c is None else bias_c.size), C.c_int(prepack))
    if rc != 0:
        raise OracleError(rc, {-1: "Columns of matrix `a` must match rows of matrix `b`",
                               -2: "Bias vector length must match rows of matrix `a`",
                               -3: "Output buffer should be large enough"}.get(rc, "panic"))
    if out_row_stride is None and out.ndim == 1:
        return out.reshape(m, n)
    return out


def pack_a_block(a, mr, rows, cols):
    r0, r1 = rows
    c0, c1 = cols
    out = np.empty(((r1 - r0 + mr - 1) // mr) * mr * (c1 - c0), dtype=np.float32)
    rs, cs = _estrides(a)
    lib().orc_pack_a_block(C.c_size_t(mr), _fp(out), _fp(a), C.c_size_t(a.shape[0]), C.c_size_t(a.shape[1]),
                           C.c_ssize_t(rs), C.c_ssize_t(cs), C.c_size_t(r0), C.c_size_t(r1), C.c_size_t(c0),
                           C.c_size_t(c1))
    return out


def pack_b_block(b, nr, rows, cols):
    r0, r1 = rows
    c0, c1 = cols
    out = np.empty(((c1 - c0 + nr - 1) // nr) * nr * (r1 - r0), dtype=np.float32)
    rs, cs = _estrides(b)
    lib().orc_pack_b_block(C.c_size_t(nr), _fp(out), _fp(b), C.c_size_t(b.shape[0]), C.c_size_t(b.shape[1]),
                           C.c_ssize_t(rs), C.c_ssize_t(cs), C.c_size_t(r0), C.c_size_t(r1), C.c_size_t(c0),
                           C.c_size_t(c1))
    return out


def conv_out_size(in_hw, k_hw, strides, padding, dilations=(1, 1)):
    """calc_output_size_and_padding.  padding = 'same' or [top,left,bottom,right]."""
    same = padding == "same"
    pads = _sz([0, 0, 0, 0] if same else padding)
    oh, ow = C.c_size_t(), C.c_size_t()
    fixed = _sz([0, 0, 0, 0])
    _check(lib().orc_conv_out_size(C.c_size_t(in_hw[0]), C.c_size_t(in_hw[1]), C.c_size_t(k_hw[0]),
                                   C.c_size_t(k_hw[1]), C.c_size_t(strides[0]), C.c_size_t(strides[1]),
                                   C.c_int(same), pads, C.c_size_t(dilations[0]), C.c_size_t(dilations[1]),
                                   C.byref(oh), C.byref(ow), fixed))
    return oh.value, ow.value, list(fixed)


def conv2d(x, w, bias=None, padding=(0, 0, 0, 0), groups=1, strides=(1, 1), dilations=(1, 1), kernel=KERNEL_AUTO,
           alloc=None):
    """conv() for NCHW / OIHW inputs, including the 1-D lift (src/ops/conv.rs:98-139).  ``alloc(shape)`` supplies
    the output buffer (the reference allocates it from its TensorPool, src/ops/conv.rs:212-220)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    if x.ndim == 3:
        if w.ndim != 3:
            raise OracleError(5, "kernel must have 3 dims (OCW)")
        if padding != "same" and len(padding) != 2:
            raise OracleError(5, "expected 2 pad values")
        if len(strides) != 1:
            raise OracleError(5, "expected 1 stride value")
        if len(dilations) != 1:
            raise OracleError(5, "expected 1 dilation value")
        n, c, wd = x.shape
        pad2 = padding if padding == "same" else (0, padding[0], 0, padding[1])
        y = conv2d(x.reshape(n, c, 1, wd), w.reshape(w.shape[0], w.shape[1], 1, w.shape[2]), bias, pad2, groups,
                   (1, strides[0]), (1, dilations[0]), kernel)
        return y.reshape(y.shape[0], y.shape[1], y.shape[3])
    if x.ndim != 4:
        raise OracleError(5, "input must have 4 dims (NCHW)")
    if w.ndim != 4:
        raise OracleError(5, "kernel must have 4 dims (OCHW)")
    if len(strides) != 2:
        raise OracleError(5, "expected 2 stride values")
    if len(dilations) != 2:
        raise OracleError(5, "expected 2 dilation values")
    same = padding == "same"
    pads = _sz([0, 0, 0, 0] if same else padding)
    b = np.ascontiguousarray(bias, dtype=np.float32) if bias is not None else None
    oshape = _sz([0, 0, 0, 0])
    args = [C.c_int(kernel), _fp(x), _sz(x.shape), _fp(w), _sz(w.shape), _fp(b), C.c_size_t(0 if b is None else b.size),
            C.c_int(same), pads, C.c_size_t(groups), _sz(strides), _sz(dilations)]
    _check(lib().orc_conv2d(*args, None, oshape))
    out = alloc(tuple(oshape)) if alloc is not None else np.empty(tuple(oshape), dtype=np.float32)
    _check(lib().orc_conv2d(*args, _fp(out), oshape))
    return out


def im2col(image, k_hw, pads, strides, dilations=(1, 1), panel_width=16, pack=True):
    """VirtualIm2Col offset tables and the packed B panels for one CHW image."""
    image = np.ascontiguousarray(image, dtype=np.float32)
    c, h, w = image.shape
    oh, ow, _ = conv_out_size((h, w), k_hw, strides, list(pads), dilations)
    n_rows, n_cols = c * k_hw[0] * k_hw[1], oh * ow
    ncp = (n_cols + panel_width - 1) // panel_width * panel_width
    rc_, ry, rx = (np.empty(n_rows, np.int32) for _ in range(3))
    cy, cx = (np.empty(ncp, np.int32) for _ in range(2))
    out = np.empty(n_rows * ncp, np.float32) if pack else None
    nr, nc = C.c_size_t(), C.c_size_t()
    ip = lambda a: a.ctypes.data_as(_i32p)
    lib().orc_im2col(_fp(image), C.c_size_t(c), C.c_size_t(h), C.c_size_t(w), C.c_size_t(k_hw[0]), C.c_size_t(k_hw[1]),
                     _sz(pads), C.c_size_t(strides[0]), C.c_size_t(strides[1]), C.c_size_t(dilations[0]),
                     C.c_size_t(dilations[1]), C.c_size_t(panel_width), ip(rc_), ip(ry), ip(rx), ip(cy), ip(cx),
                     C.byref(nr), C.byref(nc), _fp(out))
    assert nr.value == n_rows and nc.value == n_cols
    return {"row_chan": rc_, "row_y": ry, "row_x": rx, "col_y": cy, "col_x": cx, "n_rows": n_rows, "n_cols": n_cols,
            "packed": out}


def unary(op, x, lo=-np.inf, hi=np.inf, in_place=False):
    """``in_place``: write into ``x`` (UnaryFloatOp::run_in_place, src/ops/unary_elementwise.rs:39-63)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = x if in_place and x.flags.writeable else np.empty_like(x)
    lo = np.float32(np.finfo(np.float32).min if lo is None or lo == -np.inf else lo)
    hi = np.float32(np.finfo(np.float32).max if hi is None or hi == np.inf else hi)
    lib().orc_unary(C.c_int(UNARY[op]), _fp(x), _fp(out), C.c_size_t(x.size), C.c_float(lo), C.c_float(hi))
    return out


def vm_scalar(name, x):
    return lib().orc_vm_scalar({"exp": 0, "sigmoid": 1, "erf": 2, "gelu": 3}[name], C.c_float(x))


def binary(op, a, b, in_place=False):
    """``in_place``: write into ``a`` when it already has the output shape (binary_op_in_place,
    src/ops/binary_elementwise.rs:261-322)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    oshape = _sz([0] * 8)
    ond = C.c_size_t()
    rc = lib().orc_binary(C.c_int(BINARY[op]), _fp(a), _sz(a.shape), C.c_size_t(a.ndim), _fp(b), _sz(b.shape),
                          C.c_size_t(b.ndim), None, oshape, C.byref(ond))
    if rc != 0:
        raise OracleError(3, "Cannot broadcast inputs to same shape")
    oshp = tuple(oshape[:ond.value])
    out = a if in_place and a.shape == oshp and a.flags.writeable else np.empty(oshp, dtype=np.float32)
    lib().orc_binary(C.c_int(BINARY[op]), _fp(a), _sz(a.shape), C.c_size_t(a.ndim), _fp(b), _sz(b.shape),
                     C.c_size_t(b.ndim), _fp(out), oshape, C.byref(ond))
    return out


def pool2d(kind, x, kernel_size, strides, padding=(0, 0, 0, 0), count_include_pad=False):
    x = np.ascontiguousarray(x, dtype=np.float32)
    same = padding == "same"
    pads = _sz([0, 0, 0, 0] if same else padding)
    oshape = _sz([0, 0, 0, 0])
    args = [C.c_int(kind == "max"), _fp(x), _sz(x.shape), _sz(kernel_size), _sz(strides), C.c_int(same), pads,
            C.c_int(count_include_pad)]
    _check(lib().orc_pool2d(*args, None, oshape))
    out = np.empty(tuple(oshape), dtype=np.float32)
    _check(lib().orc_pool2d(*args, _fp(out), oshape))
    return out


def global_average_pool(x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, c, h, w = x.shape
    out = np.empty((n, c, 1, 1), dtype=np.float32)
    lib().orc_global_average_pool(_fp(x), C.c_size_t(n), C.c_size_t(c), C.c_size_t(h), C.c_size_t(w), _fp(out))
    return out


def matmul(a, b, kernel=KERNEL_AUTO):
    assert a.dtype == np.float32 and b.dtype == np.float32
    oshape = _sz([0] * 8)
    ond = C.c_size_t()
    args = [C.c_int(kernel), _fp(a), C.c_size_t(a.ndim), _sz(a.shape), _pd(_estrides(a)), _fp(b), C.c_size_t(b.ndim),
            _sz(b.shape), _pd(_estrides(b))]
    # Shape / error query needs the data too (cheap at test sizes).
    try_shape = np.broadcast_shapes(a.shape[:-2], b.shape[:-2]) + (a.shape[-2], b.shape[-1]) \
        if a.ndim >= 2 and b.ndim >= 2 and a.shape[-1] == b.shape[-2] else None
    cap = int(np.prod(try_shape)) if try_shape is not None else 0
    out = np.empty(cap, dtype=np.float32)
    _check(lib().orc_matmul(*args, _fp(out), C.c_size_t(cap), oshape, C.byref(ond)))
    return out.reshape(tuple(oshape[:ond.value]))


def gemm_op(a, b, c=None, alpha=1.0, beta=1.0, transpose_a=False, transpose_b=False, kernel=KERNEL_AUTO):
    assert a.dtype == np.float32 and b.dtype == np.float32
    m = a.shape[1] if transpose_a else a.shape[0]
    n = b.shape[0] if transpose_b else b.shape[1]
    out = np.empty((m, n), dtype=np.float32)
    oshape = _sz([0, 0])
    cargs = [None, C.c_size_t(0), None, None]
    if c is not None:
        c = np.asarray(c, dtype=np.float32)
        cargs = [_fp(c), C.c_size_t(c.ndim), _sz(c.shape), _pd(_estrides(c))]
    _check(lib().orc_gemm_op(C.c_int(kernel), _fp(a), _sz(a.shape), _pd(_estrides(a)), _fp(b), _sz(b.shape),
                             _pd(_estrides(b)), *cargs, C.c_float(alpha), C.c_float(beta), C.c_int(transpose_a),
                             C.c_int(transpose_b), _fp(out), C.c_size_t(out.size), oshape))
    return out


def softmax(x, axis=-1):
    """src/ops/norm.rs softmax over rten-vecmath/src/softmax.rs: exp(x - max) with the vecmath exp, then / sum."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    xm = np.moveaxis(x, axis, -1)
    m = xm.max(axis=-1, keepdims=True)
    if xm.size <= 4096:
        e = np.array([vm_scalar("exp", float(v)) for v in (xm - m).reshape(-1)], dtype=np.float32).reshape(xm.shape)
    else:
        e = np.exp((xm - m).astype(np.float64)).astype(np.float32)
    s = e.sum(axis=-1, keepdims=True, dtype=np.float32)
    return np.ascontiguousarray(np.moveaxis(e / s, -1, axis))


def layer_norm(x, scale, bias, axis, eps):
    """src/ops/norm.rs:245-299: mean, centre, inverse RMS of the centred values, scale, shift (f32 throughout)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    axis = axis + x.ndim if axis < 0 else axis
    axes = tuple(range(axis, x.ndim))
    mean = x.mean(axis=axes, keepdims=True, dtype=np.float32)
    d = x - mean
    inv = (1.0 / np.sqrt((d * d).mean(axis=axes, keepdims=True, dtype=np.float32) + np.float32(eps))).astype(np.float32)
    y = d * inv * scale
    if bias is not None:
        y = y + bias
    return y.astype(np.float32)


# ---------------------------------------------------------------- ConvTranspose (test infrastructure, like the rest)
def conv_transpose_output_size_and_padding(in_hw, k_hw, padding, strides):
    """src/ops/conv.rs:377-439.  ``padding``: "same" or a list.  Returns ((out_h, out_w), pads) where ``pads`` is the
    4-tuple the reference returns -- in ITS order: [top, left, bottom, right] for fixed padding but
    [top, bottom, left, right] for "same" (:424), which conv_transpose then reads as [top, left, ...] (:511)."""
    (in_h, in_w), (k_h, k_w), (s_h, s_w) = in_hw, k_hw, strides
    if s_h == 0 or s_w == 0:
        raise OracleError(5, "Strides must be > 0")
    if in_h == 0 or in_w == 0:
        raise OracleError(5, "Input width and height must be > 0")
    if isinstance(padding, str):
        out_h, out_w = in_h * s_h, in_w * s_w
        pad_h, pad_w = (in_h - 1) * s_h + k_h - out_h, (in_w - 1) * s_w + k_w - out_w
        if pad_h < 0 or pad_w < 0:
            raise OracleError(5, "Input is too small")
        return (out_h, out_w), (pad_h // 2, -(-pad_h // 2), pad_w // 2, -(-pad_w // 2))
    if len(padding) != 4:
        raise OracleError(5, "Wrong number of pad values")
    t, l, b, r = padding
    out_h, out_w = (in_h - 1) * s_h + k_h - (t + b), (in_w - 1) * s_w + k_w - (l + r)
    if out_h < 0 or out_w < 0:
        raise OracleError(5, "Input is too small")
    return (out_h, out_w), (t, l, b, r)


def conv_transpose(x, w, bias=None, padding=(0, 0, 0, 0), strides=(1, 1)):
    """conv_transpose() (src/ops/conv.rs:441-545): per image one GemmExecutor product
    ``kernel_mat^T [O*Kh*Kw, C] . input [C, H*W]`` followed by col2im (:329-375), whose accumulation order is bias,
    then ky-major / kx-minor.  Input NCHW (or NCW), kernel [C, O, Kh, Kw]."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    if x.ndim == 3:
        if w.ndim != 3:
            raise OracleError(5, "kernel must have 3 dims (OCW)")
        if not isinstance(padding, str):
            if len(padding) != 2:
                raise OracleError(5, "expected 2 pad values")
            padding = (0, padding[0], 0, padding[1])
        if len(strides) != 1:
            raise OracleError(5, "expected 1 stride value")
        n, c, wd = x.shape
        y = conv_transpose(x.reshape(n, c, 1, wd), w.reshape(w.shape[0], w.shape[1], 1, w.shape[2]), bias, padding,
                           (1, strides[0]))
        return y.reshape(y.shape[0], y.shape[1], y.shape[3])
    if x.ndim != 4:
        raise OracleError(5, "input must have 4 dims (NCHW)")
    if w.ndim != 4:
        raise OracleError(5, "kernel must have 4 dims (OCHW)")
    n, c, h, wd = x.shape
    kc, o, kh, kw = w.shape
    if bias is not None and np.ndim(bias) != 1:
        raise OracleError(5, "bias must have 1 dims")
    if c != kc:
        raise OracleError(3, "Input channels does not match kernel input channels")
    if len(strides) != 2:
        raise OracleError(5, "expected 2 stride values")
    (oh, ow), pads = conv_transpose_output_size_and_padding((h, wd), (kh, kw), padding, strides)
    pad_t, pad_l = pads[0], pads[1]  # what col2im is handed as [top, left] (see above for "same")
    sh, sw = strides
    out = np.empty((n, o, oh, ow), dtype=np.float32)
    kmat_t = w.reshape(c, o * kh * kw).T
    for i in range(n):
        cols = gemm(kmat_t, x[i].reshape(c, h * wd)).reshape(o, kh, kw, h, wd)
        img = np.empty((o, oh, ow), dtype=np.float32)
        img[...] = 0.0 if bias is None else np.asarray(bias, dtype=np.float32)[:, None, None]
        for ky in range(kh):
            ys = np.arange(h) * sh + ky
            yok = (ys >= pad_t) & (ys < oh + pad_t)
            for kx in range(kw):
                xs = np.arange(wd) * sw + kx
                xok = (xs >= pad_l) & (xs < ow + pad_l)
                if not yok.any() or not xok.any():
                    continue
                # every (y, x) of one tap lands on a distinct output pixel, so a vectorised add keeps the order
                img[:, (ys[yok] - pad_t)[:, None], (xs[xok] - pad_l)[None, :]] += cols[:, ky, kx][:, yok][:, :, xok]
        out[i] = img
    return out
