"""ctypes loader for ``librten_gpu.so`` (the C-ABI declared in ``include/rten_gpu.h``).

There is deliberately no fallback: if the shared library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RTG_PROF_LIB=1 selects the developer build with per-role cycle counters (`make PROF=1`), RTG_DEV_LIB=<file name> any
# other developer build next to the package (A/B experiments); never the default.
LIB_PATH = os.path.join(_HERE, os.environ.get("RTG_DEV_LIB") or (
    "librten_gpu_prof.so" if os.environ.get("RTG_PROF_LIB") == "1" else "librten_gpu.so"))

RTG_HOST = -1
RTG_F32, RTG_I32 = 0, 1
RTG_MAX_DIMS = 8
MATH_TF32, MATH_3XTF32, MATH_TF32_BF16X2 = 0, 1, 2

STATUS_NAMES = {
    0: "Ok", 1: "IncorrectInputType", 2: "IncorrectOutputType", 3: "IncompatibleInputShapes", 4: "MissingInputs",
    5: "InvalidValue", 6: "UnsupportedValue", 50: "Panic", 100: "Cuda", 101: "Nccl", 102: "Internal", 103: "Format",
}


class RtgError(RuntimeError):
    """Mirrors ``OpError`` (src/ops/mod.rs:665-705): ``kind`` is the variant name, ``msg`` the reference's message."""

    def __init__(self, code: int, msg: str):
        self.code, self.kind, self.msg = code, STATUS_NAMES.get(code, str(code)), msg
        super().__init__(f"{self.kind}: {msg}")


class Tensor(C.Structure):
    _fields_ = [
        ("data", C.c_void_p),
        ("dtype", C.c_int32),
        ("ndim", C.c_int32),
        ("shape", C.c_int64 * RTG_MAX_DIMS),
        ("strides", C.c_int64 * RTG_MAX_DIMS),
        ("device", C.c_int32),
    ]


class ConvAttrs(C.Structure):
    _fields_ = [
        ("groups", C.c_int32),
        ("pad_mode", C.c_int32),
        ("pads", C.c_int64 * 4),
        ("n_pads", C.c_int32),
        ("strides", C.c_int64 * 2),
        ("n_strides", C.c_int32),
        ("dilations", C.c_int64 * 2),
        ("n_dilations", C.c_int32),
    ]


class ConvTransposeAttrs(C.Structure):
    _fields_ = [
        ("pad_mode", C.c_int32),
        ("pads", C.c_int64 * 4),
        ("n_pads", C.c_int32),
        ("strides", C.c_int64 * 2),
        ("n_strides", C.c_int32),
    ]


class PoolAttrs(C.Structure):
    _fields_ = [
        ("kernel", C.c_int64 * 2),
        ("strides", C.c_int64 * 2),
        ("pad_mode", C.c_int32),
        ("pads", C.c_int64 * 4),
        ("count_include_pad", C.c_int32),
    ]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.rtg_last_error.restype = C.c_char_p
        _lib.rtg_ctx_last_kernel.restype = C.c_char_p
        _lib.rtg_ctx_last_kernel.argtypes = [C.c_void_p]
        for name in dir(_lib):
            pass
    return _lib


def check(status: int):
    if status != 0:
        raise RtgError(status, lib().rtg_last_error().decode())
