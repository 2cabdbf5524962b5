"""ctypes binding of libssb200.so (the C ABI declared in include/ssb200.h).

Fails loudly: a missing library, a missing symbol or a non-zero return code raises.
There is no eager/PyTorch/CPU fallback anywhere in this package."""
from __future__ import annotations

import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libssb200.so")
_lib = None


class SsbError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SsbError(f"{LIB_PATH} is missing: run `python -m silent_speech_b200.build` "
                           "(there is no fallback path)")
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.ssb_last_error.restype = ctypes.c_char_p
    return _lib


_launches = 0


def launch_count() -> int:
    """Number of C-ABI compute calls made so far; every one of them enqueues exactly one kernel."""
    return _launches


def check(code: int) -> None:
    global _launches
    _launches += 1
    if code != 0:
        raise SsbError(f"libssb200 error {code}: {lib().ssb_last_error().decode()}")


def ptr(t):
    """Device pointer of a tensor (or NULL)."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def ptr_array(tensors):
    """Host array of device pointers."""
    arr = (ctypes.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = 0 if t is None else t.data_ptr()
    return arr


i32, i64, f32, u64 = ctypes.c_int32, ctypes.c_int64, ctypes.c_float, ctypes.c_uint64


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
