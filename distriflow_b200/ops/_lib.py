"""ctypes loader for the in-tree native library `libdistriflow_b200.so`.

The library is plain CUDA C++ with an `extern "C"` surface (declared in
`include/distriflow_b200.h`); it does not link against libtorch, so it builds in seconds with
nvcc alone (`python -m distriflow_b200.build` or `__graft_entry__.build()`), ships to the GPU
box inside the source tree, and shows up in the process's loaded-library list.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# DFB_LIB_PATH: another build of the same library (A/B runs of two kernel versions on one box); default: the in-tree build
LIB_PATH = os.environ.get("DFB_LIB_PATH") or os.path.join(_HERE, "libdistriflow_b200.so")
_lib: Optional[ctypes.CDLL] = None
_err: Optional[str] = None

c_void_p, c_int, c_int64, c_float, c_size_t = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_size_t


class NativeError(RuntimeError):
    pass


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> ctypes.CDLL:
    """Load (once) and return the library; raises loudly when it is missing."""
    global _lib, _err
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NativeError(f"{LIB_PATH} not built — run `python -m distriflow_b200.build` (needs nvcc)")
    try:
        _lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    except OSError as e:
        raise NativeError(f"cannot load {LIB_PATH}: {e}") from None
    _lib.dfb_last_error.restype = ctypes.c_char_p
    return _lib


LAUNCHES = [0]        # native calls issued (each is at least one kernel launch of this library); bench.py reports it


def check(rc: int, what: str = "") -> None:
    LAUNCHES[0] += 1
    if rc != 0:
        msg = lib().dfb_last_error()
        raise NativeError(f"{what} failed (code {rc}): {msg.decode() if msg else '?'}")


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


DT_F32, DT_BF16, DT_F16 = 0, 1, 2
_DT = {torch.float32: DT_F32, torch.bfloat16: DT_BF16, torch.float16: DT_F16}


def dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise NativeError(f"unsupported dtype {t.dtype}") from None
