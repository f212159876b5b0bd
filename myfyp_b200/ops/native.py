"""ctypes binding to ``libmyfyp_b200.so`` (hand-written sm_100a kernels, C ABI in include/myfyp_b200.h).

The library is built in-tree by ``python -m myfyp_b200.build`` (plain ``nvcc``; no torch headers, so
it compiles in seconds and any C caller can use it).  It loads without a GPU; calling a kernel
without one raises.  On a CUDA tensor the ops in this package NEVER fall back to PyTorch: a missing
library is an error, so a silent eager path cannot masquerade as the native one.
"""

from __future__ import annotations

import ctypes
import os
import threading
from typing import Optional, Sequence

import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib", "libmyfyp_b200.so")
_lib: Optional[ctypes.CDLL] = None
_lock = threading.Lock()

# launch counter: bench.py reports how many of OUR kernels ran in the timed region
_launches = 0


class NativeLibraryError(RuntimeError):
    pass


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(_LIB_PATH):
                    raise NativeLibraryError(f"{_LIB_PATH} not built - run `python -m myfyp_b200.build` (needs nvcc)")
                lib = ctypes.CDLL(_LIB_PATH)
                lib.b2_last_error.restype = ctypes.c_char_p
                lib.b2_version.restype = ctypes.c_int
                _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b2_last_error()
        raise NativeLibraryError(f"{what} failed (rc={rc}): {msg.decode() if msg else '?'}")


def count_launch(n: int = 1) -> None:
    global _launches
    _launches += n


def launches() -> int:
    return _launches


def reset_launches() -> None:
    global _launches
    _launches = 0


def stream_ptr(device: torch.device) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t: Optional[torch.Tensor]) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def ptr_array(ts: Sequence[torch.Tensor]):
    arr = (ctypes.c_void_p * len(ts))()
    for i, t in enumerate(ts):
        arr[i] = t.data_ptr()
    return arr


def float_array(vals: Sequence[float]):
    return (ctypes.c_float * len(vals))(*[float(v) for v in vals])


class device_guard:
    """Make ``device`` current for the native call (kernels launch on the *current* device)."""

    def __init__(self, device: torch.device) -> None:
        self.device = device
        self._g = None

    def __enter__(self):
        self._g = torch.cuda.device(self.device)
        self._g.__enter__()
        return self

    def __exit__(self, *a):
        return self._g.__exit__(*a)
