"""ctypes loader for the native libraries (C ABI declared in include/bigdl_b200.h).

The CUDA library is mandatory on a machine with a GPU: if it is missing there, importing fails loudly instead of
silently falling back to eager PyTorch. On a CPU-only machine ``cuda_lib()`` returns ``None`` and callers use the
CPU oracle path.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

import torch

_LIBDIR = Path(__file__).resolve().parent.parent / "lib"
_cuda = None
_host = None


class NativeError(RuntimeError):
    pass


def _load(name: str):
    path = _LIBDIR / name
    if not path.exists():
        try:
            from .. import build as _build

            if name == "libbigdl_b200.so":
                _build.build_cuda()
            else:
                _build.build_host()
        except Exception as e:  # pragma: no cover - build environment problems
            raise NativeError(f"{name} is missing and could not be built: {e}") from e
    return ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)


def cuda_lib():
    """The sm_100a kernel library, or None when this process has no CUDA device."""
    global _cuda
    if _cuda is None:
        if not torch.cuda.is_available():
            return None
        _cuda = _load("libbigdl_b200.so")
        _cuda.bdl_last_error.restype = ctypes.c_char_p
    return _cuda


def host_lib():
    global _host
    if _host is None:
        _host = _load("libbigdl_b200_host.so")
    return _host


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = _cuda.bdl_last_error().decode() if _cuda is not None else "?"
        raise NativeError(f"{what} failed ({rc}): {msg}")


def stream_ptr() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


LAUNCHES = 0  # number of native kernel-launching calls made by this process (bench.py reports it)


def count_launch(n: int = 1):
    global LAUNCHES
    LAUNCHES += n
