"""DnnTensor: a device-resident, explicitly released buffer (reference: S/tensor/DnnTensor.scala, DnnStorage.scala — 64-byte
aligned off-heap memory for MKL-DNN with ``release()`` and a leak-tracking pool; only copy/zero/add/axpby are defined
on it, everything else must go through a dense Tensor).

Here the "native side" is the GPU: the buffer lives in HBM (the caching allocator's blocks are 512-byte aligned, which
covers the 128-byte TMA and 256-byte vector requirements of the kernels), ``release()`` returns it immediately instead
of waiting for garbage collection, and the same small op set is offered."""
from __future__ import annotations

import weakref

import torch

from .tensor import Tensor

_LIVE: "weakref.WeakSet[DnnTensor]" = weakref.WeakSet()


class DnnTensor:
    CACHE_LINE = 256      # bytes; the reference's ``bigdl.cache.line`` is 64 on CPUs

    def __init__(self, *sizes, dtype=torch.float32, device=None):
        if len(sizes) == 1 and isinstance(sizes[0], (list, tuple)):
            sizes = tuple(sizes[0])
        dev = torch.device(device) if device is not None else torch.device(
            "cuda" if torch.cuda.is_available() else "cpu")
        self._t = torch.zeros(tuple(int(s) for s in sizes), dtype=dtype, device=dev)
        self._released = False
        _LIVE.add(self)

    # ---- the reference's op set
    def _check(self):
        if self._released:
            raise RuntimeError("DnnTensor used after release()")
        return self._t

    def size(self, dim: int | None = None):
        s = list(self._check().shape)
        return s if dim is None else s[dim - 1]

    def nElement(self) -> int:
        return self._check().numel()

    def storageAddress(self) -> int:
        return self._check().data_ptr()

    def isReleased(self) -> bool:
        return self._released

    def release(self) -> None:
        """Free the device memory now (idempotent)."""
        if not self._released:
            self._t = None
            self._released = True

    def zero(self):
        self._check().zero_()
        return self

    def copy(self, other):
        src = other._check() if isinstance(other, DnnTensor) else (other.t if isinstance(other, Tensor) else other)
        self._check().copy_(src.reshape(self._t.shape))
        return self

    def add(self, other: "DnnTensor"):
        self._check().add_(other._check())
        return self

    def axpby(self, a: float, b: float, other: "DnnTensor"):
        """this = a * other + b * this"""
        self._check().mul_(b).add_(other._check(), alpha=a)
        return self

    def dense(self) -> Tensor:
        """Copy out into an ordinary Tensor (the reference copies DnnTensor → DenseTensor the same way)."""
        return Tensor(self._check().clone())

    def torch(self) -> torch.Tensor:
        return self._check()

    @staticmethod
    def liveCount() -> int:
        """Unreleased DnnTensors (the reference tracks them to report native leaks)."""
        return sum(1 for t in _LIVE if not t._released)
