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


class DnnStorage:
    """Aligned host storage with explicit ``release()`` (S/tensor/DnnStorage.scala): memory comes from the native pool in
    ``csrc/host/aligned_pool.cpp`` (page-aligned blocks, size-class reuse, a live-pointer table that turns a double
    release into an error and lets tests assert that nothing leaked). Only what the reference's DnnStorage supports is
    offered: ``copy`` from another storage / tensor, ``release``, the address; everything else raises like there."""

    CACHE_LINE_SIZE = 64

    def __init__(self, size: int, dtype=torch.float32):
        import ctypes

        from ..ops import _lib

        lib = _lib.host_lib()
        lib.bdlh_aligned_alloc.restype = ctypes.c_void_p
        lib.bdlh_aligned_alloc.argtypes = [ctypes.c_size_t, ctypes.c_size_t]
        lib.bdlh_aligned_free.argtypes = [ctypes.c_void_p]
        lib.bdlh_aligned_free.restype = ctypes.c_int
        self._lib, self._n, self.dtype = lib, int(size), dtype
        nbytes = max(self._n * torch.empty(0, dtype=dtype).element_size(), 1)
        self.ptr = lib.bdlh_aligned_alloc(nbytes, DnnStorage.CACHE_LINE_SIZE)
        if not self.ptr:
            raise MemoryError(f"aligned allocation of {nbytes} bytes failed")
        buf = (ctypes.c_uint8 * nbytes).from_address(self.ptr)
        self._t = torch.frombuffer(buf, dtype=dtype, count=self._n) if self._n else torch.empty(0, dtype=dtype)
        self._released = False

    def length(self) -> int:
        return self._n

    def isReleased(self) -> bool:
        return self._released

    def tensor(self) -> torch.Tensor:
        if self._released:
            raise RuntimeError("DnnStorage used after release()")
        return self._t

    def copy(self, source, offset: int = 0, sourceOffset: int = 0, length: int | None = None):
        src = source.tensor() if isinstance(source, DnnStorage) else (source.t if isinstance(source, Tensor) else source)
        src = torch.as_tensor(src).reshape(-1)
        n = (src.numel() - sourceOffset) if length is None else length
        self.tensor()[offset:offset + n].copy_(src[sourceOffset:sourceOffset + n])
        return self

    def release(self) -> None:
        if self._released:
            return
        self._t = None
        if self._lib.bdlh_aligned_free(self.ptr) != 0:
            raise RuntimeError("DnnStorage: pointer is not live in the native pool (double release?)")
        self._released = True

    def __del__(self):
        try:
            if not self._released:
                self.release()
        except Exception:
            pass

    def fill(self, *a):
        raise NotImplementedError("Not support this operation in DnnStorage")

    resize = set = array = apply = update = fill

    @staticmethod
    def stats() -> dict:
        import ctypes

        from ..ops import _lib

        lib = _lib.host_lib()
        lib.bdlh_aligned_stat.restype = ctypes.c_size_t
        lib.bdlh_aligned_stat.argtypes = [ctypes.c_int]
        names = ("live_blocks", "live_bytes", "cached_bytes", "allocations", "reused")
        return {n: int(lib.bdlh_aligned_stat(i)) for i, n in enumerate(names)}


class DnnTensor:
    CACHE_LINE = 256      # bytes; the reference's ``bigdl.cache.line`` is 64 on CPUs

    def __init__(self, *sizes, dtype=torch.float32, device=None):
        if len(sizes) == 1 and isinstance(sizes[0], (list, tuple)):
            sizes = tuple(sizes[0])
        dev = torch.device(device) if device is not None else torch.device(
            "cuda" if torch.cuda.is_available() else "cpu")
        shape = tuple(int(s) for s in sizes)
        self._storage = None
        if dev.type == "cpu":
            n = 1
            for d in shape:
                n *= d
            self._storage = DnnStorage(n, dtype)             # aligned native block, returned to the pool on release()
            self._t = self._storage.tensor().view(shape).zero_()
        else:
            self._t = torch.zeros(shape, dtype=dtype, device=dev)
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
            if self._storage is not None:
                self._storage.release()
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
