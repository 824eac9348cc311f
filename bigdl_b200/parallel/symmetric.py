"""Symmetric peer-mapped device buffers: every rank allocates the same-sized buffer and maps every peer's copy.

This is the data plane that replaces the reference's Spark BlockManager (S/parameters/AllReduceParameter.scala uses
``BlockManagerWrapper.putBytes/getLocalOrRemoteBytes``; spark/spark-version/*/BlockManagerWrapper.scala): a "block id"
becomes (peer, offset) inside a symmetric buffer and fetching a remote block becomes an NVLink load issued from
inside a kernel.

Allocation is plain ``cudaMalloc`` + CUDA IPC handles exchanged once over ``torch.distributed`` (NCCL/gloo are only
the control plane). Buffers are exposed to PyTorch zero-copy through ``__cuda_array_interface__`` so that, e.g.,
the model's bf16 compute weights *are* the buffer peers push into.
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from ..ops import _lib

_TYPESTR = {torch.uint8: "|u1", torch.int16: "<i2", torch.int32: "<i4", torch.float32: "<f4", torch.bfloat16: "<i2"}


class _RawCuda:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def tensor_from_ptr(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    """uint8 tensor aliasing raw device memory (no ownership)."""
    return torch.as_tensor(_RawCuda(ptr, nbytes), device=device)


class SymmetricBuffer:
    """``nbytes`` of device memory on every rank of ``group``; ``ptrs[r]`` is rank r's copy as seen from this process."""

    def __init__(self, nbytes: int, group=None, device: torch.device | None = None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.nbytes = (nbytes + 255) // 256 * 256
        lib = _lib.cuda_lib()
        local = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        with torch.cuda.device(self.device):
            _lib.check(lib.bdl_ipc_alloc(self.nbytes, ctypes.byref(local), handle), "bdl_ipc_alloc")
        self.local_ptr = local.value
        self.ptrs = [0] * self.world
        self.ptrs[self.rank] = self.local_ptr
        self._opened = []
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle.raw), group=group)
            for r, h in enumerate(handles):
                if r == self.rank:
                    continue
                p = ctypes.c_void_p()
                with torch.cuda.device(self.device):
                    _lib.check(lib.bdl_ipc_open(ctypes.create_string_buffer(h, 64), ctypes.byref(p)), "bdl_ipc_open")
                self.ptrs[r] = p.value
                self._opened.append(p.value)
            dist.barrier(group=group)
        self.local = tensor_from_ptr(self.local_ptr, self.nbytes, self.device)

    def view(self, dtype: torch.dtype, offset_bytes: int = 0, numel: int | None = None) -> torch.Tensor:
        esz = torch.empty(0, dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - offset_bytes) // esz
        return self.local[offset_bytes:offset_bytes + numel * esz].view(dtype)

    def peer_ptr_array(self, offset_bytes: int = 0):
        arr = (ctypes.c_void_p * self.world)()
        for r in range(self.world):
            arr[r] = self.ptrs[r] + offset_bytes
        return arr

    def close(self, collective: bool = True):
        """Unmap the peers' copies and free the local one. ``collective=False`` (finalisers) skips the barrier that
        normally keeps a rank from freeing memory a slower peer still has mapped."""
        lib = _lib.cuda_lib()
        if lib is None or self.local_ptr is None:
            return
        torch.cuda.synchronize(self.device)
        for p in self._opened:
            lib.bdl_ipc_close(ctypes.c_void_p(p))
        self._opened = []
        if collective and self.world > 1 and dist.is_initialized():
            try:
                dist.barrier(group=self.group)
            except Exception:
                pass
        lib.bdl_ipc_free(ctypes.c_void_p(self.local_ptr))
        self.local_ptr = None


class LocalPeerGroup:
    """N *virtual* ranks whose buffers all live on one GPU — the single-device stand-in used by tests (launch order
    replaces the cross-GPU flag barriers; see csrc/comm_fused.cu header)."""

    def __init__(self, world: int, nbytes: int, device: torch.device):
        self.world = world
        self.device = device
        self.nbytes = (nbytes + 255) // 256 * 256
        self.buffers = [torch.zeros(self.nbytes, dtype=torch.uint8, device=device) for _ in range(world)]
        self.ptrs = [b.data_ptr() for b in self.buffers]

    def view(self, rank: int, dtype: torch.dtype, offset_bytes: int = 0, numel: int | None = None):
        esz = torch.empty(0, dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - offset_bytes) // esz
        return self.buffers[rank][offset_bytes:offset_bytes + numel * esz].view(dtype)

    def peer_ptr_array(self, offset_bytes: int = 0):
        arr = (ctypes.c_void_p * self.world)()
        for r in range(self.world):
            arr[r] = self.ptrs[r] + offset_bytes
        return arr
