"""In-switch (NVLS / NVLink SHARP) all-reduce over a multicast-mapped symmetric buffer.

The symmetric allocation and its multicast mapping come from ``torch.distributed._symmetric_memory`` (cuMemCreate +
cuMulticastCreate / BindMem across the ranks of a process group — the VMM plumbing the reference's
``BlockManagerWrapper`` has no analogue of); the reduction itself is this repo's kernels (csrc/nvls.cu):
``multimem.ld_reduce`` makes the NVSwitch add the ranks' values and hand back one result, so a rank receives the bucket
once instead of once per peer. Used by ``DistriParameterSynchronizer`` (ParallelOptimizer's per-bucket all-reduce)
when ``bigdl.parameter.nvls`` is on and the GPUs support multicast; the peer-load kernel stays the fallback.
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from ..ops import _lib

_P, _I, _L, _F = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
_lib.declare("bdl_nvls_reduce_bf16_to_f32", [_P, _P, _L, _F, _P])
_lib.declare("bdl_nvls_allreduce_bf16_inplace", [_P, _L, _I, _I, _P])
_lib.declare("bdl_nvls_allreduce_f32", [_P, _P, _L, _F, _P])


def multicast_supported(device=None) -> bool:
    """True when this GPU can join a multicast object (NVSwitch fabric, driver support)."""
    if not torch.cuda.is_available():
        return False
    try:
        from torch._C._distributed_c10d import _SymmetricMemory
        from torch._C._autograd import DeviceType

        idx = torch.cuda.current_device() if device is None else torch.device(device).index
        return bool(_SymmetricMemory.has_multicast_support(DeviceType.CUDA, idx))
    except Exception:
        return False


class NvlsBuffer:
    """``numel`` elements of ``dtype`` at the same offset on every rank of ``group`` plus the multicast address of the
    range. ``local`` is this rank's tensor (write your contribution there); the all-reduce methods bracket the kernel
    with group barriers on the current stream (nobody reads before everybody wrote; nobody overwrites before everybody
    read)."""

    def __init__(self, numel: int, dtype=torch.bfloat16, group=None, device=None):
        import torch.distributed._symmetric_memory as symm_mem

        self.group = group if group is not None else dist.group.WORLD
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        try:
            symm_mem.enable_symm_mem_for_group(self.group.group_name)
        except Exception:
            pass
        self.local = symm_mem.empty(numel, dtype=dtype, device=device)
        self.handle = symm_mem.rendezvous(self.local, self.group)
        self.mc_ptr = int(self.handle.multicast_ptr or 0)
        if not self.mc_ptr:
            raise RuntimeError("the symmetric allocation has no multicast mapping (NVLS unavailable for this group)")
        self.rank, self.world = self.handle.rank, self.handle.world_size
        self.local.zero_()
        self._chan = 0

    def _barrier(self):
        self.handle.barrier(channel=self._chan)
        self._chan ^= 1

    def _mc(self, offset: int) -> int:
        return self.mc_ptr + offset * self.local.element_size()

    def reduce_to_f32(self, out: torch.Tensor, offset: int = 0, average: bool = True) -> torch.Tensor:
        """out[i] = (mean | sum) over ranks of local[offset + i] — one-shot: each rank pulls the reduced range itself."""
        lib = _lib.cuda_lib()
        n = out.numel()
        scale = 1.0 / self.world if average else 1.0
        self._barrier()
        if self.local.dtype == torch.bfloat16:
            _lib.check(lib.bdl_nvls_reduce_bf16_to_f32(self._mc(offset), out.data_ptr(), n, scale, _lib.stream_ptr()),
                       "bdl_nvls_reduce_bf16_to_f32")
        else:
            _lib.check(lib.bdl_nvls_allreduce_f32(self._mc(offset), out.data_ptr(), n, scale, _lib.stream_ptr()),
                       "bdl_nvls_allreduce_f32")
        _lib.count_launch()
        self._barrier()
        return out

    def all_reduce_inplace(self, offset: int = 0, numel: int | None = None) -> torch.Tensor:
        """local[offset : offset + numel] = sum over ranks, on every rank (two-shot: rank r reduces slice r in the switch
        and multicasts it back). bf16 only."""
        lib = _lib.cuda_lib()
        n = self.local.numel() - offset if numel is None else numel
        self._barrier()
        _lib.check(lib.bdl_nvls_allreduce_bf16_inplace(self._mc(offset), n, self.rank, self.world, _lib.stream_ptr()),
                   "bdl_nvls_allreduce_bf16_inplace")
        _lib.count_launch()
        self._barrier()
        return self.local[offset:offset + n]
