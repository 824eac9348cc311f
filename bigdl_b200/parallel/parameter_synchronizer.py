"""Small keyed exchange used by layers that need a cross-replica reduction inside forward/backward — sync-BatchNorm in
the reference (S/utils/ParameterSynchronizer.scala:29-106: a CyclicBarrier per key; every replica thread ``syncData``s
its tensor and ``collect``s the element-wise sum; hooked from S/nn/BatchNormalization.scala:65-76,224-235).

With one replica per GPU the replicas are ranks, so ``collect`` is an all-reduce on the replica group. The thread
variant is kept for CPU multi-replica runs (LocalOptimizer-style tests)."""
from __future__ import annotations

import threading

import torch
import torch.distributed as dist


class Event:
    """One rendezvous point: ``parties`` replicas deposit a tensor, meet at the barrier, read the sum
    (ParameterSynchronizer.scala:108-150)."""

    def __init__(self, parties: int):
        self.parties = parties
        self.barrier = threading.Barrier(parties)
        self.data: dict[int, torch.Tensor] = {}
        self.lock = threading.Lock()


class ParameterSynchronizer:
    _events: dict[str, Event] = {}
    _lock = threading.Lock()

    @classmethod
    def register(cls, key: str, parties: int):
        with cls._lock:
            if key not in cls._events:
                cls._events[key] = Event(parties)

    @classmethod
    def reset(cls, key: str | None = None):
        with cls._lock:
            if key is None:
                cls._events.clear()
            else:
                cls._events.pop(key, None)

    @classmethod
    def syncData(cls, key: str, dt: torch.Tensor, replica_id: int | None = None):
        ev = cls._events[key]
        rid = threading.get_ident() if replica_id is None else replica_id
        with ev.lock:
            ev.data[rid] = dt
        ev.barrier.wait()

    @classmethod
    def collect(cls, key: str) -> torch.Tensor:
        ev = cls._events[key]
        with ev.lock:
            out = torch.stack(list(ev.data.values())).sum(0)
        ev.barrier.wait()          # nobody overwrites its slot before everyone has read
        return out

    # ---- rank form
    @staticmethod
    def all_reduce(t: torch.Tensor, group=None) -> torch.Tensor:
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(t, group=group)
        return t


def enable_sync_batchnorm(model, group=None):
    """Turn every BatchNormalization in ``model`` into sync-BN over ``group`` (default: all ranks) — the counterpart of
    ``setParallism`` + ParameterSynchronizer in S/models/resnet/TrainImageNet.scala:143-151."""
    from ..nn.layers_core import BatchNormalization

    n = 0
    for m in model.all_modules():
        if isinstance(m, BatchNormalization):
            m._sync = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
            n += 1
    return n


_Event = Event
