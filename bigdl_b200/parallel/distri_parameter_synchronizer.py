"""Per-bucket asynchronous all-reduce overlapped with backward — the B200 form of the reference's
``BlockManagerParameterSynchronizer`` (S/utils/DistriParameterSynchronizer.scala:60-477: ``init/put/get/clear``, a
state machine FETCH→AGGREGATE→PUT per layer bucket driven by thread pools and two priority queues).

Here a bucket is a contiguous range of the flat gradient vector. ``put(name)`` is called from inside ``backward`` the
moment the bucket's gradients are final; it launches the all-reduce on a dedicated communication stream (NCCL, bf16 or
fp32 payload) or as an asynchronous gloo collective (CPU), so the transfer overlaps the rest of backward. ``get(name)``
is called lazily — right before the next forward needs that bucket's weights — and only then waits.

Priorities: the reference serves high-priority buckets (the FIRST layers, needed earliest by the next forward) before
others because its transfers are point-to-point fetches. Collectives must be issued in the same order on every rank,
so the launch order here is the deterministic ``put`` order; priority decides the order in which ``flush`` launches
buckets that were not launched from a hook, and the order ``get_all`` drains them.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch
import torch.distributed as dist


@dataclass
class _Bucket:
    name: str
    offset: int
    length: int
    priority: int
    weights: torch.Tensor
    grads: torch.Tensor
    work: object = None            # outstanding collective
    event: object = None           # CUDA event recorded on the comm stream after the collective
    payload: torch.Tensor | None = None
    launched: bool = False
    ready: bool = False
    stats: dict = field(default_factory=lambda: {"puts": 0, "gets": 0})


class DistriParameterSynchronizer:
    def __init__(self, world: int | None = None, compress: str = "auto", group=None):
        self.world = world if world is not None else (dist.get_world_size() if dist.is_initialized() else 1)
        self.group = group
        self.compress = compress
        self.buckets: dict[str, _Bucket] = {}
        self._order: list[str] = []
        self._comm_stream = None
        self.launch_log: list[str] = []        # names in launch order (tests / tracing)

    # ------------------------------------------------------------------ reference API: init / put / get / clear
    def init(self, name: str, global_size: int, priority: int = 1, weights: torch.Tensor | None = None,
             grads: torch.Tensor | None = None, offset: int = 0):
        assert weights is not None and grads is not None and weights.numel() == global_size == grads.numel()
        self.buckets[name] = _Bucket(name, offset, global_size, priority, weights, grads)
        self._order.append(name)
        if grads.is_cuda and self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream(device=grads.device)
        return self

    def _use_bf16(self, b: _Bucket) -> bool:
        if self.compress == "none":
            return False
        return b.grads.is_cuda        # gloo has no bf16 sum on every build; the CPU wire format stays fp32

    def put(self, name: str):
        """Gradients of ``name`` are final: start averaging them across ranks."""
        b = self.buckets[name]
        if b.launched:
            return
        b.launched, b.ready = True, False
        b.stats["puts"] += 1
        self.launch_log.append(name)
        if self.world == 1:
            b.ready = True
            return
        if b.grads.is_cuda:
            cs = self._comm_stream
            cs.wait_stream(torch.cuda.current_stream(b.grads.device))     # grads were written on the compute stream
            with torch.cuda.stream(cs):
                if self._use_bf16(b):
                    b.payload = b.grads.to(torch.bfloat16)
                    dist.all_reduce(b.payload, group=self.group)
                    b.grads.copy_(b.payload)
                    b.grads.div_(self.world)
                else:
                    dist.all_reduce(b.grads, group=self.group)
                    b.grads.div_(self.world)
                b.event = torch.cuda.Event()
                b.event.record(cs)
        else:
            b.work = dist.all_reduce(b.grads, group=self.group, async_op=True)

    def get(self, name: str):
        """Wait until ``name``'s averaged gradients are available; returns ``(weights, grads)`` views, or ``None``
        when the bucket has nothing outstanding."""
        b = self.buckets[name]
        if not b.launched:
            return None
        b.stats["gets"] += 1
        if self.world > 1:
            if b.event is not None:
                torch.cuda.current_stream(b.grads.device).wait_event(b.event)      # device-side wait, host does not block
                b.event, b.payload = None, None
            elif b.work is not None:
                b.work.wait()
                b.work = None
                b.grads.div_(self.world)
        b.launched, b.ready = False, True
        return b.weights, b.grads

    def flush(self):
        """Launch every bucket that has not been launched yet, highest priority first."""
        for name in sorted(self._order, key=lambda n: -self.buckets[n].priority):
            self.put(name)

    def pending(self) -> list[str]:
        return [n for n in self._order if self.buckets[n].launched]

    def clear(self):
        for b in self.buckets.values():
            if b.work is not None:
                b.work.wait()
            b.work = b.event = b.payload = None
            b.launched = False
        return self
