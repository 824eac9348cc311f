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


P2P_ALLREDUCE_CALLS = 0          # buckets reduced by the peer-memory kernel in this process (diagnostics)


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
    def __init__(self, world: int | None = None, compress: str = "auto", group=None, backend: str | None = None):
        self.world = world if world is not None else (dist.get_world_size() if dist.is_initialized() else 1)
        self.group = group
        self.compress = compress
        # "p2p": one-shot all-reduce over NVLink peer memory (csrc/comm_fused.cu, reduce + write-back mode of the fused
        # kernel, flag barriers in peer memory); "nccl": dist.all_reduce on the communication stream (the baseline)
        from ..utils import config

        self.backend = backend or config.get("bigdl.parameter.backend")
        self._p2p = None
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
                if self._nvls_ready(b):
                    self._nvls_allreduce(b)
                elif self._p2p_ready(b):
                    self._p2p_allreduce(b)
                elif self._use_bf16(b):
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

    # ------------------------------------------------------------------ NVLink peer-memory data path
    def _p2p_ready(self, b: _Bucket) -> bool:
        if self.backend not in ("auto", "p2p") or self.compress == "none" or not dist.is_initialized():
            return False
        if self._p2p is None:
            try:
                self._p2p = self._p2p_init(b.grads.device)
            except Exception as e:            # no peer access between the GPUs of this job: fall back to NCCL
                if self.backend == "p2p":
                    raise
                import logging

                logging.getLogger("bigdl_b200").warning("peer-memory all-reduce unavailable (%s); using NCCL", e)
                self._p2p = False
        return bool(self._p2p) and b.offset % 8 == 0 and b.length % 8 == 0

    def _p2p_init(self, device):
        from .symmetric import SymmetricBuffer

        P = max(b.offset + b.length for b in self.buckets.values())
        off_flags = (P * 2 + 255) // 256 * 256
        sym = SymmetricBuffer(off_flags + 256, self.group, device)
        return {"sym": sym, "g16": sym.view(torch.int16, 0, P), "g16_ptrs": sym.peer_ptr_array(0),
                "flag_ptrs": sym.peer_ptr_array(off_flags),
                "local_sync": torch.zeros(2, dtype=torch.int32, device=device),
                "epoch_dev": torch.zeros(1, dtype=torch.int64, device=device),
                "rank": dist.get_rank(self.group)}

    # ------------------------------------------------------------------ in-switch (NVLS) data path
    def _nvls_ready(self, b: _Bucket) -> bool:
        from ..utils import config

        if not config.get("bigdl.parameter.nvls") or self.compress == "none" or not dist.is_initialized():
            return False
        if getattr(self, "_nvls", None) is None:
            from . import nvls

            self._nvls = False
            if nvls.multicast_supported(b.grads.device):
                try:
                    P = max(bk.offset + bk.length for bk in self.buckets.values())
                    self._nvls = nvls.NvlsBuffer((P + 7) // 8 * 8, torch.bfloat16, self.group, b.grads.device)
                except Exception as e:
                    import logging

                    logging.getLogger("bigdl_b200").warning("NVLS all-reduce unavailable (%s); using peer loads", e)
        return bool(self._nvls) and b.offset % 8 == 0 and b.length % 8 == 0

    def _nvls_allreduce(self, b: _Bucket):
        """grads[bucket] = mean over ranks with the sum computed inside the NVSwitch (csrc/nvls.cu)."""
        from ..ops import _lib

        buf = self._nvls
        lib = _lib.cuda_lib()
        g16 = buf.local[b.offset:b.offset + b.length]
        _lib.check(lib.bdl_compress_grads(b.grads.data_ptr(), g16.data_ptr(), b.length, 0, 0, _lib.stream_ptr()),
                   "bdl_compress_grads")
        _lib.count_launch()
        buf.reduce_to_f32(b.grads.view(-1)[:b.length], offset=b.offset, average=True)
        b.stats["nvls"] = b.stats.get("nvls", 0) + 1

    def _p2p_allreduce(self, b: _Bucket):
        """grads[bucket] = mean over ranks, one kernel: every rank loads the bucket's bf16 slice of every peer over
        NVLink, sums in fp32 in rank order and writes the average back (the one-shot form of S/utils/
        DistriParameterSynchronizer.scala:234-356's fetch → aggregate → put)."""
        import ctypes

        from ..ops import _lib
        from ..optim.optim_method import pack_hp
        from .all_reduce_parameter import CM_BARRIER_BEGIN, CM_BARRIER_END, CM_REDUCE, CM_WRITE_GRAD

        p = self._p2p
        lib = _lib.cuda_lib()
        st = _lib.stream_ptr()
        g16 = p["g16"][b.offset:b.offset + b.length]
        _lib.check(lib.bdl_compress_grads(b.grads.data_ptr(), g16.data_ptr(), b.length, 0, 0, st), "bdl_compress_grads")
        hp = pack_hp(gscale=1.0 / self.world)
        buf = ctypes.create_string_buffer(hp, len(hp))
        mode = CM_REDUCE | CM_WRITE_GRAD | CM_BARRIER_BEGIN | CM_BARRIER_END
        _lib.check(lib.bdl_fused_rs_update(p["g16_ptrs"], None, p["flag_ptrs"], self.world, p["rank"], b.offset,
                                           b.length, None, b.grads.data_ptr(), None, None, 0,
                                           ctypes.cast(buf, ctypes.c_void_p), None, mode, 1, p["local_sync"].data_ptr(),
                                           None, 32, p["epoch_dev"].data_ptr(), st), "bdl_fused_rs_update")
        _lib.check(lib.bdl_counter_add(p["epoch_dev"].data_ptr(), 1, st), "bdl_counter_add")
        _lib.count_launch(3)
        b.stats["p2p"] = b.stats.get("p2p", 0) + 1
        global P2P_ALLREDUCE_CALLS
        P2P_ALLREDUCE_CALLS += 1

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


# ------------------------------------------------------------------------------------------------ reference names
SyncMeta = _Bucket          # the per-layer record of an outstanding synchronisation (DistriParameterSynchronizer.scala:464)


class SyncState:
    """The four stages of one layer's synchronisation in the reference's BlockManager protocol
    (DistriParameterSynchronizer.scala:471-477). Here a bucket moves INIT → launched (``put``: the collective is on
    the communication stream) → ready (``get`` waited on its event); the names map onto those stages."""
    INIT, FETCH_PARTITION, AGGREGATION, PUT_AGGREGATED = "0", "1", "2", "3"

    @staticmethod
    def of(bucket) -> str:
        if bucket.ready:
            return SyncState.PUT_AGGREGATED
        if bucket.launched:
            return SyncState.AGGREGATION
        return SyncState.INIT


class BlockManagerParameterSynchronizer(DistriParameterSynchronizer):
    """``new BlockManagerParameterSynchronizer(partitionID, totalPartition)`` of the reference: the same
    init / put / get / clear protocol. The Spark BlockManager transport (fetch the peers' compressed slices, sum, put
    back) is replaced by one all-reduce per bucket over NVLink peer memory, NVLS or NCCL."""

    def __init__(self, partitionID: int | None = None, totalPartition: int | None = None, **kw):
        super().__init__(world=totalPartition, **kw)
        self.partitionID = partitionID if partitionID is not None else (dist.get_rank() if dist.is_initialized() else 0)
        self.totalPartition = self.world
