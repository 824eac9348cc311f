"""AllReduceParameter: reduce-scatter of compressed gradients → sharded OptimMethod update → all-gather of compressed
weights (ZeRO-1-like), the core of synchronous data-parallel training.

Reference: S/parameters/AllReduceParameter.scala (init :155-175, getWeights :201-228, aggregateGradientPartition
:236-278, putGradients :287-309, sendWeightPartition :315-328), S/parameters/FP16CompressedTensor.scala,
S/optim/DistriOptimizer.scala:204-396 for how the pieces are sequenced.

Backends (``bigdl.parameter.backend``):
  * ``p2p``   — the product: one fused sm_100a kernel per step over NVLink peer memory (csrc/comm_fused.cu).
  * ``nccl``  — baseline: bf16 ``reduce_scatter_tensor`` + fused local update + ``all_gather_into_tensor``.
  * ``gloo``  — CPU processes (tests / plumbing); implements the reference arithmetic literally, including the
                truncate-after-every-add parity mode that tests compare bit-for-bit with a straight-line oracle.
  * world size 1 — no communication, just the flat fused update.

Compression (``bigdl.parameter.compress``): ``bf16_trunc`` is the reference's "fp16" codec (keep the top 16 bits of the
fp32 word), ``bf16_rn`` rounds to nearest even (default on GPU), ``fp32`` sends uncompressed (gloo/nccl only).
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist

from ..nn.flat import FlatParams, partition_range
from ..ops import elementwise as ew
from ..utils import config

CM_REDUCE, CM_WRITE_GRAD, CM_UPDATE, CM_PUBLISH, CM_BARRIER_BEGIN, CM_BARRIER_END, CM_PARITY, CM_SUMSQ, CM_DIVIDE = (
    1, 2, 4, 8, 16, 32, 64, 128, 256)


def aligned_partition(total: int, parts: int, rank: int, align: int = 8) -> tuple[int, int]:
    """Partition [0,total) into ``parts`` contiguous slices whose boundaries are multiples of ``align`` — the
    reference's rule (size/N, first size%N slices get +1) applied to units of ``align`` elements."""
    if total % align:
        raise ValueError("total must be a multiple of align")
    s, l = partition_range(total // align, parts, rank)
    return s * align, l * align


def _trunc_to_bf16_bits(x: torch.Tensor) -> torch.Tensor:
    return (x.contiguous().view(torch.int32) >> 16).to(torch.int16)


def _bits_to_f32(b: torch.Tensor) -> torch.Tensor:
    return (b.to(torch.int32) << 16).view(torch.float32)


class AllReduceParameter:
    def __init__(self, flat: FlatParams, group=None, backend: str | None = None, compress: str | None = None,
                 align: int = 8, root=None, method_ranges=None):
        """``root``: the model whose parameters ``flat`` holds — lets the p2p backend cut gradient buckets at its
        top-level layers and overlap their exchange with backward. ``method_ranges``: ``[(offset, length)]`` of the
        per-submodule OptimMethods (sorted, tiling the vector); bucket borders then respect them."""
        self.flat = flat
        self.root, self.method_ranges = root, method_ranges
        self._plane = None
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = flat.weights.device
        self.P = flat.numel
        self.align = align if self.P % align == 0 else 1
        self.start, self.length = aligned_partition(self.P, self.world, self.rank, self.align)
        backend = backend or config.get("bigdl.parameter.backend")
        if backend == "auto":
            if self.world == 1:
                backend = "local"
            elif self.device.type == "cuda":
                backend = "p2p"
            else:
                backend = "gloo"
        self.backend = backend
        compress = compress or config.get("bigdl.parameter.compress")
        self.compress = compress
        self.truncate = compress == "bf16_trunc"
        self.epoch = 0
        self.weight_partition: torch.Tensor | None = None   # fp32 master copy of my slice
        self.gradient_partition: torch.Tensor | None = None
        self._sym = None
        self._inited = False

    # ------------------------------------------------------------------ set-up
    def init(self):
        if self._inited:
            return self
        s, l = self.start, self.length
        self.weight_partition = self.flat.weights[s:s + l].clone()
        self.gradient_partition = torch.zeros(l, dtype=torch.float32, device=self.device)
        if self.backend == "p2p" and config.get("bigdl.parameter.p2p") == "bucketed":
            self._init_bucketed()
        elif self.backend == "p2p":
            self._init_p2p()
        elif self.backend == "nccl":
            # equal-sized shards for the NCCL collectives
            self._shard = (self.P + self.world - 1) // self.world
            self._shard = (self._shard + 7) // 8 * 8
            self._g16 = torch.zeros(self._shard * self.world, dtype=torch.bfloat16, device=self.device)
            self._gpart16 = torch.zeros(self._shard, dtype=torch.bfloat16, device=self.device)
            self._w16_all = torch.zeros(self._shard * self.world, dtype=torch.bfloat16, device=self.device)
            self.start = min(self.rank * self._shard, self.P)
            self.length = max(0, min(self._shard, self.P - self.start))
            self.weight_partition = self.flat.weights[self.start:self.start + self.length].clone()
            self.gradient_partition = torch.zeros(self.length, dtype=torch.float32, device=self.device)
        elif self.backend == "gloo" and self.compress != "fp32":
            # every rank starts from the *compressed* initial weights, like the reference's first getWeights()
            with torch.no_grad():
                w = self.flat.weights
                w.copy_(_bits_to_f32(_trunc_to_bf16_bits(w)) if self.truncate else w.to(torch.bfloat16).float())
        if self.flat.compute is not None:
            self.flat.refresh_compute(force=True)
        self._inited = True
        return self

    def _init_p2p(self):
        from .symmetric import SymmetricBuffer

        P = self.P
        self._off_g16 = 0
        self._off_w16 = (P * 2 + 255) // 256 * 256
        self._off_flags = self._off_w16 + (P * 2 + 255) // 256 * 256
        total = self._off_flags + 256
        self._sym = SymmetricBuffer(total, self.group, self.device)
        self.g16 = self._sym.view(torch.int16, self._off_g16, P)
        w16 = self._sym.view(torch.bfloat16, self._off_w16, P)
        self._rebind_compute(w16)
        self._local_sync = torch.zeros(2, dtype=torch.int32, device=self.device)
        self._epoch_dev = torch.zeros(1, dtype=torch.int64, device=self.device)   # barrier generation, device resident
        self._sumsq = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._g16_ptrs = self._sym.peer_ptr_array(self._off_g16)
        self._w16_ptrs = self._sym.peer_ptr_array(self._off_w16)
        self._flag_ptrs = self._sym.peer_ptr_array(self._off_flags)

    def _init_bucketed(self):
        """p2p backend, bucketed form (parallel/bucketed.py): the rank owns one sub-slice of every bucket; its master
        weights / averaged gradients / optimizer state are the concatenation of those sub-slices."""
        from .bucketed import BucketedP2P

        pl = BucketedP2P(self, self.root, self.method_ranges)
        self._plane = pl
        self._sym = pl.sym
        self.g16 = pl.g16
        parts = [self.flat.weights[s:s + l] for s, l, _ in pl.my_slices()]
        self.weight_partition = torch.cat(parts).clone() if parts else torch.zeros(0, device=self.device)
        self.gradient_partition = torch.zeros(pl.part_len, dtype=torch.float32, device=self.device)
        self.start, self.length = -1, pl.part_len          # not one contiguous range any more
        self._rebind_compute(pl.w16)
        self._epoch_dev = pl.epoch_dev

    @property
    def overlapped(self) -> bool:
        """True when the gradient exchange is launched from inside backward (begin_step / end_step protocol)."""
        return self._plane is not None

    def method_target(self, i: int = 0) -> torch.Tensor:
        """The fp32 vector OptimMethod ``i`` updates on this rank."""
        if self._plane is None:
            return self.update_target()
        off, ln = self._plane.method_partition(i) if self.method_ranges else (0, self._plane.part_len)
        return self.weight_partition[off:off + ln]

    def begin_step(self, methods, avg_numbers: int | None = None, l2_clip: float | None = None, hp_dev: int | None = None,
                   plans=None):
        """Call BEFORE forward (bucketed p2p backend). ``methods``: the OptimMethod(s), in ``method_ranges`` order.
        Resolves this step's hyper-parameters, issues the just-in-time weight pulls and arms the backward hooks."""
        if not self._inited:
            self.init()
        avg = float(avg_numbers if avg_numbers is not None else self.world)
        if plans is None:
            plans = []
            for i, om in enumerate(methods):
                om.gradient_scale = avg if self.truncate else 1.0 / avg
                plans.append(om.native_plan(self.method_target(i), None))
        self._plane.begin_step(plans, hp_dev, avg, l2_clip)

    def end_step(self):
        self._plane.end_step()

    def ensure_weights_current(self):
        """Outside a training step (validation, checkpoint, end of training): pull every remote slice now."""
        if self._plane is not None and self._inited:
            self._plane.pull_all()

    def _rebind_compute(self, new_compute: torch.Tensor):
        """Make ``new_compute`` (bf16 [P]) the model's compute-weight buffer and re-point every layer's view."""
        from ..nn.flat import _physical_view

        fl = self.flat
        if fl.compute is not None:
            new_compute.copy_(fl.compute)
        fl.compute = new_compute
        for m, n, off, numel in fl.entries:
            p = m._parameters[n]
            view = _physical_view(new_compute, off, p)
            m._compute[n] = view
        # shared parameters held by other modules
        by_id = {id(m._parameters[n]): m._compute[n] for m, n, _, _ in fl.entries}
        root_modules = {id(m): m for m, _, _, _ in fl.entries}
        for m in root_modules.values():
            for n, p in m._parameters.items():
                if id(p) in by_id:
                    m._compute[n] = by_id[id(p)]
        fl.dirty = True
        fl.refresh_compute(force=True)

    # ------------------------------------------------------------------ the step
    def step(self, optim_method, avg_numbers: int | None = None, l2_clip: float | None = None):
        """One synchronous parameter update: gradients in ``flat.grads`` (local fp32) → updated weights everywhere.
        ``avg_numbers`` is the number of replicas whose gradients are summed (default: world size)."""
        if not self._inited:
            self.init()
        avg = float(avg_numbers if avg_numbers is not None else self.world)
        if self.backend == "local":
            return self._step_local(optim_method, avg, l2_clip)
        if self._plane is not None:
            # gradients are already complete (called after backward): same kernels, nothing left to overlap with
            self.begin_step([optim_method], avg_numbers, l2_clip)
            return self.end_step()
        if self.backend == "p2p":
            return self._step_p2p(optim_method, avg, l2_clip)
        if self.backend == "nccl":
            return self._step_nccl(optim_method, avg, l2_clip)
        return self._step_gloo(optim_method, avg, l2_clip)

    # ---- graph-capturable form: every per-step scalar comes from device memory
    def update_target(self) -> torch.Tensor:
        """The fp32 vector the OptimMethod updates on this rank (whole vector locally, my slice(s) when sharded)."""
        return self.flat.weights if self.backend == "local" else self.weight_partition

    def grad_scale_for(self, world: int) -> float:
        if self.backend == "p2p" and self.truncate:
            return float(world)          # parity mode divides (CM_DIVIDE)
        return 1.0 / world

    def step_device_hp(self, plan, hp_dev: int, epoch_dev: int):
        """``step`` with the OptimHP read from ``hp_dev`` and the barrier generation from ``epoch_dev`` — safe to
        capture in a CUDA graph. ``plan`` = (kind, _, s0, s1) from ``OptimMethod.native_plan``."""
        from ..ops import _lib
        from ..optim.optim_method import native_update

        if not self._inited:
            self.init()
        kind, _, s0, s1 = plan
        fl = self.flat
        if self.backend == "local":
            native_update(kind, fl.weights, fl.grads, s0, s1, None, fl.compute, self.truncate, zero_grad=True,
                          hp_dev=hp_dev)
            fl.dirty = False
            return
        if self.backend != "p2p" or self._plane is not None:
            raise NotImplementedError("graph-captured steps run on the local and p2p parameter backends "
                                      "(bucketed p2p: begin_step / end_step)")
        lib = _lib.cuda_lib()
        st = _lib.stream_ptr()
        _lib.check(lib.bdl_compress_grads(fl.grads.data_ptr(), self.g16.data_ptr(), self.P, int(self.truncate), 1, st),
                   "bdl_compress_grads")
        parity = CM_PARITY | CM_DIVIDE if self.truncate else 0
        mode = CM_REDUCE | CM_UPDATE | CM_PUBLISH | CM_BARRIER_BEGIN | CM_BARRIER_END | parity
        _lib.check(lib.bdl_fused_rs_update(self._g16_ptrs, self._w16_ptrs, self._flag_ptrs, self.world, self.rank,
                                           self.start, self.length, self.weight_partition.data_ptr(),
                                           self.gradient_partition.data_ptr(), _lib.ptr(s0), _lib.ptr(s1), kind, None,
                                           hp_dev, mode, 1, self._local_sync.data_ptr(), self._sumsq.data_ptr(), 0,
                                           self._epoch_dev.data_ptr(), st), "bdl_fused_rs_update")
        _lib.check(lib.bdl_counter_add(self._epoch_dev.data_ptr(), 1, st), "bdl_counter_add")
        _lib.count_launch(3)
        fl.dirty = False

    # ---- world size 1
    def _step_local(self, om, avg, l2_clip):
        fl = self.flat
        om.gradient_scale = 1.0 / avg
        if l2_clip is not None:
            om.gradient_scale *= self._clip_factor(ew.sumsq(fl.grads), avg, l2_clip)
        om.update(fl.weights, fl.grads, w16=fl.compute, truncate=self.truncate)
        fl.grads.zero_()
        fl.dirty = False if fl.compute is not None else fl.dirty

    @staticmethod
    def _clip_factor(sumsq: torch.Tensor, avg: float, max_norm: float) -> float:
        nrm = float(sumsq.sqrt()) / avg
        return max_norm / nrm if nrm > max_norm else 1.0

    # ---- fused NVLink kernel
    def _step_p2p(self, om, avg, l2_clip):
        from ..ops import _lib
        from ..optim.optim_method import pack_hp

        lib = _lib.cuda_lib()
        fl = self.flat
        st = _lib.stream_ptr()
        _lib.check(lib.bdl_compress_grads(fl.grads.data_ptr(), self.g16.data_ptr(), self.P, int(self.truncate), 1, st),
                   "bdl_compress_grads")
        _lib.count_launch()
        parity = CM_PARITY | CM_DIVIDE if self.truncate else 0
        self.epoch += 1
        if l2_clip is None:
            om.gradient_scale = avg if self.truncate else 1.0 / avg
            kind, hp, s0, s1 = om.native_plan(self.weight_partition, None)
            mode = CM_REDUCE | CM_UPDATE | CM_PUBLISH | CM_BARRIER_BEGIN | CM_BARRIER_END | parity
            self._launch(lib, kind, hp, s0, s1, mode, st)
        else:
            # phase 1: reduce + per-slice sum of squares; scalar all-reduce; phase 2: clip-scaled update + publish
            self._sumsq.zero_()
            hp1 = pack_hp(gscale=avg if self.truncate else 1.0 / avg)
            mode1 = CM_REDUCE | CM_WRITE_GRAD | CM_SUMSQ | CM_BARRIER_BEGIN | parity
            self._launch(lib, 0, hp1, None, None, mode1, st)
            dist.all_reduce(self._sumsq, group=self.group)
            nrm = float(self._sumsq.sqrt())
            om.gradient_scale = l2_clip / nrm if nrm > l2_clip else 1.0
            kind, hp, s0, s1 = om.native_plan(self.weight_partition, None)
            self.epoch += 1
            self._launch(lib, kind, hp, s0, s1, CM_UPDATE | CM_PUBLISH | CM_BARRIER_END | (parity & CM_PARITY), st)
        fl.dirty = False

    def _launch(self, lib, kind, hp, s0, s1, mode, st):
        from ..ops import _lib

        buf = ctypes.create_string_buffer(hp, len(hp))
        # generation = 1 + device counter; the counter is bumped after every launch so eager and graph-replayed steps
        # share one monotone sequence on every rank
        _lib.check(lib.bdl_fused_rs_update(self._g16_ptrs, self._w16_ptrs, self._flag_ptrs, self.world, self.rank,
                                           self.start, self.length, self.weight_partition.data_ptr(),
                                           self.gradient_partition.data_ptr(), _lib.ptr(s0), _lib.ptr(s1), kind,
                                           ctypes.cast(buf, ctypes.c_void_p), None, mode, 1,
                                           self._local_sync.data_ptr(), self._sumsq.data_ptr(), 0,
                                           self._epoch_dev.data_ptr(), st), "bdl_fused_rs_update")
        _lib.check(lib.bdl_counter_add(self._epoch_dev.data_ptr(), 1, st), "bdl_counter_add")
        _lib.count_launch(2)

    # ---- NCCL baseline
    def _step_nccl(self, om, avg, l2_clip):
        fl = self.flat
        P = self.P
        ew.cast_f32_to_bf16(fl.grads, self._g16[:P], truncate=self.truncate)
        fl.grads.zero_()
        dist.reduce_scatter_tensor(self._gpart16, self._g16, group=self.group)
        gp = self.gradient_partition
        if self.length:
            ew.cast_bf16_to_f32(self._gpart16[:self.length], gp)
        om.gradient_scale = 1.0 / avg
        if l2_clip is not None:
            ss = ew.sumsq(gp) if self.length else torch.zeros(1, device=self.device)
            dist.all_reduce(ss, group=self.group)
            om.gradient_scale *= self._clip_factor(ss, avg, l2_clip)
        mine = self._w16_all[self.rank * self._shard:(self.rank + 1) * self._shard]
        if self.length:
            om.update(self.weight_partition, gp, w16=mine[:self.length], truncate=self.truncate)
        dist.all_gather_into_tensor(self._w16_all, mine, group=self.group)
        fl.compute.copy_(self._w16_all[:P])
        fl.dirty = False

    # ---- CPU processes (reference arithmetic, literal)
    def _step_gloo(self, om, avg, l2_clip):
        fl = self.flat
        s, l = self.start, self.length
        if self.compress == "fp32":
            g = fl.grads.clone()
            dist.all_reduce(g, group=self.group)
            gp = g[s:s + l] / avg
        else:
            bits = _trunc_to_bf16_bits(fl.grads) if self.truncate else \
                fl.grads.to(torch.bfloat16).view(torch.int16)
            raw = bits.contiguous().view(torch.uint8)          # gloo has no int16: ship the payload as bytes
            gathered_raw = [torch.empty_like(raw) for _ in range(self.world)]
            dist.all_gather(gathered_raw, raw, group=self.group)
            gathered = [t.view(torch.int16) for t in gathered_raw]
            if self.truncate:
                acc = _bits_to_f32(gathered[0][s:s + l])
                for r in range(1, self.world):      # left fold in partition order, re-truncating after every add
                    acc = _bits_to_f32(_trunc_to_bf16_bits(acc + _bits_to_f32(gathered[r][s:s + l])))
            else:
                acc = torch.zeros(l)
                for r in range(self.world):
                    acc = acc + _bits_to_f32(gathered[r][s:s + l])
            gp = acc / avg
        self.gradient_partition = gp
        fl.grads.zero_()
        om.gradient_scale = 1.0
        if l2_clip is not None:
            ss = (gp.double() ** 2).sum().float().reshape(1)
            dist.all_reduce(ss, group=self.group)
            nrm = float(ss.sqrt())
            om.gradient_scale = l2_clip / nrm if nrm > l2_clip else 1.0
        if l:
            om.update(self.weight_partition, gp)
        # publish my slice (compressed), fetch everyone's, expand into the full local weight vector
        if self.compress == "fp32":
            pub = self.weight_partition
        elif self.truncate:
            pub = _trunc_to_bf16_bits(self.weight_partition)
        else:
            pub = self.weight_partition.to(torch.bfloat16).view(torch.int16)
        parts = [None] * self.world
        dist.all_gather_object(parts, pub, group=self.group)
        with torch.no_grad():
            for r, blk in enumerate(parts):
                rs, rl = aligned_partition(self.P, self.world, r, self.align)
                if rl:
                    fl.weights[rs:rs + rl] = blk if self.compress == "fp32" else _bits_to_f32(blk)
        fl.dirty = True

    # ------------------------------------------------------------------ reference-named pieces (API parity)
    def getWeights(self, local_parameter: torch.Tensor | None = None):
        """Make the full weight vector current on this rank (the compute copy on GPU, fp32 on CPU)."""
        if local_parameter is not None and self.flat.compute is not None:
            ew.cast_bf16_to_f32(self.flat.compute, local_parameter)
        return self.flat.compute if self.flat.compute is not None else self.flat.weights

    def gather_master_weights(self) -> torch.Tensor:
        """Collect every rank's fp32 master slice into the full fp32 weight vector (checkpoint / getModel,
        S/optim/DistriOptimizer.scala:658-697)."""
        fl = self.flat
        if self.world == 1 or not self._inited or self.backend == "local":
            return fl.weights
        if fl.weights.is_cuda and dist.get_backend(self.group) == "nccl":
            # one device-side all-gather of the (padded) fp32 partitions; no pickling through the host
            lens = [self._partition_len(r) for r in range(self.world)]
            pad = max(lens)
            mine = torch.zeros(pad, dtype=torch.float32, device=self.device)
            mine[:self.weight_partition.numel()] = self.weight_partition
            every = torch.empty(pad * self.world, dtype=torch.float32, device=self.device)
            dist.all_gather_into_tensor(every, mine, group=self.group)
            with torch.no_grad():
                for r in range(self.world):
                    blk = every[r * pad:r * pad + lens[r]]
                    for s, l, po in self._partition_slices(r):
                        if l:
                            fl.weights[s:s + l] = blk[po:po + l]
            self.ensure_weights_current()
            return fl.weights
        parts = [None] * self.world
        dist.all_gather_object(parts, (self.rank, self.weight_partition.detach().cpu()), group=self.group)
        with torch.no_grad():
            for r, blk in parts:
                for s, l, po in self._partition_slices(r):
                    if l:
                        fl.weights[s:s + l] = blk[po:po + l].to(fl.weights.device)
        return fl.weights

    def _partition_slices(self, rank: int):
        """[(global start, length, offset inside rank's partition)] for ``rank``."""
        if self._plane is not None:
            return self._plane.all_slices(rank)
        if self.backend == "nccl":
            s = min(rank * self._shard, self.P)
            return [(s, max(0, min(self._shard, self.P - s)), 0)]
        s, l = aligned_partition(self.P, self.world, rank, self.align)
        return [(s, l, 0)]

    def _partition_len(self, rank: int) -> int:
        return sum(l for _, l, _ in self._partition_slices(rank))

    def close(self):
        """Free the symmetric buffer. The model's bf16 compute weights live inside it on the p2p backend, so they are
        first re-homed in private memory (every layer's view is re-pointed) — nothing may alias freed device memory."""
        if self._sym is not None:
            fl = self.flat
            self.ensure_weights_current()
            if fl.compute is not None and getattr(self, "g16", None) is not None:
                self._rebind_compute(torch.empty(self.P, dtype=torch.bfloat16, device=self.device))
            self.g16 = None
            if self._plane is not None:
                self._plane.remove_hooks()
                self._plane.sym = None
                self._plane = None
            self._sym.close()
            self._sym = None
            self._inited = False

    def __del__(self):
        try:
            # only un-shared buffers are freed by the finaliser: a peer may still have a shared one mapped, and those
            # are released collectively by close() (the optimizer does that before it builds the next one)
            if self._sym is not None and self._sym.local_ptr is not None and self.world == 1:
                self._sym.close(collective=False)
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass


class FutureResult:
    """Handle on asynchronous transfers (S/parameters/AllReduceParameter.scala:331-333): ``waitResult()`` blocks until
    every one has finished and returns their values. Accepts ``concurrent.futures`` futures, ``torch.distributed``
    work handles (``wait()``), CUDA events (``synchronize()``) or plain values."""

    def __init__(self, futures):
        self.futures = list(futures)

    def waitResult(self):
        out = []
        for f in self.futures:
            if hasattr(f, "result") and callable(f.result):
                out.append(f.result())
            elif hasattr(f, "wait") and callable(f.wait):
                out.append(f.wait())
            elif hasattr(f, "synchronize") and callable(f.synchronize):
                out.append(f.synchronize())
            else:
                out.append(f)
        return out


class Util:
    """S/parameters/Util.scala:23-52."""

    @staticmethod
    def getSumsquareInParallel(parameters: torch.Tensor, parallelism: int = 1) -> float:
        """Σ x² of a flat tensor in fp64. On a GPU this is one launch of the library's reduction (the same one the
        L2-norm gradient clipping uses); on the host torch's own threaded reduction replaces the reference's manual
        split into ``parallelism`` tasks."""
        flat = parameters.reshape(-1)
        if flat.is_cuda and flat.dtype == torch.float32 and flat.is_contiguous():
            from ..ops import elementwise as ew

            return float(ew.sumsq(flat))
        return float(flat.double().pow(2).sum())
