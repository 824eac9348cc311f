"""Bucketed parameter plane of DistriOptimizer on NVLink peer memory.

The reference's iteration is ``getWeights`` (blocking all-gather) → forward/backward → ``putGradients`` →
``aggregateGradientPartition`` → sharded update → ``sendWeightPartition`` (S/parameters/AllReduceParameter.scala:201-328,
sequenced by S/optim/DistriOptimizer.scala:204-396); its overlapped variant launches the gradient exchange per layer
bucket from inside backward (S/optim/ParallelOptimizer.scala:481-513, AbstractModule.asyncGradient). Here both ends of
the iteration are overlapped with compute:

* **reduce-scatter + update, bucket by bucket inside backward.** The flat vector is cut into buckets (contiguous
  ranges in forward order, bucket borders on top-level module borders). Every bucket is partitioned among the ranks;
  the moment backward has produced a bucket's gradients, ONE kernel on the communication stream compresses them to
  bf16, exchanges them over NVLink (P2P loads of this rank's sub-slice from every peer), averages, applies the
  OptimMethod to the fp32 master sub-slice and writes the new bf16 weights into this rank's compute buffer
  (csrc/comm_fused.cu ``bucket_rs_update_kernel``). It then raises a per-bucket "ready" generation in every peer's
  flag area and returns — nobody waits at the end of the step.
* **all-gather, pulled just in time.** At the start of the next forward one small kernel per bucket (forward order,
  communication stream) acquires the owners' ready flags and copies the remote sub-slices into the local compute buffer
  (``pull_gather_kernel``); a layer only waits for its own bucket's event, so late buckets' weights travel while the
  early layers compute. Large ``Linear`` weights skip the pull: their forward GEMM reads the remote row blocks straight
  from the owners' memory with plain (non-multicast) TMA loads and writes them through to the local copy
  (``ops/gemm.py::gemm_bf16_gathered``).

Everything varying per step lives in device memory (generation counter, OptimHP table, clip scale), so the whole
iteration is capturable in one CUDA graph.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, field

import torch
import torch.distributed as dist

from ..nn.flat import partition_range

CM_REDUCE, CM_WRITE_GRAD, CM_UPDATE, CM_PUBLISH, CM_BARRIER_BEGIN, CM_BARRIER_END, CM_PARITY, CM_SUMSQ, CM_DIVIDE = (
    1, 2, 4, 8, 16, 32, 64, 128, 256)
CM_COMPRESS, CM_PUBLISH_LOCAL, CM_SIGNAL_READY, CM_CLIP_DEV = 512, 1024, 2048, 4096


# ====================================================================== layout (pure Python, CPU-testable)
@dataclass
class Bucket:
    index: int
    off: int                      # first element of the bucket in the flat vector
    length: int
    first: object                 # module whose call marks "forward reaches this bucket" / "backward is past it"
    method: int = 0               # index of the OptimMethod that owns the bucket
    slices: list = field(default_factory=list)     # per rank: (global start, length)
    poff: list = field(default_factory=list)       # per rank: offset of its sub-slice inside that rank's partition
    gathered: list = field(default_factory=list)   # (start, length) ranges whose all-gather a fused GEMM performs


def _align_partition(length: int, parts: int, rank: int, align: int = 8):
    s, l = partition_range(length // align, parts, rank)
    return s * align, l * align


def plan_units(root, flat):
    """Bucket borders may only sit where the whole backward of everything after the border has been issued on the main
    stream: the direct children of nested ``Sequential`` containers (a ``Concat`` runs its branches on side streams, so
    it stays one unit). Returns ``[(module, offset, length)]`` tiling ``[0, P)`` in forward order, or a single unit
    when the model's parameters are not laid out in call order (shared parameters, graphs)."""
    from ..nn.module import Sequential

    P = flat.numel
    units = []

    def walk(m):
        if type(m) is Sequential and m.modules:
            for c in m.modules:
                walk(c)
            return
        off, ln = flat.offsets_of(m)
        if ln > 0:
            units.append((m, off, ln))

    walk(root)
    ok = bool(units)
    pos = 0
    for _, off, ln in units:
        if off < pos:
            ok = False
            break
        pos = off + ln
    if not ok:
        return [(root, 0, P)]
    out = []
    for i, (m, off, _) in enumerate(units):
        end = units[i + 1][1] if i + 1 < len(units) else P
        out.append((m, 0 if i == 0 else off, end - (0 if i == 0 else off)))
    return out


def build_buckets(units, world: int, cap: int, first_cap: int | None = None, method_ranges=None, align: int = 8):
    """Group ``units`` into buckets of about ``cap`` elements (the first one ``first_cap``: its exchange cannot overlap
    backward, so it is kept small), never crossing an OptimMethod range border, and partition each among ``world`` ranks.
    ``method_ranges``: ``[(offset, length)]`` sorted, tiling the vector (one entry per OptimMethod) or None."""
    first_cap = cap if first_cap is None else first_cap
    borders = set()
    if method_ranges:
        for off, _ in method_ranges:
            borders.add(off)
    buckets: list[Bucket] = []
    cur = None
    for m, off, ln in units:
        limit = first_cap if not buckets or (cur is buckets[0]) else cap
        if cur is None or cur.length >= limit or off in borders:
            cur = Bucket(len(buckets), off, 0, m)
            buckets.append(cur)
        cur.length = off + ln - cur.off
    ptot = [0] * world
    for b in buckets:
        if b.off % align or b.length % align:
            raise ValueError("bucket borders must be multiples of the alignment")
        if method_ranges:
            for i, (moff, mlen) in enumerate(method_ranges):
                if moff <= b.off < moff + max(mlen, 1):
                    b.method = i
        for r in range(world):
            s, l = _align_partition(b.length, world, r, align)
            b.slices.append((b.off + s, l))
            b.poff.append(ptot[r])
            ptot[r] += l
    return buckets, ptot


# ====================================================================== engine
class BucketedP2P:
    """Owns the symmetric buffers, the bucket layout, the hooks and the launches. Created by ``AllReduceParameter`` for
    the ``p2p`` backend; ``sync`` is that object (it keeps the reference-named surface)."""

    def __init__(self, sync, root=None, method_ranges=None):
        from ..utils import config
        from .symmetric import SymmetricBuffer

        self.sync = sync
        fl = sync.flat
        self.flat = fl
        self.world, self.rank, self.device = sync.world, sync.rank, sync.device
        P = sync.P
        cap = int(float(config.get("bigdl.parameter.bucketMB")) * (1 << 20)) // 4
        first_cap = int(float(config.get("bigdl.parameter.firstBucketMB")) * (1 << 20)) // 4
        units = plan_units(root, fl) if root is not None else [(None, 0, P)]
        post_backward = root is not None and any(
            m.wRegularizer is not None or m.bRegularizer is not None or m.scaleW != 1.0 or m.scaleB != 1.0
            for m in root.all_modules())
        if not bool(config.get("bigdl.parameter.overlap")) or post_backward:
            # regularisers / gradient scales touch the gradients AFTER backward: nothing may leave before that
            # (pieces without a module get no hook: they are exchanged in end_step and pulled before forward starts)
            head = units[0][0]
            units = [(head, 0, P)] if not method_ranges else \
                [(head if i == 0 else None, o, l) for i, (o, l) in enumerate(method_ranges) if l]
        self.method_ranges = method_ranges
        self.buckets, ptot = build_buckets(units, self.world, cap, first_cap, method_ranges)
        self.part_len = ptot[self.rank]
        B = len(self.buckets)
        self.nb = B
        # symmetric buffer: bf16 gradients | bf16 weights | flags (3N words per bucket + N for the clip exchange) | clip slots
        self._off_g16 = 0
        self._off_w16 = (P * 2 + 255) // 256 * 256
        self._off_flags = self._off_w16 + (P * 2 + 255) // 256 * 256
        nflags = (3 * B + 1) * self.world
        self._off_slots = self._off_flags + (nflags * 4 + 255) // 256 * 256
        total = self._off_slots + 2 * self.world * 4 + 256
        self.sym = SymmetricBuffer(total, sync.group, self.device)
        self.g16 = self.sym.view(torch.int16, self._off_g16, P)
        self.w16 = self.sym.view(torch.bfloat16, self._off_w16, P)
        self._g16_ptrs = self.sym.peer_ptr_array(self._off_g16)
        self._w16_ptrs = self.sym.peer_ptr_array(self._off_w16)
        self._flag_ptrs = self.sym.peer_ptr_array(self._off_flags)
        self._slot_ptrs = self.sym.peer_ptr_array(self._off_slots)
        self._my_flags = self.sym.local_ptr + self._off_flags
        dev = self.device
        # 4 words per bucket: the in-grid barrier word holds a generation, and generations DEcrease from one launch to
        # the next inside a step (buckets go out last-first), so the launches must not share it
        self.local_sync = torch.zeros(4 * B, dtype=torch.int32, device=dev)
        self.epoch_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.clip_scale = torch.ones(1, dtype=torch.float32, device=dev)
        self.grad_norm = torch.zeros(1, dtype=torch.float32, device=dev)
        self.comm = torch.cuda.Stream(device=dev, priority=-1)
        self.max_blocks = int(config.get("bigdl.parameter.commBlocks"))
        # "push": the owner stores its new bf16 sub-slice into every peer's buffer right after the update (posted NVLink
        # writes, overlapped with backward); consumers only acquire the ready flag. "pull": consumers copy it themselves.
        self.publish = CM_PUBLISH if config.get("bigdl.parameter.allgather") == "push" else CM_PUBLISH_LOCAL
        self._hooked = []
        self._launched = [False] * B
        self._pull_events = [None] * B
        # "dry" instrument (bench.py exposed_comm_pct): the same launches with the peers taken out — every kernel sees a
        # world of one rank (its own buffers), so no NVLink traffic and no waiting, only the local work
        self.dry = False
        me = self.rank
        self._self_tables = tuple((ctypes.c_void_p * 1)(t[me]) for t in (self._g16_ptrs, self._w16_ptrs, self._flag_ptrs))
        self._step = None              # per-step launch context (set by begin_step)
        self._in_step = False
        self.stats = {"bucket_launches": 0, "pull_launches": 0, "overlapped": 0, "gathered_gemms": 0}
        self._gather_mods = []
        self._plan_gathered(float(config.get("bigdl.parameter.gatherGemmMB")))
        self._install_hooks()

    # ------------------------------------------------------------------ layout helpers
    def my_slices(self):
        """[(global start, length, partition offset)] of this rank, bucket order."""
        return [(b.slices[self.rank][0], b.slices[self.rank][1], b.poff[self.rank]) for b in self.buckets]

    def all_slices(self, rank: int):
        return [(b.slices[rank][0], b.slices[rank][1], b.poff[rank]) for b in self.buckets]

    def method_partition(self, mi: int):
        """(offset, length) inside this rank's partition arrays of the sub-slices that OptimMethod ``mi`` owns."""
        mine = [b for b in self.buckets if b.method == mi]
        if not mine:
            return 0, 0
        off = mine[0].poff[self.rank]
        return off, sum(b.slices[self.rank][1] for b in mine)

    # ------------------------------------------------------------------ all-gather fused into the consumer GEMM
    GATHER_BLOCK_N = 128

    def _plan_gathered(self, min_mb: float):
        """Large ``Linear`` weights are not pushed to the peers after the update: the rows a rank owns (whole 128-row
        tiles inside its sub-slice) stay in its buffer and every other rank's forward GEMM loads them from there
        (csrc/gemm_tcgen05.cu kAG). One such range per (bucket, rank) — the kernel's publish "hole"."""
        from ..nn.layers_core import Linear

        if self.world == 1 or min_mb <= 0 or self.world > 8:
            return
        bn = self.GATHER_BLOCK_N
        by_bucket: dict[int, list] = {}
        for m, name, off, numel in self.flat.entries:
            if not (isinstance(m, Linear) and name == "weight") or numel * 2 < min_mb * (1 << 20) or m.input_size % 8:
                continue
            b = next((bk for bk in self.buckets if bk.off <= off < bk.off + bk.length), None)
            if b is None or off + numel > b.off + b.length or b.gathered:
                continue                     # one gathered matrix per bucket (the kernel takes one hole per launch)
            K, N = m.input_size, m.output_size
            t0, t1, holes = [0] * self.world, [0] * self.world, []
            for r in range(self.world):
                s, l = b.slices[r]
                lo, hi = max(s, off), min(s + l, off + numel)
                if hi <= lo:
                    continue
                first = -(-(lo - off) // (K * bn))             # first whole tile inside the slice
                last = (hi - off) // (K * bn)                    # one past the last whole tile
                if hi == off + numel:
                    last = -(-N // bn)                           # the ragged last tile belongs to whoever owns its rows
                if last > first:
                    t0[r], t1[r] = first, last
                    hs = off + first * bn * K
                    holes.append((r, hs, min(last * bn * K, numel) - first * bn * K))
            if not holes:
                continue
            b.gathered = sorted((hs, hl) for _, hs, hl in holes)
            self._hole_of = getattr(self, "_hole_of", {})
            for r, hs, hl in holes:
                self._hole_of[(b.index, r)] = (hs, hl)
            plan = {"nranks": self.world, "rank": self.rank, "block_n": bn, "t0": t0, "t1": t1,
                    "ptrs": [self.sym.ptrs[r] + self._off_w16 + 2 * off for r in range(self.world)]}
            m.__dict__["_gather_plan"] = self._make_gather(plan)
            self._gather_mods.append(m)

    def _make_gather(self, plan):
        def gather():
            # only inside a training step is the local copy stale; outside, ensure_weights_current() has filled it
            if not self._in_step or not torch.is_grad_enabled():
                return None
            self.stats["gathered_gemms"] += 1
            return plan
        return gather

    # ------------------------------------------------------------------ hooks
    def _install_hooks(self):
        for b in self.buckets:
            m = b.first
            if m is None:
                continue
            prev = m.__dict__.get("_pre_call_hook")
            m.__dict__["_pre_call_hook"] = self._make_hook(b.index, prev)
            self._hooked.append((m, prev))

    def remove_hooks(self):
        for m, prev in self._hooked:
            if prev is None:
                m.__dict__.pop("_pre_call_hook", None)
            else:
                m.__dict__["_pre_call_hook"] = prev
        self._hooked = []
        for m in self._gather_mods:
            m.__dict__.pop("_gather_plan", None)
        self._gather_mods = []

    def _make_hook(self, i: int, prev):
        def hook(module, x):
            if prev is not None:
                x = prev(module, x)
            if not self._in_step:
                return x
            ev = self._pull_events[i]
            if ev is not None:                 # this bucket's remote weights: wait (device side) for their pull
                torch.cuda.current_stream(self.device).wait_event(ev)
                self._pull_events[i] = None
            if (i > 0 and torch.is_grad_enabled() and module.isTraining() and isinstance(x, torch.Tensor)
                    and x.requires_grad):
                return _BucketBoundary.apply(x, self, i)
            return x
        return hook

    def _on_boundary(self, i: int):
        """Backward has passed bucket ``i``'s first layer input: buckets i.. hold final gradients."""
        if not self._in_step:
            return
        for k in range(self.nb - 1, i - 1, -1):
            if not self._launched[k]:
                self._launch_bucket(k, overlapped=True)

    # ------------------------------------------------------------------ one step
    def begin_step(self, plans, hp_dev: int | None, avg: float, l2_clip: float | None):
        """``plans``: per OptimMethod ``(kind, hp_bytes, s0, s1)`` whose state tensors cover that method's part of this
        rank's partition; ``hp_dev``: device address of the OptimHP table (graph replay) or None (host values)."""
        from ..ops import _lib

        self._step = {"plans": plans, "hp_dev": hp_dev, "avg": avg, "clip": l2_clip}
        self._launched = [False] * self.nb
        self._in_step = True
        main = torch.cuda.current_stream(self.device)
        self.comm.wait_stream(main)
        lib, w = _lib.cuda_lib(), self.world
        if w > 1 and not self.dry:
            with torch.cuda.stream(self.comm):
                st = _lib.stream_ptr()
                for b in self.buckets:
                    segs = self._pull_segments(b)
                    if not segs:
                        self._pull_events[b.index] = None
                        continue
                    arr = (ctypes.c_longlong * (3 * len(segs)))(*[v for s in segs for v in s])
                    _lib.check(lib.bdl_pull_gather(self._w16_ptrs, self._my_flags, w, self.rank, arr, len(segs), self.nb,
                                                   b.index + 1, 1, self.epoch_dev.data_ptr(), 3 * w * b.index,
                                                   self.max_blocks, st), "bdl_pull_gather")
                    _lib.count_launch()
                    self.stats["pull_launches"] += 1
                    ev = torch.cuda.Event()
                    ev.record(self.comm)
                    self._pull_events[b.index] = ev
            for b in self.buckets:             # no hook to wait in: the forward may not start before these arrive
                if b.first is None and self._pull_events[b.index] is not None:
                    main.wait_event(self._pull_events[b.index])
                    self._pull_events[b.index] = None

    def _pull_segments(self, b: Bucket):
        """(peer, start, length) to fetch before forward may use bucket ``b``; length 0 = only wait for that owner's
        ready flag (its slice was pushed, or a fused GEMM gathers it)."""
        segs = []
        for k in range(1, self.world):
            r = (self.rank + k) % self.world
            s, l = b.slices[r]
            if self.publish == CM_PUBLISH or l == 0:
                segs.append((r, s, 0))
                continue
            pieces = [(gs, gl) for gs, gl in self._subtract(s, l, b.gathered) if gl]
            segs.extend((r, gs, gl) for gs, gl in pieces)
            if not pieces:
                segs.append((r, s, 0))
        return segs

    @staticmethod
    def _subtract(s: int, l: int, holes):
        """[s, s+l) minus the ``holes`` ranges (sorted, disjoint)."""
        out, pos, end = [], s, s + l
        for hs, hl in holes:
            he = hs + hl
            if he <= pos or hs >= end:
                continue
            if hs > pos:
                out.append((pos, hs - pos))
            pos = max(pos, he)
        if pos < end:
            out.append((pos, end - pos))
        return out

    def _launch_bucket(self, k: int, overlapped: bool, mode_override: int | None = None):
        from ..ops import _lib

        st_ctx = self._step
        b = self.buckets[k]
        sync = self.sync
        lib = _lib.cuda_lib()
        kind, hp, s0, s1 = st_ctx["plans"][b.method]
        moff, _ = self.method_partition(b.method)
        start, length = b.slices[self.rank]
        poff = b.poff[self.rank]
        parity = (CM_PARITY | CM_DIVIDE) if sync.truncate else 0
        clip = st_ctx["clip"]
        if mode_override is not None:
            mode = mode_override
        elif clip is None:
            mode = (CM_COMPRESS | CM_REDUCE | CM_UPDATE | self.publish | CM_BARRIER_BEGIN | CM_SIGNAL_READY | parity)
        else:       # phase 1 of global-norm clipping: reduce + keep the averaged gradient + partial sum of squares
            mode = CM_COMPRESS | CM_REDUCE | CM_WRITE_GRAD | CM_SUMSQ | CM_BARRIER_BEGIN | parity
        hp_dev = st_ctx["hp_dev"]
        hsz = lib.bdl_optim_hp_size()
        buf = ctypes.create_string_buffer(hp, len(hp)) if hp_dev is None else None

        def at(t, o):
            return None if t is None else t.data_ptr() + 4 * o

        hole = getattr(self, "_hole_of", {}).get((k, self.rank), (0, 0))
        main = torch.cuda.current_stream(self.device)
        self.comm.wait_stream(main)
        with torch.cuda.stream(self.comm):
            g16t, w16t, flt = self._self_tables if self.dry else (self._g16_ptrs, self._w16_ptrs, self._flag_ptrs)
            _lib.check(lib.bdl_bucket_rs_update(
                g16t, w16t, flt, 1 if self.dry else self.world, 0 if self.dry else self.rank, b.off, b.length, start, length,
                self.flat.grads.data_ptr(), at(sync.weight_partition, poff), at(sync.gradient_partition, poff),
                at(s0, poff - moff), at(s1, poff - moff), kind,
                ctypes.cast(buf, ctypes.c_void_p) if buf is not None else None,
                None if hp_dev is None else hp_dev + hsz * b.method, mode, self.nb, k + 1,
                self.epoch_dev.data_ptr(), 3 * self.world * k, self.local_sync.data_ptr() + 16 * k, self.sumsq.data_ptr(),
                self.clip_scale.data_ptr(), hole[0], hole[1],
                # inside backward: enough CTAs for the bucket's bytes (one per 256 KB of fp32 gradients), capped so the
                # compute stream keeps most SMs; after backward: every SM
                min(self.max_blocks, max(16, b.length * 4 // (256 << 10))) if overlapped else 0,
                _lib.stream_ptr()), "bdl_bucket_rs_update")
        _lib.count_launch()
        self._launched[k] = True
        self.stats["bucket_launches"] += 1
        if overlapped:
            self.stats["overlapped"] += 1

    def end_step(self):
        """Launch what backward could not (bucket 0: its input carries no gradient), finish global-norm clipping, join
        the communication stream and advance the generation."""
        from ..ops import _lib

        lib = _lib.cuda_lib()
        st_ctx = self._step
        for k in range(self.nb - 1, -1, -1):
            if not self._launched[k]:
                self._launch_bucket(k, overlapped=False)
        if st_ctx["clip"] is not None:
            main = torch.cuda.current_stream(self.device)
            self.comm.wait_stream(main)
            with torch.cuda.stream(self.comm):
                _lib.check(lib.bdl_peer_clip_scale(self._slot_ptrs, self._flag_ptrs, self.world, self.rank, 1, 1,
                                                   self.epoch_dev.data_ptr(), 3 * self.world * self.nb,
                                                   self.sumsq.data_ptr(), float(st_ctx["clip"]),
                                                   self.clip_scale.data_ptr(), self.grad_norm.data_ptr(),
                                                   _lib.stream_ptr()), "bdl_peer_clip_scale")
                _lib.count_launch()
            parity = CM_PARITY if self.sync.truncate else 0
            for k in range(self.nb - 1, -1, -1):
                self._launch_bucket(k, overlapped=False,
                                    mode_override=CM_UPDATE | self.publish | CM_SIGNAL_READY | CM_CLIP_DEV | parity)
        main = torch.cuda.current_stream(self.device)
        main.wait_stream(self.comm)
        _lib.check(lib.bdl_counter_add(self.epoch_dev.data_ptr(), 1, _lib.stream_ptr()), "bdl_counter_add")
        _lib.count_launch()
        self._in_step = False
        self.flat.dirty = False

    def pull_all(self):
        """Make every remote slice current in the local compute buffer NOW (validation, checkpoints, end of training):
        the just-in-time pulls only run at the start of a training step."""
        if self.world == 1:
            return
        from ..ops import _lib

        lib = _lib.cuda_lib()
        main = torch.cuda.current_stream(self.device)
        self.comm.wait_stream(main)
        with torch.cuda.stream(self.comm):
            for b in self.buckets:
                segs = []
                for k in range(1, self.world):
                    r = (self.rank + k) % self.world
                    s, l = b.slices[r]
                    if self.publish == CM_PUBLISH:      # pushed already; only the ranges left to a fused GEMM remain
                        holes = [(max(s, hs), min(s + l, hs + hl) - max(s, hs)) for hs, hl in b.gathered]
                        holes = [(hs, hl) for hs, hl in holes if hl > 0]
                        segs.extend((r, hs, hl) for hs, hl in holes)
                        if not holes:
                            segs.append((r, s, 0))
                    elif l:
                        segs.append((r, s, l))
                if not segs:
                    continue
                arr = (ctypes.c_longlong * (3 * len(segs)))(*[v for s in segs for v in s])
                _lib.check(lib.bdl_pull_gather(self._w16_ptrs, self._my_flags, self.world, self.rank, arr, len(segs),
                                               self.nb, b.index + 1, 1, self.epoch_dev.data_ptr(),
                                               3 * self.world * b.index, self.max_blocks, _lib.stream_ptr()),
                           "bdl_pull_gather")
                _lib.count_launch()
        main.wait_stream(self.comm)

    def close(self):
        self.remove_hooks()
        if self.sym is not None:
            self.sym.close()
            self.sym = None


class _BucketBoundary(torch.autograd.Function):
    """Identity in forward; its backward runs once every layer called after it has run its own backward."""

    @staticmethod
    def forward(ctx, x, plane, index):
        ctx.plane, ctx.index = plane, index
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        ctx.plane._on_boundary(ctx.index)
        return g, None, None
