"""ParallelOptimizer — layer-wise, bucketed, asynchronous data-parallel training (S/optim/ParallelOptimizer.scala:
48-791; hooks in S/nn/abstractnn/AbstractModule.scala:282-297 ``asyncGradient`` and :1184-1200 ``updateParameter``).

Differences from DistriOptimizer, as in the reference: gradients are all-reduced per bucket of layers *while backward
is still running*, every rank keeps the full (non-sharded) optimizer state and applies the update for a bucket lazily,
just before the next forward reaches that bucket's first layer; each bucket owns a clone of the OptimMethod.

Mechanics on the autograd substrate: the first parameterised layer of every bucket gets a pre-call hook. In forward
the hook (1) drains that bucket's pending all-reduce and applies its optimizer step, then (2) routes the layer input
through ``_BucketBoundary``, an identity whose backward runs after every layer created later has finished its
backward (the autograd engine executes nodes in reverse creation order on one device) — exactly the moment all
buckets from this one onwards have final gradients, so it calls ``sync.put``.
"""
from __future__ import annotations

import torch

from ..nn.module import Module
from ..parallel.distri_parameter_synchronizer import DistriParameterSynchronizer
from ..utils import config
from .optimizer import DistriOptimizer, _train_loop


class _BucketBoundary(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, strategy, index):
        ctx.strategy, ctx.index = strategy, index
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        ctx.strategy._on_boundary(ctx.index)
        return g, None, None


class ParallelStrategy:
    """Owns the buckets, the synchroniser and the per-bucket optimizer clones for one training run."""

    def __init__(self, model: Module, optim_method, world: int, bucket_bytes: int | None = None, compress="auto"):
        self.model, self.world = model, world
        flat = model._flat
        self.flat = flat
        cap = bucket_bytes if bucket_bytes is not None else int(float(config.get("bigdl.parallel.bucketMB")) * (1 << 20))
        leaves = []                      # (module, flat offset, length) in forward/flat order
        for m, off, ln in flat.module_ranges():
            if ln:
                leaves.append((m, off, ln))
        self.buckets: list[dict] = []
        cur = None
        for m, off, ln in leaves:
            if cur is None or (cur["len"] * 4 >= cap):
                cur = {"first": m, "off": off, "len": 0, "modules": []}
                self.buckets.append(cur)
            cur["len"] = off + ln - cur["off"]
            cur["modules"].append(m)
        n = len(self.buckets)
        self.sync = DistriParameterSynchronizer(world, compress=compress)
        self.methods = []
        for i, b in enumerate(self.buckets):
            b["name"] = f"bucket{i}"
            w = flat.weights[b["off"]:b["off"] + b["len"]]
            g = flat.grads[b["off"]:b["off"] + b["len"]]
            # priority = #layers − forward index: the first layers are needed first (ParallelOptimizer.scala:494-505)
            self.sync.init(b["name"], b["len"], priority=n - i, weights=w, grads=g, offset=b["off"])
            self.methods.append(optim_method if i == 0 else optim_method.clone())
            b["first"].__dict__["_pre_call_hook"] = self._make_hook(i)
        self._launched_upto = n          # buckets with index >= this have been put in the current backward
        self.updates = 0

    def _make_hook(self, i: int):
        def hook(module, x):
            self._apply_pending(i)
            if torch.is_grad_enabled() and module.isTraining() and isinstance(x, torch.Tensor) and x.requires_grad:
                return _BucketBoundary.apply(x, self, i)
            return x
        return hook

    def _on_boundary(self, i: int):
        # everything from bucket i to the end has final gradients; launch the ones not launched yet (last first —
        # the deterministic order every rank follows)
        for k in range(self._launched_upto - 1, i - 1, -1):
            self.sync.put(self.buckets[k]["name"])
        self._launched_upto = min(self._launched_upto, i)

    def after_backward(self):
        """Buckets whose boundary could not fire (the first layer's input carries no gradient) are launched here."""
        self.sync.flush()
        self._launched_upto = len(self.buckets)

    def _apply_pending(self, i: int):
        b = self.buckets[i]
        got = self.sync.get(b["name"])
        if got is None:
            return
        w, g = got
        om = self.methods[i]
        fl = self.flat
        w16 = None if fl.compute is None else fl.compute[b["off"]:b["off"] + b["len"]].data
        # ``.data``: the flat buffer's autograd version counter is shared by every parameter view, and layers of
        # EARLIER buckets have already saved their weights for this iteration's backward — updating a later bucket's
        # slice must not invalidate them (the slices are disjoint, so this is safe).
        om.update(w.data, g.data, w16=w16)
        g.data.zero_()
        if fl.compute is None:
            fl.dirty = True
        self.updates += 1

    def drain(self):
        for i in range(len(self.buckets)):
            self._apply_pending(i)

    def sync_states(self, lead_state):
        for om in self.methods[1:]:
            om.state["epoch"], om.state["neval"] = lead_state["epoch"], lead_state["neval"]

    def detach(self):
        for b in self.buckets:
            b["first"].__dict__.pop("_pre_call_hook", None)
        self.sync.clear()


class ParallelOptimizer(DistriOptimizer):
    """Drop-in replacement for DistriOptimizer with bucket-wise overlapped synchronisation."""

    def __init__(self, *a, bucket_mb: float | None = None, **kw):
        super().__init__(*a, **kw)
        self._bucket_mb = bucket_mb

    def optimize(self):
        return _train_loop(self, distributed=True, strategy_factory=self._make_strategy)

    def _make_strategy(self, model, world):
        single = self._single_method()
        if single is None:
            raise NotImplementedError("ParallelOptimizer clones ONE optim method per bucket; use setOptimMethod")
        cap = None if self._bucket_mb is None else int(self._bucket_mb * (1 << 20))
        return ParallelStrategy(model, single, world, bucket_bytes=cap)
