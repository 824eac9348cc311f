"""CUDA-graph captured training step.

A training iteration of a convnet is hundreds of kernel launches; issued eagerly from Python the host cannot keep a
B200 busy (Inception-v1: 12.5 ms of host work for 13 ms of step). The whole iteration — forward, criterion, backward,
gradient compression, the fused NVLink reduce-scatter + OptimMethod update + all-gather — is therefore captured once
into a CUDA graph and replayed; per step the host only (1) copies the next batch into the graph's static input
buffers, (2) computes the step's scalar hyper-parameters (learning-rate schedule, Adam bias correction, …) and ships
them to the device-resident ``OptimHP`` the kernels read, (3) replays.

Everything that varies between replays lives in device memory: ``OptimHP`` (csrc/optim_update.cuh), the dropout
Philox position and the barrier generation of the peer-memory flags (advanced by ``bdl_counter_add`` inside the graph).

This plays the role of the reference's "one Spark job per iteration" scheduling fix discussed in its white paper
(docs/docs/whitepaper.md:171-177, Drizzle group scheduling): amortise per-iteration launch overhead.
"""
from __future__ import annotations

import ctypes

import torch

from ..ops import _lib

_RING = 8


class GraphedTrainStep:
    def __init__(self, model, criterion, sync, optim_method, example_input: torch.Tensor, example_target: torch.Tensor,
                 world: int = 1, warmup_iters: int = 3, l2_clip: float | None = None):
        # one OptimMethod, or the per-submodule list in the order of ``sync.method_ranges``
        self.oms = list(optim_method) if isinstance(optim_method, (list, tuple)) else [optim_method]
        self.model, self.criterion, self.sync, self.om, self.world = model, criterion, sync, self.oms[0], world
        self.l2_clip = l2_clip
        dev = example_input.device
        self.device = dev
        from ..nn.criterion import fuse_final_logsoftmax

        fuse_final_logsoftmax(model, criterion)          # LogSoftMax + ClassNLL -> fused cross-entropy kernels
        self.x = torch.empty_like(example_input)
        self.y = torch.empty_like(example_target)
        self.x.copy_(example_input)
        self.y.copy_(example_target)
        lib = _lib.cuda_lib()
        self.hp_size = lib.bdl_optim_hp_size()
        nm = len(self.oms)
        self.hp_dev = torch.zeros(self.hp_size * nm, dtype=torch.uint8, device=dev)
        self.hp_host = [torch.zeros(self.hp_size * nm, dtype=torch.uint8).pin_memory() for _ in range(_RING)]
        self._ring = 0
        self.epoch_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        self.loss = None
        self.graph = None
        self._plan = None
        # eager warm-up on a side stream: lazily created buffers (conv workspace, concat caches, optimizer state)
        # must exist before capture. The warm-up iterations are real updates, so everything they touch is
        # snapshotted first and restored (in place — the graph keeps the addresses) afterwards.
        snap = self._snapshot()
        # Autograd ties every leaf's gradient accumulator to the stream that was current when the leaf first entered
        # a graph; accumulators born on the legacy default stream make the engine touch that stream at the end of
        # backward, which invalidates a capture. Re-leaf the parameters so their accumulators are created (during
        # warm-up) on the one side stream that is also used for the capture.
        self._releaf_parameters()
        s = torch.cuda.Stream(dev)
        self._stream = s
        s.wait_stream(torch.cuda.current_stream(dev))
        try:
            with torch.cuda.stream(s):
                for _ in range(warmup_iters):
                    self._push_hp()
                    self._body()
                self._push_hp()
            s.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, stream=s):
                self.loss = self._body()
            self.launches_per_replay = self._count_body_launches
        except Exception as e:  # noqa: BLE001 - capture is an optimisation; the caller falls back to eager steps
            self.graph = None
            torch.cuda.synchronize(dev)
            self._restore(snap)
            raise NotImplementedError(f"CUDA graph capture failed: {str(e).splitlines()[0]}") from e
        torch.cuda.current_stream(dev).wait_stream(s)
        self._restore(snap)

    def _releaf_parameters(self):
        seen = {}
        for m in self.model.all_modules():
            for n, p in list(m._parameters.items()):
                q = seen.get(id(p))
                if q is None:
                    q = p.detach()
                    q.requires_grad_(p.requires_grad)
                    q.grad = p.grad
                    seen[id(p)] = q
                m._parameters[n] = q

    def _snapshot(self):
        import copy

        fl = self.sync.flat
        mods = list(self.model.all_modules())
        return {
            "weights": fl.weights.clone(), "compute": None if fl.compute is None else fl.compute.clone(),
            "part": None if self.sync.weight_partition is None else self.sync.weight_partition.clone(),
            "buffers": [{k: b.clone() for k, b in m._buffers.items() if isinstance(b, torch.Tensor)} for m in mods],
            "oms": [{"om_state": om.state.clone(),
                     "om_attrs": {k: copy.deepcopy(v) for k, v in om.__dict__.items() if k not in ("state", "_hist")},
                     "hist_keys": set(om._hist.keys()), "hist": {k: v.clone() for k, v in om._hist.items()}}
                    for om in self.oms],
        }

    def _restore(self, snap):
        import torch.distributed as dist

        torch.cuda.synchronize(self.device)
        if self.world > 1 and dist.is_initialized():
            dist.barrier()
        fl = self.sync.flat
        with torch.no_grad():
            fl.weights.copy_(snap["weights"])
            if fl.compute is not None:
                fl.compute.copy_(snap["compute"])
            if snap["part"] is not None:
                self.sync.weight_partition.copy_(snap["part"])
            fl.grads.zero_()
            for m, saved in zip(self.model.all_modules(), snap["buffers"]):
                for k, b in saved.items():
                    m._buffers[k].copy_(b)
            for om, sn in zip(self.oms, snap["oms"]):
                for k, v in om._hist.items():
                    if k in sn["hist"]:
                        v.copy_(sn["hist"][k])
                    else:
                        v.zero_()               # created during warm-up: logically "not there yet"
        for om, sn in zip(self.oms, snap["oms"]):
            fresh = not sn["hist_keys"]
            # restore IN PLACE: the training loop holds an alias of this Table (epoch / neval bookkeeping), so the
            # object must stay the same or epoch-based schedules would read a frozen copy under graph replay
            live = om.state
            for k in list(live.keys()):
                del live[k]
            for k, v in sn["om_state"].items():
                live[k] = v
            for k, v in sn["om_attrs"].items():
                om.__dict__[k] = v
            if fresh and om._hist:
                om._hist_fresh = True      # first real step must behave like the very first update (SGD: v = g)
        torch.cuda.synchronize(self.device)
        if self.world > 1 and dist.is_initialized():
            dist.barrier()

    # one iteration, identical in warm-up, capture and (implicitly) replay
    def _body(self):
        before = _lib.LAUNCHES
        overlapped = self.sync.overlapped
        if overlapped:      # arm the per-bucket exchange: pulls of remote weights now, reduce+update from inside backward
            self.sync.begin_step(self.oms, self.world, self.l2_clip, hp_dev=self.hp_dev.data_ptr(), plans=self._plan)
        with torch.enable_grad():
            out = self.model._call(self.x)
            loss = self.criterion._call(out, self.y)
        seed = torch.ones((), dtype=loss.dtype, device=loss.device)
        seed._bdl_is_one = True
        torch.autograd.backward(loss, seed)
        from ..ops.functions import join_side_streams

        join_side_streams()
        self.model._apply_regularizers()
        if overlapped:
            self.sync.end_step()
        else:
            self.sync.step_device_hp(self._plan[0], self.hp_dev.data_ptr(), self.epoch_dev.data_ptr())
        self._count_body_launches = _lib.LAUNCHES - before
        return loss.detach()

    def _push_hp(self):
        """Advance the OptimMethod's scalar state by one step on the host and ship the resulting OptimHP."""
        plans = []
        for i, om in enumerate(self.oms):
            om.gradient_scale = self.sync.grad_scale_for(self.world)
            plans.append(om.native_plan(self.sync.method_target(i), None))
        self._plan = plans
        buf = self.hp_host[self._ring % _RING]
        self._ring += 1
        buf.copy_(torch.frombuffer(bytearray(b"".join(p[1] for p in plans)), dtype=torch.uint8))
        self.hp_dev.copy_(buf, non_blocking=True)

    def run(self, x_host: torch.Tensor | None = None, y_host: torch.Tensor | None = None) -> torch.Tensor:
        """One training step. ``x_host``/``y_host`` (ideally pinned) are copied into the static input buffers."""
        if x_host is not None:
            self.x.copy_(x_host, non_blocking=True)
        if y_host is not None:
            self.y.copy_(y_host, non_blocking=True)
        cur = torch.cuda.current_stream(self.device)
        self._stream.wait_stream(cur)            # inputs / hyper-parameters were enqueued on the caller's stream
        self._push_hp()
        self._stream.wait_stream(cur)
        with torch.cuda.stream(self._stream):
            self.graph.replay()
        cur.wait_stream(self._stream)
        _lib.count_launch(self.launches_per_replay)
        return self.loss
