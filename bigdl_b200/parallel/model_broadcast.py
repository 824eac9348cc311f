"""Model broadcast (S/models/utils/ModelBroadcast.scala:39-200): ship the structure once and the weights once; the
receivers rebuild the model around the received flat buffer so replicas on one host can share it.

B200 form: rank 0 serialises the module tree WITHOUT tensors (``.bigdl`` protobuf, weights stripped), broadcasts those
bytes, then broadcasts the flat fp32 weight vector as a tensor collective (NCCL over NVLink for GPU models, gloo on CPU).
``value(initGradient, shareWeight)`` mirrors the reference signature."""
from __future__ import annotations

import io
import os
import tempfile

import torch
import torch.distributed as dist


class ModelBroadcast:
    def __init__(self, src: int = 0, group=None):
        self.src, self.group = src, group
        self._model = None

    def broadcast(self, model=None):
        """Call on every rank; ``model`` is required on ``src`` and ignored elsewhere."""
        from ..nn.module import Module

        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        if world == 1:
            self._model = model
            return self
        rank = dist.get_rank(self.group)
        if rank == self.src:
            with tempfile.TemporaryDirectory() as d:
                p = os.path.join(d, "m.bigdl")
                model.saveModule(p, overWrite=True)
                blob = open(p, "rb").read()
            objs = [blob, str(model.device())]
        else:
            objs = [None, None]
        dist.broadcast_object_list(objs, src=self.src, group=self.group)
        if rank != self.src:
            with tempfile.TemporaryDirectory() as d:
                p = os.path.join(d, "m.bigdl")
                open(p, "wb").write(objs[0])
                model = Module.loadModule(p)
        w, _ = model.getParameters()
        if w.numel():
            dist.broadcast(w, src=self.src, group=self.group)       # weights travel once, as one flat tensor
            model._mark_weights_dirty()
        for m in model.all_modules():                                # running statistics and other buffers
            for name, b in m._buffers.items():
                if isinstance(b, torch.Tensor) and b.numel() and b.is_floating_point():
                    dist.broadcast(b, src=self.src, group=self.group)
        self._model = model
        return self

    def value(self, initGradient: bool = False, shareWeight: bool = True):
        m = self._model
        if not shareWeight:
            m = m.cloneModule()
        if initGradient:
            m.zeroGradParameters()
        return m
