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
    def __new__(cls, *a, **kw):
        # ``ModelBroadcast()`` asks the configured factory (bigdl.ModelBroadcastFactory), as the reference's companion
        # ``apply`` does (ModelBroadcast.scala:75-93)
        if cls is ModelBroadcast and not kw.pop("_direct", False):
            return ModelBroadcastFactory.configured().create(*a, **kw)
        return object.__new__(cls)

    def __init__(self, src: int = 0, group=None, _direct: bool = False):
        import uuid as _uuid

        self.src, self.group = src, group
        self._model = None
        self._uuid = str(_uuid.uuid4())

    def uuid(self) -> str:
        return self._uuid

    def broadcast(self, model=None):
        """Call on every rank; ``model`` is required on ``src`` and ignored elsewhere."""
        from ..nn.module import Module

        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        if world == 1:
            self._model = model
            CachedModels.add(self._uuid, model)
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
        CachedModels.add(self._uuid, model)
        return self

    def value(self, initGradient: bool = False, shareWeight: bool = True):
        m = self._model
        if not shareWeight:
            m = m.cloneModule()
        if initGradient:
            m.zeroGradParameters()
        return m


class ModelBroadcastImp(ModelBroadcast):
    """The stock implementation (ModelBroadcast.scala:95-200). ``applyProtoBuffer`` chose between Java serialisation
    and protobuf for the structure in the reference; the structure always travels as ``.bigdl`` protobuf bytes here,
    so the flag only records what was asked for."""

    def __init__(self, applyProtoBuffer: bool = False, src: int = 0, group=None):
        if getattr(self, "_built", False):
            return          # ``ModelBroadcast()`` → factory → this object; Python then calls __init__ a second time
        super().__init__(src, group, _direct=True)
        self.applyProtoBuffer = applyProtoBuffer
        self._built = True


class ModelBroadcastFactory:
    """Override how models are broadcast: subclass, implement ``create()`` and name the class in the
    ``bigdl.ModelBroadcastFactory`` property as ``package.module.ClassName`` (ModelBroadcastFactory.scala:23-42)."""

    def create(self, *a, **kw) -> ModelBroadcast:  # pragma: no cover - interface
        raise NotImplementedError

    @staticmethod
    def configured() -> "ModelBroadcastFactory":
        from ..utils import config

        name = str(config.get("bigdl.ModelBroadcastFactory", "default") or "default")
        if name in ("default", "DefaultModelBroadcastFactory"):
            return DefaultModelBroadcastFactory()
        if name in ("protobuf", "ProtoBufferModelBroadcastFactory"):
            return ProtoBufferModelBroadcastFactory()
        import importlib

        mod, _, cls = name.rpartition(".")
        return getattr(importlib.import_module(mod), cls)()


class DefaultModelBroadcastFactory(ModelBroadcastFactory):
    def create(self, *a, **kw):
        return ModelBroadcastImp(False, *a, **kw)


class ProtoBufferModelBroadcastFactory(ModelBroadcastFactory):
    def create(self, *a, **kw):
        return ModelBroadcastImp(True, *a, **kw)


class ModelInfo:
    """(uuid, model) pair whose pickled form carries only the uuid; the receiving process looks the model up in
    ``CachedModels`` (ModelBroadcast.scala:278-299)."""

    def __init__(self, uuid: str, model=None):
        self.uuid, self.model = uuid, model

    def __getstate__(self):
        if self.model is not None:
            CachedModels.add(self.uuid, self.model)
        return {"uuid": self.uuid}

    def __setstate__(self, st):
        self.uuid = st["uuid"]
        found = CachedModels.get(self.uuid)
        self.model = found[0] if found else None


class CachedModels:
    """Process-wide registry of broadcast models by uuid, so stale generations can be released
    (ModelBroadcast.scala:301-350)."""

    import threading as _threading

    _lock = _threading.Lock()
    _models: dict = {}

    @classmethod
    def add(cls, uuid: str, model):
        with cls._lock:
            held = cls._models.setdefault(uuid, [])
            if not any(m is model for m in held):
                held.append(model)

    @classmethod
    def get(cls, uuid: str):
        with cls._lock:
            return list(cls._models.get(uuid, []))

    @classmethod
    def deleteAll(cls, currentKey: str):
        """Release every generation except ``currentKey``."""
        with cls._lock:
            for k in [k for k in cls._models if k != currentKey]:
                for m in cls._models.pop(k):
                    if hasattr(m, "release"):
                        m.release()

    @classmethod
    def deleteKey(cls, key: str):
        with cls._lock:
            for m in cls._models.pop(key, []):
                if hasattr(m, "release"):
                    m.release()
