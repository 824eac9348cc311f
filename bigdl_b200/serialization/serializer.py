"""The reference's entry-point objects of the ``.bigdl`` serializer (S/utils/serializer/ModuleLoader.scala:47-230,
ModulePersister :230-370, ModuleSerializer.scala:68-190) over ``module_io``: load / save files, copy trained weights
into a model definition by layer name, dump a weight-free text definition, register custom layer classes."""
from __future__ import annotations

import torch

from . import module_io as _io


class ModuleLoader:
    @staticmethod
    def loadFromFile(modelPath: str, weightPath: str | None = None):
        return _io.load_module(modelPath, weightPath)

    @staticmethod
    def loadFromDefinition(definition, modelPath: str, layers=None):
        """Copy ``weight`` / ``bias`` from the model stored at ``modelPath`` into ``definition`` for every layer name in
        ``layers`` (default: every layer of ``definition``). A named layer that has parameters in the definition must
        exist in the file."""
        loaded = _io.load_module(modelPath)
        names = set(layers) if layers is not None else {m.getName() for m in definition.all_modules()}
        mine, theirs = definition.getParametersTable(), loaded.getParametersTable()
        with torch.no_grad():
            for name in names:
                if name not in mine:
                    continue
                if name not in theirs:
                    raise KeyError(f"{name} does not exist in loaded module")
                for p in ("weight", "bias"):
                    if p in mine[name] and mine[name][p] is not None:
                        src = theirs[name][p]
                        mine[name][p].copy_(src.reshape(mine[name][p].shape).to(mine[name][p].dtype))
        definition._mark_weights_dirty()
        return definition


class ModulePersister:
    @staticmethod
    def saveToFile(modelPath: str, weightPath: str | None, module, overwrite: bool = False):
        _io.save_module(module, modelPath, weightPath, overwrite)

    @staticmethod
    def saveModelDefinitionToFile(definitionPath: str, module, overwrite: bool = False):
        """The protobuf TEXT form of the model with every tensor payload removed — a readable structure dump."""
        import os

        from google.protobuf import text_format

        if os.path.exists(definitionPath) and not overwrite:
            raise FileExistsError(f"{definitionPath} already exists (pass overwrite=True)")
        from .snapshot import _portable_copy

        root = _io._module_to_proto(_portable_copy(module), _io._Ctx())

        def strip(p):
            p.ClearField("weight")
            p.ClearField("bias")
            del p.parameters[:]
            p.hasParameters = False
            for s in p.subModules:
                strip(s)
        strip(root)
        _io._write_bytes(definitionPath, [text_format.MessageToString(root).encode()])


class ModuleSerializable:
    """Marker for layers with their own (de)serialisation hooks. A custom layer takes part in ``saveModule`` /
    ``loadModule`` once its class is registered; constructor arguments are captured automatically, and a layer may
    define ``_ser_children()`` / ``_deser_children(mods)`` when its children cannot be rebuilt from them."""


class ModuleSerializer:
    @staticmethod
    def registerModule(moduleType, serializer=None):
        """``registerModule("my.pkg.MyLayer", MyLayer)`` or ``registerModule(MyLayer)``."""
        if isinstance(moduleType, str):
            if serializer is None:
                raise ValueError("registerModule(name, cls): the layer class is required")
            return _io.register_module_class(serializer, moduleType)
        return _io.register_module_class(moduleType)

    @staticmethod
    def serialize(module) -> bytes:
        """The module as ``.bigdl`` bytes (tensors inline)."""
        import os
        import tempfile

        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "m.bigdl")
            _io.save_module(module, p, None, True)
            with open(p, "rb") as f:
                return f.read()

    @staticmethod
    def load(data: bytes):
        import os
        import tempfile

        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "m.bigdl")
            with open(p, "wb") as f:
                f.write(data)
            return _io.load_module(p)
