"""Constructor reflection (reference: S/utils/ReflectionUtils.scala — finds a layer's primary constructor, reads its
arguments off an instance and re-instantiates from an argument map; used by the module serializer and by the blas ↔ IR
conversion). Python layers record their constructor arguments when they are built (``Module._ctor_args``), so the
same three operations are direct."""
from __future__ import annotations

import inspect


class ReflectionUtils:
    @staticmethod
    def getPrimCtorParams(cls) -> list[str]:
        """Names of the constructor parameters, in order (without ``self``)."""
        return [n for n in inspect.signature(cls.__init__).parameters if n != "self"]

    @staticmethod
    def getCtorArgs(module) -> dict:
        """The arguments ``module`` was constructed with (defaults filled in)."""
        return dict(module.__dict__.get("_ctor_args", {}) or {})

    @staticmethod
    def reflection(cls, args: dict):
        """Instantiate ``cls`` from an argument map, ignoring keys the constructor does not take."""
        sig = inspect.signature(cls.__init__)
        accepts_kw = any(p.kind is inspect.Parameter.VAR_KEYWORD for p in sig.parameters.values())
        kw = {k: v for k, v in args.items() if accepts_kw or k in sig.parameters}
        var_pos = [n for n, p in sig.parameters.items() if p.kind is inspect.Parameter.VAR_POSITIONAL]
        pos = []
        for n in var_pos:
            pos = list(kw.pop(n, ()))
        return cls(*pos, **kw)

    @staticmethod
    def reflectToIR(module):
        """(op name, constructor arguments) — what BlasToIR stores per node."""
        return type(module).__name__, ReflectionUtils.getCtorArgs(module)

    @staticmethod
    def reflectFromIR(op: str, args: dict):
        """Rebuild a layer from its IR description, looking the class up in ``nn.layer``."""
        from ..nn import layer as L

        cls = getattr(L, op, None)
        if cls is None:
            raise KeyError(f"no layer named {op}")
        return ReflectionUtils.reflection(cls, args)
