"""Scale bookkeeping for quantised inference on FLOAT modules (reference: S/nn/MklInt8Convertible.scala:53-500 —
``calcScales(input)`` runs a forward pass and records, for every module in the tree, the abs-max of its input(s), its
output(s) and its weights; masks select the granularity: 0 = one scale per tensor, bit d set = one scale per index of
dimension d (the reference uses weight mask 1 → per output channel). The scales are persisted in the ``.bigdl`` file
(bigdl.proto fields 17-23, see serialization/module_io.py) and consumed by ``Quantizer.quantize``, which turns recorded
input scales into static activation scales of the fp8 layers.

The methods are attached to ``Module`` at import time so every layer and container has them, like the Scala trait."""
from __future__ import annotations

import torch

from ..utils.table import Table
from .module import Module


def _state(m: Module) -> dict:
    s = m.__dict__.get("_mkl_int8")
    if s is None:
        s = {"inputDimMask": 0, "outputDimMask": 0, "weightDimMask": 0, "inputScales": [], "outputScales": [],
             "weightScales": []}
        m.__dict__["_mkl_int8"] = s
    return s


def _tensor_scale(t: torch.Tensor, mask: int) -> list[float]:
    """abs-max over every dimension whose mask bit is clear (MklInt8Convertible.calcTensorScale :159-175)."""
    x = t.detach().float().abs()
    if mask == 0 or x.dim() == 0:
        return [float(x.max())] if x.numel() else [0.0]
    keep = [d for d in range(x.dim()) if (mask >> d) & 1]
    reduce = [d for d in range(x.dim()) if d not in keep]
    if reduce:
        x = x.amax(dim=reduce)
    return [float(v) for v in x.reshape(-1)]


def _activity_scales(act, mask: int) -> list[list[float]]:
    if isinstance(act, torch.Tensor):
        return [_tensor_scale(act, mask)]
    if isinstance(act, Table):
        out = []
        for v in act.to_list():
            out.extend(_activity_scales(v, mask))
        return out
    return []


def _merge(old: list, new: list) -> list:
    """Running maximum across calibration batches."""
    if len(old) != len(new):
        return new
    return [[max(a, b) for a, b in zip(o, n)] if len(o) == len(n) else n for o, n in zip(old, new)]


def calcScales(self: Module, inputActvt):
    """Forward ``inputActvt`` and record input/output/weight scales of every module in the tree."""
    if inputActvt is None:
        raise ValueError("calcScales needs a non-empty input activity")
    mods = list(self.all_modules())
    originals = {}

    def wrap(m):
        orig = m.updateOutput

        def recording(x, _m=m, _orig=orig):
            y = _orig(x)
            st = _state(_m)
            st["inputScales"] = _merge(st["inputScales"], _activity_scales(x, st["inputDimMask"]))
            st["outputScales"] = _merge(st["outputScales"], _activity_scales(y, st["outputDimMask"]))
            return y
        originals[id(m)] = orig
        m.__dict__["updateOutput"] = recording

    for m in mods:
        wrap(m)
    try:
        was = self.isTraining()
        self.evaluate()
        self.forward(inputActvt)
        if was:
            self.training()
    finally:
        for m in mods:
            m.__dict__.pop("updateOutput", None)
    for m in mods:
        w = m._parameters.get("weight")
        if w is not None and w.dim() >= 2:
            st = _state(m)
            st["weightScales"] = [_tensor_scale(w, st["weightDimMask"])]
    return self


def _getter(key):
    def get(self):
        return [list(v) for v in _state(self)[key]]
    return get


def _setter(key):
    def set_(self, scales):
        _state(self)[key] = [list(v) for v in scales]
        return self
    return set_


def _mask_getter(key):
    def get(self):
        return _state(self)[key]
    return get


def _mask_setter(key):
    def set_(self, mask: int, overrideSubmodules: bool = False):
        _state(self)[key] = int(mask)
        if overrideSubmodules:
            for m in self.all_modules():
                _state(m)[key] = int(mask)
        return self
    return set_


Module.calcScales = calcScales
Module.calc_scales = calcScales
for _name, _key in (("Input", "inputScales"), ("Output", "outputScales"), ("Weight", "weightScales")):
    setattr(Module, f"get{_name}Scales", _getter(_key))
    setattr(Module, f"set{_name}Scales", _setter(_key))
for _name, _key in (("Input", "inputDimMask"), ("Output", "outputDimMask"), ("Weight", "weightDimMask")):
    setattr(Module, f"get{_name}DimMask", _mask_getter(_key))
    setattr(Module, f"set{_name}DimMask", _mask_setter(_key))
