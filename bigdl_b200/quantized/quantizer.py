"""Model quantiser (S/nn/quantized/Quantizer.scala:27-140, Quantizable.scala): walks containers, graphs and recurrent
cells and swaps Linear / SpatialConvolution / SpatialDilatedConvolution for their fp8 versions. ``calibrate`` is the
B200 form of ``MklInt8Convertible.calcScales`` (S/nn/MklInt8Convertible.scala:53-98): it runs sample batches, records the
abs-max of every quantised layer's input, freezes those as static scales and links producer→consumer pairs so the
producer's GEMM epilogue writes e4m3 for the consumer directly."""
from __future__ import annotations

import torch

from ..nn import layers_core as LC
from ..nn.module import Container, Graph, Module, Sequential
from . import layers as QL


class Quantizer:
    _registry: dict[type, type] = {}

    @classmethod
    def register(cls, float_cls, quant_cls):
        if float_cls in cls._registry:
            raise ValueError(f"Module: {float_cls.__name__} has been registered.")
        cls._registry[float_cls] = quant_cls

    @classmethod
    def quantize(cls, model: Module, fold_bn: bool = True, block_scale: bool | None = None) -> Module:
        """Returns a quantised deep copy in evaluate mode (AbstractModule.quantize clones first, S/nn/abstractnn/
        AbstractModule.scala:920-928)."""
        from ..nn.fusion import fold_bn as do_fold, fuse, unfuse

        from ..nn.flat import unflatten_parameters

        m = model.cloneModule().evaluate()
        dev = m.device()
        unflatten_parameters(m)
        # a model that already lived on the GPU carries fusion flags (e.g. BN+ReLU); they must be recomputed AFTER the BN
        # folding, otherwise a folded BN would still apply "its" ReLU — a no-op on e4m3 bytes handed over by a conv
        unfuse(m)
        if fold_bn:
            do_fold(m)
        fuse(m)
        _fuse_through_identities(m)
        out = cls._convert(m)
        out.evaluate()
        if block_scale is None:
            from ..utils import config

            block_scale = bool(config.get("bigdl.quantized.blockScale"))
        if block_scale:
            # MX block scaling for the fully-connected layers (tcgen05 kind::mxf8f6f4.block_scale): per-32-element
            # power-of-two scales computed on the fly, so these layers need neither abs-max passes nor calibration
            from .layers import Linear as QLinear

            for l in cls.quantized_layers(out):
                if isinstance(l, QLinear):
                    l.enable_block_scale()
        for l in cls.quantized_layers(out):
            l.__dict__.pop("_w_float", None)
        if any(l.static_scale for l in cls.quantized_layers(out)):
            cls.link_chains(out)
        if dev.type == "cuda" and out.torch_parameters():
            out.getParameters()
        return out

    @classmethod
    def _convert(cls, m: Module) -> Module:
        q = cls._registry.get(type(m))
        if q is not None:
            qm = q.from_float(m)
            st = m.__dict__.get("_mkl_int8")          # scales recorded by calcScales (MklInt8Convertible)
            if st and st.get("inputScales") and st["inputScales"][0]:
                qm.setInputScales([[max(st["inputScales"][0])]])
            return qm
        if isinstance(m, Graph):
            return GraphQuantizer.quantize(m)
        if isinstance(m, Container):
            return ContainerQuantizer.quantize(m)
        if isinstance(m.__dict__.get("cell"), Module):           # Recurrent-style wrappers own a single cell
            return CellQuantizer.quantize(m)
        return ModuleQuantizer.quantize(m)

    # ---------------------------------------------------------------- calibration
    @staticmethod
    def quantized_layers(model: Module) -> list[QL.QuantizedModule]:
        return [m for m in model.all_modules() if isinstance(m, QL.QuantizedModule)]

    @classmethod
    def calibrate(cls, model: Module, batches, chain: bool = True) -> Module:
        """Run ``batches`` (iterable of inputs) through the quantised model recording input abs-max per layer, then
        switch every layer to static scales and link requantising epilogues."""
        qs = cls.quantized_layers(model)
        for q in qs:
            q._buffers["input_absmax"].zero_()
            q.calibrating, q._emit_q = True, None
        for b in batches:
            model.forward(b)
        for q in qs:
            q.calibrating, q.static_scale, q._colscale = False, True, None
            q.__dict__.pop("_in_qscale", None)
        if chain:
            cls.link_chains(model)
        return model

    @staticmethod
    def link_chains(model: Module) -> int:
        """producer._emit_q = consumer wherever the producer's only user is a statically-scaled quantised layer whose
        input channel count needs no padding (so the e4m3 tensor can be consumed in place). Returns the link count."""
        n = 0
        for a, b in _effective_pairs(model):
            if isinstance(a, QL.QuantizedModule) and isinstance(b, QL.QuantizedModule) and b.static_scale \
                    and not getattr(b, "block_scale", False) and not getattr(a, "block_scale", False):
                if isinstance(b, QL.SpatialConvolution):
                    ok = isinstance(a, QL.SpatialConvolution) and a.n_output_plane == b.n_input_plane and \
                        b.n_input_plane % 16 == 0 and b.n_group == 1 and b.data_format == a.data_format
                else:
                    ok = isinstance(a, QL.Linear) and a.output_size == b.input_size and b.input_size % 16 == 0
                if ok:
                    a._emit_q = b
                    n += 1
        # branches of a ConcatTable read the same tensor: give their first quantised layers ONE scale buffer so the
        # second branch reuses the first one's quantised copy (layers.py caches it on the tensor)
        from ..nn.layers_table import ConcatTable

        for ct in model.all_modules():
            if not isinstance(ct, ConcatTable):
                continue
            firsts = []
            for br in ct.modules:
                f = br
                while isinstance(f, Sequential):
                    mods = [c for c in f.modules if not _is_identity(c)]
                    f = mods[0] if mods else None
                if isinstance(f, QL.SpatialConvolution) and f.static_scale:
                    firsts.append(f)
            if len(firsts) >= 2:
                shared = torch.stack([f._buffers["input_absmax"] for f in firsts]).amax(dim=0)
                for f in firsts:
                    f._buffers["input_absmax"] = shared
                    f._colscale = None
                    f.__dict__.pop("_in_qscale", None)
        return n


def _is_identity(m: Module) -> bool:
    return (type(m) is LC.ReLU and getattr(m, "_fused", False)) or \
        (isinstance(m, LC.BatchNormalization) and getattr(m, "_folded", False)) or type(m).__name__ == "Identity"


def _effective_pairs(model: Module):
    """(producer, consumer) pairs with single-use edges, looking through modules that became identities."""
    for m in model.all_modules():
        if isinstance(m, Sequential):
            mods = [c for c in m.modules if not _is_identity(c)]
            for a, b in zip(mods, mods[1:]):
                yield a, b
        elif isinstance(m, Graph):
            for n in m.exec_order:
                if _is_identity(n.element) or n in m.output_nodes or len(n.next) != 1:
                    continue
                nxt = n.next[0]
                while _is_identity(nxt.element) and len(nxt.next) == 1 and len(nxt.prev) == 1 and \
                        nxt not in m.output_nodes:
                    nxt = nxt.next[0]
                if len(nxt.prev) == 1 and not _is_identity(nxt.element):
                    yield n.element, nxt.element


def _fuse_through_identities(model: Module):
    """After conv+BN folding the BN is an identity, so a ReLU behind it can still go into the conv epilogue."""
    for a, b in _effective_pairs(model):
        if type(b) is LC.ReLU and not getattr(b, "_fused", False) and \
                isinstance(a, (LC.SpatialConvolution, LC.Linear)) and not a._fuse_relu:
            a._fuse_relu, b._fused = True, True


Quantizer.register(LC.Linear, QL.Linear)
Quantizer.register(LC.SpatialConvolution, QL.SpatialConvolution)
Quantizer.register(LC.SpatialDilatedConvolution, QL.SpatialDilatedConvolution)


# ------------------------------------------------------------------------------------------------ per-kind quantisers
class Quantizable:
    """What a layer's companion implements to take part in ``Quantizer.quantize`` (S/nn/quantized/Quantizable.scala):
    ``quantize(module)`` returns the module to use in its place. Register an implementation for your own layer class
    with ``Quantizer.register(FloatLayer, QuantLayer)`` where ``QuantLayer.from_float(layer)`` builds the replacement."""

    @staticmethod
    def quantize(module: Module) -> Module:  # pragma: no cover - interface
        raise NotImplementedError


class ModuleQuantizer(Quantizable):
    """Leaf layers without a registered quantised twin stay as they are."""

    @staticmethod
    def quantize(module: Module) -> Module:
        return module


class ContainerQuantizer(Quantizable):
    """Replace every child in place."""

    @staticmethod
    def quantize(module: Module) -> Module:
        module.modules = [Quantizer._convert(c) for c in module.modules]
        if hasattr(module, "_after_children_replaced"):
            module._after_children_replaced()
        return module


class GraphQuantizer(Quantizable):
    """Replace the element of every node, then rebuild the execution order."""

    @staticmethod
    def quantize(module: Module) -> Module:
        for n in module.exec_order:
            new = Quantizer._convert(n.element)
            if new is not n.element:
                n.element = new
        module._build()
        return module


class CellQuantizer(Quantizable):
    """Wrappers that own one inner ``cell`` module."""

    @staticmethod
    def quantize(module: Module) -> Module:
        cell = module.__dict__.get("cell")
        if isinstance(cell, Module):
            module.cell = Quantizer._convert(cell)
        return module
