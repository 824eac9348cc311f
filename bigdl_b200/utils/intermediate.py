"""Intermediate representation between engines (reference: S/utils/intermediate/{IRGraph,IRElement,BlasToIR,IRToBlas,
IRToDnn,IRConverter,ConversionUtils}.scala, 1,325 LoC: a blas model is lowered to an engine-neutral IR graph and
re-emitted as dnn layers when ``bigdl.engineType=mkldnn``; ``ConversionUtils.convert`` is called by the optimizers,
predictors and evaluators before they run a model).

Here both "engines" are the same layer classes, so the IR is a light, serialisable description of the graph — op type +
constructor arguments + edges — that is useful on its own (export, inspection, rebuilding a model on another device or
with different fusion settings). ``ConversionUtils.convert`` does what the reference's does at the same call sites:
prepare the model for the active engine (GPU placement, fusion, optional quantisation switch)."""
from __future__ import annotations

from dataclasses import dataclass, field

import torch

from . import config


@dataclass
class IRElement:
    name: str
    op: str                         # layer class name, e.g. "SpatialConvolution"
    attrs: dict = field(default_factory=dict)
    weights: dict = field(default_factory=dict)     # parameter name → tensor (shared with the source model)

    def getName(self):
        return self.name

    def getOp(self):
        return self.op


@dataclass
class IRNode:
    element: IRElement
    prev: list = field(default_factory=list)        # indices into IRGraph.nodes
    edge_index: list = field(default_factory=list)


class IRGraph:
    """Topologically ordered nodes + input/output node indices (IRGraph.scala:30-220)."""

    def __init__(self, nodes, inputs, outputs, name="graph"):
        self.nodes, self.inputs, self.outputs, self.name = nodes, inputs, outputs, name

    def __len__(self):
        return len(self.nodes)

    def summary(self) -> str:
        lines = []
        for i, n in enumerate(self.nodes):
            lines.append(f"{i:3d} {n.element.op:28s} {n.element.name:32s} <- {n.prev}")
        return "\n".join(lines)


class BlasToIR:
    """Module tree → IRGraph. Sequential chains are flattened; Graph edges are kept; any other container (Concat,
    ConcatTable, Recurrent, …) stays one IR node holding the whole sub-module, as BlasToIR does for unsupported layers."""

    @staticmethod
    def convert(model) -> IRGraph:
        from ..nn.module import Graph, Sequential

        nodes: list[IRNode] = []

        def element(m):
            attrs = dict(m.__dict__.get("_ctor_args", {}) or {})
            return IRElement(m.getName(), type(m).__name__, attrs,
                             {k: v for k, v in m._parameters.items()} if not m.modules else {"__module__": m})

        def add(m, prev, edge_index=None):
            if isinstance(m, Sequential) and type(m).__name__ == "Sequential":
                cur = prev
                for c in m.modules:
                    cur = [add(c, cur)]
                return cur[0]
            if m.modules:
                el = IRElement(m.getName(), type(m).__name__, {}, {"__module__": m})
            else:
                el = element(m)
                el.weights["__module__"] = m
            nodes.append(IRNode(el, list(prev), list(edge_index or [None] * len(prev))))
            return len(nodes) - 1

        if isinstance(model, Graph):
            idx = {}
            for n in model.exec_order:
                idx[id(n)] = add(n.element, [idx[id(p)] for p in n.prev], list(n.edge_index))
            return IRGraph(nodes, [idx[id(n)] for n in model.input_nodes], [idx[id(n)] for n in model.output_nodes],
                           model.getName())
        last = add(model, [])
        first = 0
        return IRGraph(nodes, [first], [last], model.getName())


class IRToBlas:
    """IRGraph → Graph model reusing the source modules (and therefore their weights)."""

    @staticmethod
    def convert(ir: IRGraph):
        from ..nn.module import Graph, Input, Node

        made: list = []
        entry = None
        for i, n in enumerate(ir.nodes):
            m = n.element.weights.get("__module__")
            if m is None:
                raise ValueError(f"IR node {n.element.name} has no source module to rebuild from")
            node = Node(m, [])
            for p, ei in zip(n.prev, n.edge_index):
                made[p].add(node, ei)
            made.append(node)
        ins = [made[i] for i in ir.inputs]
        # a flattened Sequential has no explicit Input node: prepend one
        fixed_ins = []
        for n in ins:
            if type(n.element).__name__ != "Input":
                inp = Input()
                inp.add(n)
                fixed_ins.append(inp)
            else:
                fixed_ins.append(n)
        return Graph(fixed_ins, [made[i] for i in ir.outputs])


class IRToDnn(IRToBlas):
    """IRGraph → DnnGraph (compile/fusion capable)."""

    @staticmethod
    def convert(ir: IRGraph):
        from ..nn.mkldnn import DnnGraph

        g = IRToBlas.convert(ir)
        return DnnGraph(g.input_nodes, g.output_nodes)


class ConversionUtils:
    @staticmethod
    def convert(model, device=None):
        """Prepare ``model`` for the active engine (ConversionUtils.scala:31-85): with ``bigdl.engineType=b200`` and a GPU
        present the model is placed on it (flat parameters, bf16 compute copy) and the fusion pass runs; with
        ``bigdl.quantize`` set (the reference's ``bigdl.mkldnn.int8`` switch) an inference model is quantised first."""
        dev = torch.device(device) if device is not None else (
            torch.device("cuda") if torch.cuda.is_available() and config.get("bigdl.engineType") == "b200"
            else torch.device("cpu"))
        if config.get("bigdl.quantize", False) and not model.isTraining():
            model = model.quantize()
        if model.device() != dev:
            model = model.to(dev)
        if dev.type == "cuda":
            from ..nn.fusion import fuse

            fuse(model)
        return model

    @staticmethod
    def convertBack(model):
        """Undo engine-specific preparation: fusion flags cleared (weights are shared, so nothing to copy back)."""
        from ..nn.fusion import unfuse

        return unfuse(model)


# ------------------------------------------------------------------------------------------------ typed operators
class IROperator:
    """Engine-neutral description of ONE layer (IRElement.scala:27-140): its kind is the class, its settings are the
    attributes (the constructor arguments of the source layer). ``IRElement.operator()`` builds it from an element."""

    layer: str | None = None                      # the nn.layer class this operator stands for

    def __init__(self, **attrs):
        self.__dict__.update(attrs)

    def name(self) -> str:
        return type(self).__name__

    def attrs(self) -> dict:
        return dict(self.__dict__)

    def __eq__(self, o):
        return type(o) is type(self) and {k: v for k, v in o.__dict__.items() if not hasattr(v, "shape")} == \
            {k: v for k, v in self.__dict__.items() if not hasattr(v, "shape")}

    def __hash__(self):
        return hash(type(self).__name__)

    def __repr__(self):
        return f"{type(self).__name__}({', '.join(f'{k}={v!r}' for k, v in self.__dict__.items() if not hasattr(v, 'shape'))})"


def _ir(name, layer):
    return type(name, (IROperator,), {"layer": layer, "__doc__": f"IR form of ``nn.{layer}``."})


IRSpatialMaxPooling = _ir("IRSpatialMaxPooling", "SpatialMaxPooling")
IRSpatialAveragePooling = _ir("IRSpatialAveragePooling", "SpatialAveragePooling")
IRSpatialConvolution = _ir("IRSpatialConvolution", "SpatialConvolution")
IRSpatialShareConvolution = _ir("IRSpatialShareConvolution", "SpatialShareConvolution")
IRSpatialBatchNormalization = _ir("IRSpatialBatchNormalization", "SpatialBatchNormalization")
IRIdentity = _ir("IRIdentity", "Identity")
IRDropout = _ir("IRDropout", "Dropout")
IRReLU = _ir("IRReLU", "ReLU")
IRLinear = _ir("IRLinear", "Linear")
IRSqueeze = _ir("IRSqueeze", "Squeeze")
IRSpatialCrossMapLRN = _ir("IRSpatialCrossMapLRN", "SpatialCrossMapLRN")
IRSoftMax = _ir("IRSoftMax", "SoftMax")
IRSelectTable = _ir("IRSelectTable", "SelectTable")
IRCAddTable = _ir("IRCAddTable", "CAddTable")
IRJoinTable = _ir("IRJoinTable", "JoinTable")
IRConcatTable = _ir("IRConcatTable", "ConcatTable")
IRInput = _ir("IRInput", "Input")


class IRGeneralModule(IROperator):
    """Any layer without a dedicated operator: carries the source module itself (IRElement.scala:139-140)."""

    def __init__(self, model=None, **attrs):
        super().__init__(**attrs)
        self.model = model


_IR_OF_LAYER = {c.layer: c for c in (IRSpatialMaxPooling, IRSpatialAveragePooling, IRSpatialConvolution,
                                     IRSpatialShareConvolution, IRSpatialBatchNormalization, IRIdentity, IRDropout,
                                     IRReLU, IRLinear, IRSqueeze, IRSpatialCrossMapLRN, IRSoftMax, IRSelectTable,
                                     IRCAddTable, IRJoinTable, IRConcatTable, IRInput)}


def _operator(self) -> IROperator:
    cls = _IR_OF_LAYER.get(self.op)
    if cls is None:
        return IRGeneralModule(self.weights.get("__module__"), **self.attrs)
    return cls(**{k: v for k, v in self.attrs.items()})


IRElement.operator = _operator


class ConvertBase:
    """Shared shape of the two directions (BlasToIR / IRToBlas / IRToDnn extend it in the reference):
    ``convertLayerCheck(layer)`` says whether a dedicated mapping exists, ``convertLayer(layer)`` performs it."""

    @staticmethod
    def convertLayerCheck(layer) -> bool:
        name = layer.op if isinstance(layer, IRElement) else type(layer).__name__
        return name in _IR_OF_LAYER

    @staticmethod
    def convertLayer(layer):
        if isinstance(layer, IRElement):
            return layer.weights.get("__module__")
        attrs = dict(layer.__dict__.get("_ctor_args", {}) or {})
        return IRElement(layer.getName(), type(layer).__name__, attrs, {"__module__": layer})


class IRConverter:
    """``IRConverter(irGraph).toGraph()`` (IRConverter.scala:31-140): emit the model for the active engine —
    a ``DnnGraph`` (fusion-capable) when ``bigdl.engineType`` is ``mkldnn`` / ``b200``, a plain ``Graph`` otherwise."""

    def __init__(self, IRgraph: IRGraph):
        self.ir = IRgraph

    def toGraph(self):
        engine = str(config.get("bigdl.engineType"))
        return IRToDnn.convert(self.ir) if engine in ("mkldnn", "b200") else IRToBlas.convert(self.ir)
