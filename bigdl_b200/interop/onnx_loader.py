"""ONNX model import (reference: P/contrib/onnx/{onnx_loader,ops_converter,ops_mapping,converter_utils}.py and
S/python/api/PythonBigDLOnnx.scala; the reference maps 14 operators — the list here is a superset).

``load(path)`` / ``load_model_proto(bytes)`` → Graph. Initializers that feed weight slots (Conv W/B, Gemm B/C,
BatchNormalization scale/B/mean/var, MatMul B) become trainable ``Variable``s or buffers; other initializers and
``Constant`` nodes become constants. Layout is ONNX's NCHW, which is also the framework's logical layout."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from ..nn import onnx as OX
from ..nn import tf as TFO
from ..nn.module import Graph, Identity, Node
from ..utils.table import Table
from . import onnx_pb as P
from .tensorflow import _Fn, _ints

_OPS: dict = {}


def op(*names):
    def deco(fn):
        for n in names:
            _OPS[n] = fn
        return fn
    return deco


def _attrs(node) -> dict:
    out = {}
    for a in node.attribute:
        if a.HasField("t"):
            out[a.name] = P.to_numpy(a.t)
        elif len(a.ints):
            out[a.name] = [int(v) for v in a.ints]
        elif len(a.floats):
            out[a.name] = [float(v) for v in a.floats]
        elif a.HasField("s"):
            out[a.name] = a.s.decode()
        elif a.HasField("f"):
            out[a.name] = float(a.f)
        elif a.HasField("i"):
            out[a.name] = int(a.i)
        else:
            out[a.name] = []
    return out


def _pads2d(a, k):
    pads = a.get("pads", [0, 0, 0, 0])
    auto = a.get("auto_pad", "NOTSET")
    return pads, auto


def _apply_pads(x, pads, value=0.0):
    if any(pads):
        n = len(pads) // 2
        flat = []
        for i in reversed(range(n)):
            flat += [pads[i], pads[i + n]]
        return F.pad(x, flat, value=value)
    return x


@op("Conv")
def _conv(a):
    s, d, g = a.get("strides", [1, 1]), a.get("dilations", [1, 1]), a.get("group", 1)
    pads, auto = _pads2d(a, None)

    def run(x, w, *b):
        if auto in ("SAME_UPPER", "SAME_LOWER"):
            from .tensorflow import _same_pad

            x = F.pad(x, _same_pad(x, w.shape[2:], s, d))
        else:
            x = _apply_pads(x, pads)
        fn = F.conv2d if w.dim() == 4 else (F.conv1d if w.dim() == 3 else F.conv3d)
        return fn(x, w.to(x.dtype), b[0].to(x.dtype) if b else None, s, 0, d, g)
    return _Fn(run)


@op("ConvTranspose")
def _convt(a):
    s, g = a.get("strides", [1, 1]), a.get("group", 1)
    pads, op_ = a.get("pads", [0, 0, 0, 0]), a.get("output_padding", [0, 0])
    return _Fn(lambda x, w, *b: F.conv_transpose2d(x, w, b[0] if b else None, s, (pads[0], pads[1]), tuple(op_), g))


def _pool(kind):
    def load(a):
        k, s = a["kernel_shape"], a.get("strides", [1] * len(a["kernel_shape"]))
        pads, ceil = a.get("pads", [0] * 2 * len(k)), bool(a.get("ceil_mode", 0))
        incl = bool(a.get("count_include_pad", 0))

        def run(x):
            if kind == "max":
                return F.max_pool2d(_apply_pads(x, pads, float("-inf")), k, s, 0, 1, ceil)
            if any(pads) and not incl:
                ones = _apply_pads(torch.ones_like(x[:1, :1]), pads)
                return F.avg_pool2d(_apply_pads(x, pads), k, s, 0, ceil) / F.avg_pool2d(ones, k, s, 0, ceil)
            return F.avg_pool2d(_apply_pads(x, pads), k, s, 0, ceil)
        return _Fn(run, 1)
    return load


_OPS["MaxPool"], _OPS["AveragePool"] = _pool("max"), _pool("avg")


@op("GlobalAveragePool")
def _gap(a): return _Fn(lambda x: x.mean(tuple(range(2, x.dim())), keepdim=True), 1)


@op("GlobalMaxPool")
def _gmp(a): return _Fn(lambda x: x.amax(tuple(range(2, x.dim())), keepdim=True), 1)


@op("BatchNormalization")
def _bn(a):
    eps = a.get("epsilon", 1e-5)
    return _Fn(lambda x, sc, b, m, v: F.batch_norm(x, m, v, sc, b, False, 0.0, eps), 5)


@op("Gemm")
def _gemm(a):
    return OX.Gemm(a.get("alpha", 1.0), a.get("beta", 1.0), bool(a.get("transA", 0)), bool(a.get("transB", 0)))


@op("MatMul")
def _matmul(a): return _Fn(lambda x, y: torch.matmul(x, y), 2)


for _n, _f in {"Relu": torch.relu, "Sigmoid": torch.sigmoid, "Tanh": torch.tanh, "Exp": torch.exp, "Log": torch.log,
               "Sqrt": torch.sqrt, "Abs": torch.abs, "Neg": torch.neg, "Floor": torch.floor, "Ceil": torch.ceil,
               "Identity": lambda x: x, "Softplus": F.softplus, "Softsign": F.softsign, "Erf": torch.erf,
               "Reciprocal": torch.reciprocal, "Not": torch.logical_not}.items():
    _OPS[_n] = (lambda f: (lambda a: _Fn(f, 1)))(_f)

for _n, _f in {"Add": torch.add, "Sub": torch.sub, "Mul": torch.mul, "Div": torch.div, "Pow": torch.pow,
               "Greater": torch.gt, "Less": torch.lt, "Equal": torch.eq, "And": torch.logical_and,
               "Or": torch.logical_or}.items():
    _OPS[_n] = (lambda f: (lambda a: _Fn(f, 2)))(_f)


@op("Sum")
def _sum(a): return _Fn(lambda *xs: sum(xs[1:], xs[0]))


@op("Max")
def _max(a): return _Fn(lambda *xs: torch.stack(torch.broadcast_tensors(*xs)).amax(0))


@op("Min")
def _min(a): return _Fn(lambda *xs: torch.stack(torch.broadcast_tensors(*xs)).amin(0))


@op("Mean")
def _mean(a): return _Fn(lambda *xs: torch.stack(torch.broadcast_tensors(*xs)).mean(0))


@op("LeakyRelu")
def _lrelu(a): return _Fn(lambda x: F.leaky_relu(x, a.get("alpha", 0.01)), 1)


@op("Elu")
def _elu(a): return _Fn(lambda x: F.elu(x, a.get("alpha", 1.0)), 1)


@op("PRelu")
def _prelu(a): return _Fn(lambda x, s: torch.where(x >= 0, x, x * s), 2)


@op("Clip")
def _clip(a):
    return _Fn(lambda x, *mm: torch.clamp(x, float(mm[0]) if mm else a.get("min"), float(mm[1]) if len(mm) > 1 else a.get("max")))


@op("Softmax")
def _softmax(a): return _Fn(lambda x: torch.softmax(x, a.get("axis", -1)), 1)


@op("LogSoftmax")
def _logsoftmax(a): return _Fn(lambda x: torch.log_softmax(x, a.get("axis", -1)), 1)


@op("LRN")
def _lrn(a):
    return _Fn(lambda x: F.local_response_norm(x, a["size"], a.get("alpha", 1e-4), a.get("beta", 0.75), a.get("bias", 1.0)), 1)


@op("Dropout")
def _dropout(a): return Identity()


@op("Flatten")
def _flatten(a):
    ax = a.get("axis", 1)
    return _Fn(lambda x: x.reshape(int(np.prod(x.shape[:ax])) if ax else 1, -1), 1)


@op("Reshape")
def _reshape(a): return OX.Reshape(a.get("shape"))


@op("Shape")
def _shape(a): return OX.Shape()


@op("Transpose")
def _transpose(a): return _Fn(lambda x: x.permute(a["perm"]) if "perm" in a else x.permute(*reversed(range(x.dim()))), 1)


@op("Concat")
def _concat(a): return _Fn(lambda *xs: torch.cat(list(xs), a.get("axis", 1)))


@op("Unsqueeze")
def _unsq(a):
    def run(x, *axes):
        ax = _ints(axes[0]) if axes else a["axes"]
        for d in sorted(ax):
            x = x.unsqueeze(d)
        return x
    return _Fn(run)


@op("Squeeze")
def _sq(a):
    def run(x, *axes):
        ax = _ints(axes[0]) if axes else a.get("axes")
        if ax is None:
            return x.squeeze()
        for d in sorted((v % x.dim() for v in ax), reverse=True):
            x = x.squeeze(d)
        return x
    return _Fn(run)


@op("Gather")
def _gather(a):
    ax = a.get("axis", 0)
    return _Fn(lambda x, i: torch.index_select(x, ax, i.long().reshape(-1)).reshape(
        x.shape[:ax] + tuple(i.shape) + x.shape[ax + 1:]), 2)


@op("Cast")
def _cast(a):
    dt = {1: torch.float32, 6: torch.int32, 7: torch.int64, 9: torch.bool, 10: torch.float16, 11: torch.float64}[a["to"]]
    return _Fn(lambda x: x.to(dt), 1)


@op("Pad")
def _pad(a):
    def run(x, *extra):
        pads = _ints(extra[0]) if extra else a["pads"]
        value = float(extra[1]) if len(extra) > 1 else a.get("value", 0.0)
        return _apply_pads(x, pads, value)
    return _Fn(run)


def _reduce(fn):
    def load(a):
        keep = bool(a.get("keepdims", 1))

        def run(x, *axes):
            ax = _ints(axes[0]) if axes else a.get("axes", list(range(x.dim())))
            return fn(x, tuple(ax), keep)
        return _Fn(run)
    return load


_OPS["ReduceMean"] = _reduce(lambda x, ax, k: x.mean(ax, keepdim=k))
_OPS["ReduceSum"] = _reduce(lambda x, ax, k: x.sum(ax, keepdim=k))
_OPS["ReduceMax"] = _reduce(lambda x, ax, k: x.amax(ax, keepdim=k))
_OPS["ReduceMin"] = _reduce(lambda x, ax, k: x.amin(ax, keepdim=k))

SUPPORTED_OPS = sorted(list(_OPS) + ["Constant"])
_TRAINABLE_SLOTS = {"Conv": [1, 2], "ConvTranspose": [1, 2], "Gemm": [1, 2], "MatMul": [1], "BatchNormalization": [1, 2],
                    "PRelu": [1]}


def load_model_proto(data: bytes) -> Graph:
    model = P.ModelProto()
    model.ParseFromString(data)
    g = model.graph
    inits = {t.name: P.to_numpy(t) for t in g.initializer}
    trainable = set()
    for n in g.node:
        for slot in _TRAINABLE_SLOTS.get(n.op_type, []):
            if slot < len(n.input) and n.input[slot] in inits:
                trainable.add(n.input[slot])
    producers: dict[str, tuple[Node, int | None]] = {}
    in_nodes = []
    for vi in g.input:
        if vi.name in inits:
            continue
        node = Node(Identity().setName(vi.name), [])
        producers[vi.name] = (node, None)
        in_nodes.append(node)
    for name, arr in inits.items():
        if name in trainable and arr.dtype.kind == "f":
            m = TFO.Variable(torch.from_numpy(arr.astype(np.float32)))
        else:
            m = TFO.Const(torch.from_numpy(arr))
        m.setName(name)
        producers[name] = (Node(m, []), None)
    for i, n in enumerate(g.node):
        a = _attrs(n)
        if n.op_type == "Constant":
            m = TFO.Const(torch.from_numpy(np.asarray(a["value"])))
        else:
            loader = _OPS.get(n.op_type)
            if loader is None:
                raise ValueError(f"unsupported ONNX operator {n.op_type} (node {n.name or i})")
            m = loader(a)
        m.setName(n.name or f"{n.op_type}_{i}")
        node = Node(m, [])
        for inp in n.input:
            if inp == "":
                continue
            if inp not in producers:
                raise ValueError(f"ONNX graph is not topologically sorted: {inp} used before definition")
            src, idx = producers[inp]
            src.add(node, idx)
        multi = len(n.output) > 1 and n.op_type not in ("Dropout", "BatchNormalization", "MaxPool")
        for k, out in enumerate(n.output):
            producers[out] = (node, (k + 1) if multi else None)
    out_nodes = [producers[o.name][0] for o in g.output]
    return Graph(in_nodes, out_nodes)


def load(path: str) -> Graph:
    with open(path, "rb") as f:
        return load_model_proto(f.read())


load_model = load
