"""Node-level ONNX operator converters with the reference's calling convention (P/contrib/onnx/{ops_converter,
ops_mapping,converter_utils}.py): ``convert(inputs, prev_modules, attrs, outputs) -> (graph node, [output shapes])``
where ``inputs`` is a list of ``(value or None, shape)`` pairs and ``prev_modules`` the predecessor graph nodes.
``interop/onnx_loader.py`` is the loader the framework itself uses (a superset of operators, no shape bookkeeping);
these functions keep user code written against the reference's extension point (``_convert_map``) working."""
from __future__ import annotations

import math

import numpy as np
import torch


# ------------------------------------------------------------------ converter_utils
def calc_output_shape(input, kernel, padding=0, stride=1, dilation=1, ceil_mode=False):
    eff = kernel + (kernel - 1) * (dilation - 1)
    size = (input + 2 * padding - eff) / stride + 1
    return int(math.ceil(size) if ceil_mode else math.floor(size))


def parse_node_attr(node_proto) -> dict:
    """Attribute name -> python value (scalars, tuples for repeated fields, the raw message for tensors / graphs)."""
    out = {}
    for a in node_proto.attribute:
        if len(a.floats) or len(a.ints) or len(a.strings):
            vals = list(a.floats) or list(a.ints) or [s.decode("utf-8") if isinstance(s, bytes) else s for s in a.strings]
            out[a.name] = tuple(vals)
        elif a.HasField("t"):
            out[a.name] = a.t
        elif hasattr(a, "g") and a.HasField("g"):
            out[a.name] = a.g
        elif a.HasField("s"):
            out[a.name] = a.s.decode("utf-8") if isinstance(a.s, bytes) else a.s
        elif a.HasField("f"):
            out[a.name] = a.f
        elif a.HasField("i"):
            out[a.name] = a.i
        else:
            raise ValueError(f"Cannot parse attribute: {a.name}")
    return out


def parse_tensor_data(tensor_proto) -> np.ndarray:
    from . import onnx_pb as P

    arr = np.asarray(P.to_numpy(tensor_proto))
    return arr.reshape(tuple(tensor_proto.dims)) if len(tensor_proto.dims) else arr.reshape(1)


# ------------------------------------------------------------------ ops_converter
def _t(v):
    return None if v is None else torch.as_tensor(np.asarray(v))


def _set(module, **tensors):
    with torch.no_grad():
        for name, val in tensors.items():
            if val is not None and name in module._parameters:
                module._parameters[name].copy_(_t(val).reshape(module._parameters[name].shape).float())
            elif val is not None and name in module._buffers:
                module._buffers[name].copy_(_t(val).reshape(module._buffers[name].shape).float())
    return module


def _pool_geometry(attrs):
    kw, kh = (int(v) for v in attrs.get("kernel_shape", (1, 1))[:2])
    sw, sh = (int(v) for v in attrs.get("strides", (1, 1))[:2])
    pw, ph = (int(v) for v in attrs.get("pads", (0, 0))[:2])
    return kw, kh, sw, sh, pw, ph


def average_pool(inputs, prev_modules, attrs, outputs):
    from ..nn import layer as L
    _, shape = inputs[0]
    kw, kh, sw, sh, pw, ph = _pool_geometry(attrs)
    ceil = bool(attrs.get("ceil_mode", 0))
    m = L.SpatialAveragePooling(kw, kh, sw, sh, pw, ph, ceil_mode=ceil, count_include_pad=bool(attrs.get("count_include_pad", 0)))
    out = (shape[0], shape[1], calc_output_shape(shape[-2], kh, ph, sh, ceil_mode=ceil),
           calc_output_shape(shape[-1], kw, pw, sw, ceil_mode=ceil))
    return m(prev_modules[0]), [out]


def max_pool(inputs, prev_modules, attrs, outputs):
    from ..nn import layer as L
    _, shape = inputs[0]
    kw, kh, sw, sh, pw, ph = _pool_geometry(attrs)
    ceil = bool(attrs.get("ceil_mode", 0))
    m = L.SpatialMaxPooling(kw, kh, sw, sh, pw, ph)
    if ceil:
        m.ceil()
    out = (shape[0], shape[1], calc_output_shape(shape[-2], kh, ph, sh, ceil_mode=ceil),
           calc_output_shape(shape[-1], kw, pw, sw, ceil_mode=ceil))
    return m(prev_modules[0]), [out]


def batch_norm(inputs, prev_modules, attrs, outputs):
    from ..nn import layer as L
    _, shape = inputs[0]
    scale, bias, mean, var = (inputs[i][0] for i in range(1, 5))
    m = L.SpatialBatchNormalization(int(shape[1]), float(attrs.get("epsilon", 1e-5)), float(attrs.get("momentum", 0.9)))
    _set(m, weight=scale, bias=bias, runningMean=mean, runningVar=var)
    return m(prev_modules[0]), [shape]


def concat(inputs, prev_modules, attrs, outputs):
    from ..nn import layer as L
    axis = int(attrs.get("axis", 1))
    shapes = [s for _, s in inputs]
    out = list(shapes[0])
    out[axis] = sum(int(s[axis]) for s in shapes)
    return L.JoinTable(axis + 1, len(out))(*prev_modules), [tuple(out)]


def constant(inputs, prev_modules, attrs, outputs):
    from ..nn import tf as T
    val = attrs["value"]
    arr = parse_tensor_data(val) if not isinstance(val, np.ndarray) else val
    node = T.Const(torch.as_tensor(arr))()
    return node, [tuple(arr.shape)]


def conv(inputs, prev_modules, attrs, outputs):
    from ..nn import layer as L
    _, shape = inputs[0]
    w, wshape = inputs[1]
    b = inputs[2][0] if len(inputs) > 2 and inputs[2] is not None else None
    pw, ph = (int(v) for v in attrs.get("pads", (0, 0))[:2])
    kw, kh = (int(v) for v in attrs.get("kernel_shape", tuple(wshape[-2:]))[:2])
    sw, sh = (int(v) for v in attrs.get("strides", (1, 1))[:2])
    group = int(attrs.get("group", 1))
    m = L.SpatialConvolution(int(shape[1]), int(wshape[0]), kw, kh, sw, sh, pw, ph, n_group=group, with_bias=b is not None)
    _set(m, weight=w, bias=b)
    out = (shape[0], wshape[0], calc_output_shape(shape[-2], kh, ph, sh), calc_output_shape(shape[-1], kw, pw, sw))
    return m(prev_modules[0]), [out]


def gather(inputs, prev_modules, attrs, outputs):
    from ..nn import ops as O
    _, shape = inputs[0]
    _, ishape = inputs[1]
    axis = int(attrs.get("axis", 0))
    out = tuple(shape[:axis]) + tuple(ishape) + tuple(shape[axis + 1:])
    return O.Gather(axis)(*prev_modules), [out]


def gemm(inputs, prev_modules, attrs, outputs):
    from ..nn import onnx as OX
    _, ashape = inputs[0]
    b, bshape = inputs[1]
    c = inputs[2][0] if len(inputs) > 2 else None
    ta, tb = int(attrs.get("transA", 0)), int(attrs.get("transB", 0))
    m = OX.Gemm(float(attrs.get("alpha", 1.0)), float(attrs.get("beta", 1.0)), ta, tb, _t(b), _t(c))
    rows = ashape[1] if ta else ashape[0]
    cols = bshape[0] if tb else bshape[1]
    return m(prev_modules[0]), [(rows, cols)]


def relu(inputs, prev_modules, attrs, outputs):
    from ..nn import layer as L
    return L.ReLU()(prev_modules[0]), [inputs[0][1]]


def reshape(inputs, prev_modules, attrs, outputs):
    from ..nn import onnx as OX
    target = inputs[1][0] if len(inputs) > 1 else attrs.get("shape")
    dims = None if target is None else [int(v) for v in np.asarray(target).reshape(-1)]
    return OX.Reshape(dims)(prev_modules[0]), [tuple(dims) if dims is not None else None]


def shape(inputs, prev_modules, attrs, outputs):
    from ..nn import onnx as OX
    return OX.Shape()(prev_modules[0]), [(len(inputs[0][1]),)]


def softmax(inputs, prev_modules, attrs, outputs):
    from ..nn import layer as L
    return L.SoftMax()(prev_modules[0]), [inputs[0][1]]


def _sum(inputs, prev_modules, attrs, outputs):
    from ..nn import layer as L
    return L.CAddTable()(*prev_modules), [inputs[0][1]]


def unsqueeze(inputs, prev_modules, attrs, outputs):
    from ..nn import layer as L
    axes = [int(a) for a in attrs.get("axes", (0,))]
    out = list(inputs[0][1])
    node = prev_modules[0]
    for a in sorted(axes):
        out.insert(a, 1)
        node = L.Unsqueeze(a + 1)(node)
    return node, [tuple(out)]


# ------------------------------------------------------------------ ops_mapping
_convert_map = {
    "Constant": constant, "Conv": conv, "BatchNormalization": batch_norm, "Relu": relu, "MaxPool": max_pool,
    "AveragePool": average_pool, "Sum": _sum, "Reshape": reshape, "Gemm": gemm, "Softmax": softmax, "Shape": shape,
    "Gather": gather, "Unsqueeze": unsqueeze, "Concat": concat,
}
