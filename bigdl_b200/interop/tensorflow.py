"""TensorFlow frozen-graph import and export.

Reference: S/utils/tf/TensorflowLoader.scala:55-400 (``load(pb, inputs, outputs, byteOrder, binFile)``: parse the
GraphDef, keep the sub-graph between the named inputs and outputs, convert it into a BigDL graph),
S/utils/tf/TensorflowToBigDL.scala (854 LoC of sub-graph patterns) plus 161 per-op loaders under S/utils/tf/loaders,
S/utils/tf/TensorflowSaver.scala + BigDLToTensorflow.scala (export), docs/docs/APIGuide/tensorflow_ops_list.md.

Design here: one small loader per TensorFlow op, registered in ``_LOADERS``; a loader receives the NodeDef and returns a
module whose inputs are the node's data inputs (as a Table when there are several). Weights reach layers as ordinary
inputs: a ``Const`` feeding a filter / kernel / bias / batch-norm slot becomes a trainable ``Variable`` module, so an
imported graph can be fine-tuned without pattern matching whole sub-graphs. Graphs containing Switch / Merge / Enter /
Exit / NextIteration run on ``DynamicGraph``. Tensors keep TensorFlow's layouts (NHWC unless the node says NCHW)."""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

from ..nn import ops as O
from ..nn import tf as TFO
from ..nn.module import Graph, Identity, Module, Node
from ..nn.ops import Operation, _ins
from ..utils.table import Table
from . import tf_pb as P

_LOADERS: dict = {}


def op(*names):
    def deco(fn):
        for n in names:
            _LOADERS[n] = fn
        return fn
    return deco


class _Fn(Module):
    """Differentiable module defined by a closure over its (tensor) inputs."""

    def __init__(self, fn, n_in: int | None = None, label: str = ""):
        super().__init__()
        self.fn, self.n_in, self.label = fn, n_in, label

    def updateOutput(self, input):
        args = _ins(input) if (isinstance(input, (Table, list, tuple)) or self.n_in not in (None, 1)) else [input]
        return self.fn(*args)


def _attr(node, name, default=None):
    if name not in node.attr:
        return default
    a = node.attr[name]
    if a.HasField("list"):
        v = a.list
        for lf in ("i", "f", "s", "b", "type"):
            if len(getattr(v, lf)):
                return list(getattr(v, lf))
        return []
    for f in ("s", "i", "f", "b", "type", "shape", "tensor"):
        if a.HasField(f):
            v = getattr(a, f)
            return v.decode() if f == "s" else v
    return default


def _ints(t):
    return [int(v) for v in torch.as_tensor(t).reshape(-1).tolist()]


def _nchw(node) -> bool:
    return _attr(node, "data_format", "NHWC") == "NCHW"


def _same_pad(x, k, s, d=(1, 1)):
    H, W = x.shape[2:]
    eh, ew = (k[0] - 1) * d[0] + 1, (k[1] - 1) * d[1] + 1
    th = max((math.ceil(H / s[0]) - 1) * s[0] + eh - H, 0)
    tw = max((math.ceil(W / s[1]) - 1) * s[1] + ew - W, 0)
    return (tw // 2, tw - tw // 2, th // 2, th - th // 2)


def _hw(vals, nchw):
    return (vals[2], vals[3]) if nchw else (vals[1], vals[2])


# ------------------------------------------------------------------------------------------------ op loaders
@op("Placeholder", "PlaceholderV2")
def _placeholder(n): return Identity()


@op("PlaceholderWithDefault", "Identity", "StopGradient", "PreventGradient", "Snapshot", "CheckNumerics")
def _identity(n):
    if n.op in ("StopGradient", "PreventGradient"):
        return _Fn(lambda x: x.detach(), 1)
    return Identity()


@op("Const")
def _const(n):
    return TFO.Const(_np_to_activity(P.tensor_to_numpy(n.attr["value"].tensor)))


def _np_to_activity(a):
    if isinstance(a, np.ndarray) and a.dtype != object:
        return torch.from_numpy(a)
    return a


@op("NoOp")
def _noop(n): return TFO.ControlDependency()


@op("Assert")
def _assert(n): return TFO.Assert()


def _bin(fn):
    return lambda n: _Fn(fn, 2)


for _name, _f in {"Add": torch.add, "AddV2": torch.add, "Sub": torch.sub, "Mul": torch.mul, "RealDiv": torch.div,
                  "Div": torch.div, "Maximum": torch.maximum, "Minimum": torch.minimum, "Pow": torch.pow,
                  "SquaredDifference": lambda a, b: (a - b) ** 2, "FloorDiv": torch.floor_divide,
                  "FloorMod": torch.remainder, "Mod": torch.fmod,
                  "TruncateDiv": lambda a, b: torch.div(a, b, rounding_mode="trunc"),
                  "Greater": torch.gt, "GreaterEqual": torch.ge, "Less": torch.lt, "LessEqual": torch.le,
                  "Equal": torch.eq, "NotEqual": torch.ne, "LogicalAnd": torch.logical_and,
                  "LogicalOr": torch.logical_or}.items():
    _LOADERS[_name] = _bin(_f)

for _name, _f in {"Neg": torch.neg, "Abs": torch.abs, "Sqrt": torch.sqrt, "Rsqrt": torch.rsqrt, "Square": torch.square,
                  "Exp": torch.exp, "Expm1": torch.expm1, "Log": torch.log, "Log1p": torch.log1p, "Floor": torch.floor,
                  "Ceil": torch.ceil, "Round": torch.round, "Rint": torch.round, "Sign": torch.sign, "Erf": torch.erf,
                  "Erfc": torch.erfc, "Lgamma": torch.lgamma, "Digamma": torch.digamma,
                  "Reciprocal": torch.reciprocal, "Inv": torch.reciprocal, "LogicalNot": torch.logical_not,
                  "IsNan": torch.isnan, "IsInf": torch.isinf, "IsFinite": torch.isfinite, "Relu": torch.relu,
                  "Relu6": lambda x: torch.clamp(x, 0, 6), "Elu": F.elu, "Sigmoid": torch.sigmoid, "Tanh": torch.tanh,
                  "Softplus": F.softplus, "Softsign": F.softsign, "Softmax": lambda x: torch.softmax(x, -1),
                  "LogSoftmax": lambda x: torch.log_softmax(x, -1), "L2Loss": lambda x: (x * x).sum() / 2,
                  "ZerosLike": torch.zeros_like, "OnesLike": torch.ones_like}.items():
    _LOADERS[_name] = (lambda f: (lambda n: _Fn(f, 1)))(_f)


@op("AddN")
def _addn(n): return _Fn(lambda *xs: sum(xs[1:], xs[0]))


@op("MatMul")
def _matmul(n):
    ta, tb = bool(_attr(n, "transpose_a", False)), bool(_attr(n, "transpose_b", False))
    return _Fn(lambda a, b: (a.t() if ta else a) @ (b.t() if tb else b), 2)


@op("BatchMatMul", "BatchMatMulV2")
def _bmm(n):
    return O.BatchMatMul(bool(_attr(n, "adj_x", False)), bool(_attr(n, "adj_y", False)))


@op("BiasAdd")
def _biasadd(n):
    if _nchw(n):
        return _Fn(lambda x, b: x + b.reshape([1, -1] + [1] * (x.dim() - 2)), 2)
    return TFO.BiasAdd()


@op("Conv2D")
def _conv2d(n):
    nchw = _nchw(n)
    s, d = _hw(_attr(n, "strides"), nchw), _hw(_attr(n, "dilations", [1, 1, 1, 1]), nchw)
    same = _attr(n, "padding") == "SAME"

    def run(x, w):                                  # w: [kh, kw, in, out]
        xc = x if nchw else x.permute(0, 3, 1, 2)
        wt = w.permute(3, 2, 0, 1)
        if same:
            xc = F.pad(xc, _same_pad(xc, wt.shape[2:], s, d))
        y = F.conv2d(xc, wt.to(xc.dtype), None, s, 0, d)
        return y if nchw else y.permute(0, 2, 3, 1)
    return _Fn(run, 2)


@op("DepthwiseConv2dNative")
def _dwconv(n):
    nchw = _nchw(n)
    s = _hw(_attr(n, "strides"), nchw)
    same = _attr(n, "padding") == "SAME"

    def run(x, w):                                  # w: [kh, kw, in, mult]
        xc = x if nchw else x.permute(0, 3, 1, 2)
        kh, kw, cin, mult = w.shape
        wt = w.permute(2, 3, 0, 1).reshape(cin * mult, 1, kh, kw)
        if same:
            xc = F.pad(xc, _same_pad(xc, (kh, kw), s))
        y = F.conv2d(xc, wt, None, s, 0, 1, cin)
        return y if nchw else y.permute(0, 2, 3, 1)
    return _Fn(run, 2)


@op("Conv2DBackpropInput")
def _deconv(n):
    nchw = _nchw(n)
    s = _hw(_attr(n, "strides"), nchw)
    same = _attr(n, "padding") == "SAME"

    def run(out_shape, w, x):                       # w: [kh, kw, out, in]
        xc = x if nchw else x.permute(0, 3, 1, 2)
        oshape = _ints(out_shape)
        oh, ow = (oshape[2], oshape[3]) if nchw else (oshape[1], oshape[2])
        kh, kw = w.shape[:2]
        wt = w.permute(3, 2, 0, 1)
        ph = max((xc.shape[2] - 1) * s[0] + kh - oh, 0) if same else 0
        pw = max((xc.shape[3] - 1) * s[1] + kw - ow, 0) if same else 0
        y = F.conv_transpose2d(xc, wt, None, s, 0)
        y = y[:, :, ph // 2:ph // 2 + oh, pw // 2:pw // 2 + ow]
        return y if nchw else y.permute(0, 2, 3, 1)
    return _Fn(run, 3)


def _pool(kind):
    def load(n):
        nchw = _nchw(n)
        k, s = _hw(_attr(n, "ksize"), nchw), _hw(_attr(n, "strides"), nchw)
        same = _attr(n, "padding") == "SAME"

        def run(x):
            xc = x if nchw else x.permute(0, 3, 1, 2)
            if kind == "max":
                if same:
                    xc = F.pad(xc, _same_pad(xc, k, s), value=float("-inf"))
                y = F.max_pool2d(xc, k, s)
            else:
                if same:
                    pad = _same_pad(xc, k, s)
                    ones = F.pad(torch.ones_like(xc[:1, :1]), pad)
                    y = F.avg_pool2d(F.pad(xc, pad), k, s) / F.avg_pool2d(ones, k, s)
                else:
                    y = F.avg_pool2d(xc, k, s)
            return y if nchw else y.permute(0, 2, 3, 1)
        return _Fn(run, 1)
    return load


_LOADERS["MaxPool"], _LOADERS["AvgPool"] = _pool("max"), _pool("avg")


@op("LRN")
def _lrn(n):
    r, bias, alpha, beta = (int(_attr(n, "depth_radius", 5)), float(_attr(n, "bias", 1.0)),
                            float(_attr(n, "alpha", 1.0)), float(_attr(n, "beta", 0.5)))

    def run(x):                                     # NHWC; TF's alpha is per element, torch's is per window
        size = 2 * r + 1
        return F.local_response_norm(x.permute(0, 3, 1, 2), size, alpha * size, beta, bias).permute(0, 2, 3, 1)
    return _Fn(run, 1)


@op("FusedBatchNorm", "FusedBatchNormV2", "FusedBatchNormV3")
def _fbn(n):
    """Five outputs like the TF kernel: y, batch mean, batch variance (Bessel-corrected, what TF feeds the moving
    average), and the two reserve spaces a later FusedBatchNormGrad reads — the mean and the biased variance."""
    eps, nchw = float(_attr(n, "epsilon", 1e-3)), _nchw(n)
    training = bool(_attr(n, "is_training", False))

    def run(x, scale, offset, mean, var):
        xc = x if nchw else x.permute(0, 3, 1, 2)
        if training or mean.numel() == 0:
            dims = [0] + list(range(2, xc.dim()))
            xf = xc.detach().float()
            mu, v = xf.mean(dims), xf.var(dims, unbiased=False)
            cnt = xf.numel() // max(xf.shape[1], 1)
            y = F.batch_norm(xc, None, None, scale, offset, True, 0.0, eps)
            stats = (mu, v * (cnt / max(cnt - 1, 1)), mu, v)
        else:
            y = F.batch_norm(xc, mean, var, scale, offset, False, 0.0, eps)
            stats = (mean, var, mean, var)
        return Table(y if nchw else y.permute(0, 2, 3, 1), *stats)
    return _Fn(run, 5)


@op("Reshape")
def _reshape(n): return _Fn(lambda x, s: x.reshape(_ints(s)), 2)


@op("Squeeze")
def _squeeze(n):
    dims = _attr(n, "squeeze_dims", []) or []

    def run(x):
        for d in sorted((int(v) % x.dim() for v in dims), reverse=True):
            x = x.squeeze(d)
        return x if dims else x.squeeze()
    return _Fn(run, 1)


@op("ExpandDims")
def _expand(n): return _Fn(lambda x, d: x.unsqueeze(_ints(d)[0] if _ints(d)[0] >= 0 else x.dim() + 1 + _ints(d)[0]), 2)


@op("Transpose")
def _transpose(n): return _Fn(lambda x, p: x.permute(_ints(p)), 2)


@op("ConcatV2")
def _concatv2(n): return _Fn(lambda *xs: torch.cat(list(xs[:-1]), _ints(xs[-1])[0]))


@op("Concat")
def _concat(n): return _Fn(lambda *xs: torch.cat(list(xs[1:]), _ints(xs[0])[0]))


@op("Pack")
def _pack(n):
    axis = int(_attr(n, "axis", 0))
    return _Fn(lambda *xs: torch.stack(list(xs), axis))


@op("Unpack")
def _unpack(n):
    axis = int(_attr(n, "axis", 0))
    return _Fn(lambda x: Table(*torch.unbind(x, axis)), 1)


@op("Split")
def _split(n):
    num = int(_attr(n, "num_split", 1))
    return _Fn(lambda d, x: Table(*torch.chunk(x, num, _ints(d)[0])), 2)


@op("SplitV")
def _splitv(n): return _Fn(lambda x, sizes, d: Table(*torch.split(x, _ints(sizes), _ints(d)[0])), 3)


@op("Slice")
def _slice(n):
    def run(x, begin, size):
        for d, (b, s) in enumerate(zip(_ints(begin), _ints(size))):
            x = x.narrow(d, b, x.shape[d] - b if s == -1 else s)
        return x
    return _Fn(run, 3)


@op("StridedSlice")
def _strided(n):
    bm, em, sm = int(_attr(n, "begin_mask", 0)), int(_attr(n, "end_mask", 0)), int(_attr(n, "shrink_axis_mask", 0))
    nm, elm = int(_attr(n, "new_axis_mask", 0)), int(_attr(n, "ellipsis_mask", 0))

    def run(x, begin, end, strides):
        b, e, s = _ints(begin), _ints(end), _ints(strides)
        idx = []
        for i in range(len(b)):
            if elm >> i & 1:
                idx.append(Ellipsis)
            elif nm >> i & 1:
                idx.append(None)
            elif sm >> i & 1:
                idx.append(b[i])
            else:
                idx.append(slice(None if bm >> i & 1 else b[i], None if em >> i & 1 else e[i], s[i]))
        if any(isinstance(i, slice) and i.step is not None and i.step < 0 for i in idx):
            raise ValueError("StridedSlice with negative strides is not supported")
        return x[tuple(idx)]
    return _Fn(run, 4)


@op("Pad", "PadV2", "MirrorPad")
def _pad(n):
    mode = {"Pad": "constant", "PadV2": "constant"}.get(n.op, "reflect" if _attr(n, "mode", "REFLECT") == "REFLECT" else "replicate")

    def run(x, pads, *value):
        flat = []
        for lo, hi in reversed(torch.as_tensor(pads).long().reshape(-1, 2).tolist()):
            flat += [lo, hi]
        if mode == "constant":
            return F.pad(x, flat, value=float(value[0]) if value else 0.0)
        return F.pad(x, flat, mode=mode)
    return _Fn(run)


@op("Tile")
def _tile(n): return _Fn(lambda x, m: x.repeat(_ints(m)), 2)


@op("Gather", "GatherV2")
def _gather(n):
    def run(x, idx, *axis):
        a = _ints(axis[0])[0] if axis else 0
        return torch.index_select(x, a, idx.long().reshape(-1)).reshape(x.shape[:a] + tuple(idx.shape) + x.shape[a + 1:])
    return _Fn(run)


@op("Shape")
def _shape(n): return TFO.Shape()


@op("Rank")
def _rank(n): return O.Rank()


@op("Size")
def _size(n): return _Fn(lambda x: torch.tensor(x.numel(), dtype=torch.int32), 1)


@op("Fill")
def _fill(n): return TFO.Fill()


@op("Range")
def _range(n): return O.RangeOps()


@op("Cast")
def _cast(n):
    dt = {1: torch.float32, 2: torch.float64, 3: torch.int32, 4: torch.uint8, 5: torch.int16, 6: torch.int8,
          9: torch.int64, 10: torch.bool, 19: torch.float16, 14: torch.bfloat16}[int(n.attr["DstT"].type)]
    return _Fn(lambda x: x.to(dt), 1)


def _reduce(fn):
    def load(n):
        keep = bool(_attr(n, "keep_dims", False) or _attr(n, "keepdims", False))

        def run(x, axes):
            for a in sorted({a % x.dim() for a in _ints(axes)}, reverse=True):
                x = fn(x, a, keep)
            return x
        return _Fn(run, 2)
    return load


_LOADERS["Mean"] = _reduce(lambda x, a, k: x.mean(a, keepdim=k))
_LOADERS["Sum"] = _reduce(lambda x, a, k: x.sum(a, keepdim=k))
_LOADERS["Prod"] = _reduce(lambda x, a, k: x.prod(a, keepdim=k))
_LOADERS["Max"] = _reduce(lambda x, a, k: x.max(a, keepdim=k)[0])
_LOADERS["Min"] = _reduce(lambda x, a, k: x.min(a, keepdim=k)[0])
_LOADERS["All"] = _reduce(lambda x, a, k: x.bool().all(a, keepdim=k))
_LOADERS["Any"] = _reduce(lambda x, a, k: x.bool().any(a, keepdim=k))


@op("ArgMax")
def _argmax(n): return _Fn(lambda x, a: x.argmax(_ints(a)[0]), 2)


@op("Select", "SelectV2")
def _select(n): return O.SelectOp()


@op("OneHot")
def _onehot(n): return O.OneHot(int(_attr(n, "axis", -1)))


@op("TopKV2", "TopK")
def _topk(n):
    return _Fn(lambda x, *k: Table(*[t if i == 0 else t.to(torch.int32) for i, t in
                                     enumerate(x.topk(_ints(k[0])[0] if k else int(_attr(n, "k", 1)), -1))]))


@op("InTopK", "InTopKV2")
def _intopk(n):
    k = int(_attr(n, "k", 1))
    return _Fn(lambda p, t, *kk: (p.topk(_ints(kk[0])[0] if kk else k, 1)[1] == t.long().reshape(-1, 1)).any(1))


@op("ResizeBilinear")
def _resize(n): return O.ResizeBilinearOps(bool(_attr(n, "align_corners", False)))


@op("SegmentSum")
def _segsum(n): return O.SegmentSum()


@op("Substr")
def _substr(n): return O.Substr()


@op("RandomUniform")
def _randu(n): return O.RandomUniform()


@op("TruncatedNormal")
def _truncn(n): return O.TruncatedNormal()


@op("DecodeJpeg", "DecodePng", "DecodeBmp", "DecodeGif")
def _decode(n): return TFO.DecodeImage(int(_attr(n, "channels", 0) or 0))


@op("Switch")
def _switch(n): return TFO.SwitchOps()


@op("Merge")
def _merge(n): return TFO.MergeOps()


@op("Enter")
def _enter(n): return TFO.Enter(_attr(n, "frame_name", ""), bool(_attr(n, "is_constant", False)))


@op("Exit")
def _exit(n): return TFO.Exit()


@op("NextIteration")
def _next(n): return TFO.NextIteration()


@op("LoopCond")
def _loopcond(n): return TFO.LoopCondition()


@op("InvertPermutation")
def _invert_perm(n): return TFO.InvertPermutation()


@op("ConcatOffset")
def _concat_offset(n): return TFO.ConcatOffset()


@op("RefEnter")
def _ref_enter(n): return TFO.Enter(_attr(n, "frame_name", ""), bool(_attr(n, "is_constant", False)))


@op("RefSwitch")
def _ref_switch(n): return TFO.SwitchOps()


@op("RefMerge")
def _ref_merge(n): return TFO.MergeOps()


@op("RefExit")
def _ref_exit(n): return TFO.Exit()


@op("RefNextIteration")
def _ref_next(n): return TFO.NextIteration()


@op("ControlTrigger")
def _control_trigger(n): return TFO.ControlDependency()


# ---- TensorArray / Stack resources (S/utils/tf/loaders/{TensorArrayV3,…,StackV2,StackPushV2,StackPopV2}.scala)
@op("TensorArrayV3")
def _ta_create(n): return TFO.TensorArrayCreator(bool(_attr(n, "dynamic_size", False)))


@op("TensorArrayWriteV3")
def _ta_write(n): return TFO.TensorArrayWrite()


@op("TensorArrayReadV3")
def _ta_read(n): return TFO.TensorArrayRead()


@op("TensorArrayGatherV3")
def _ta_gather(n): return TFO.TensorArrayGather()


@op("TensorArrayScatterV3")
def _ta_scatter(n): return TFO.TensorArrayScatter()


@op("TensorArrayConcatV3")
def _ta_concat(n): return TFO.TensorArrayConcat()


@op("TensorArraySplitV3")
def _ta_split(n): return TFO.TensorArraySplit()


@op("TensorArraySizeV3")
def _ta_size(n): return TFO.TensorArraySize()


@op("TensorArrayGradV3")
def _ta_grad(n): return TFO.TensorArrayGrad(_attr(n, "source", ""))


@op("TensorArrayCloseV3")
def _ta_close(n): return TFO.TensorArrayClose()


@op("StackV2", "Stack")
def _stack_create(n): return TFO.StackCreator()


@op("StackPushV2", "StackPush")
def _stack_push(n): return TFO.StackPush()


@op("StackPopV2", "StackPop")
def _stack_pop(n): return TFO.StackPop()


@op("VariableV2", "Variable", "VarHandleOp")
def _variable(n): return None          # resolved from binFile by the loader


from . import tensorflow_grad as _grad_ops  # noqa: E402,F401  (registers the *Grad / Conv3D / Dilation2D loaders)

SUPPORTED_OPS = sorted(_LOADERS)
_WEIGHT_SLOTS = {"Conv2D": [1], "DepthwiseConv2dNative": [1], "MatMul": [1], "BiasAdd": [1], "Conv2DBackpropInput": [1],
                 "FusedBatchNorm": [1, 2], "FusedBatchNormV2": [1, 2], "FusedBatchNormV3": [1, 2]}


_MULTI_OUTPUT_OPS = ("Switch", "Split", "SplitV", "Unpack", "TopKV2", "TopK", "ParseExample", "ParseSingleExample",
                     "BroadcastGradientArgs", "SoftmaxCrossEntropyWithLogits", "FusedBatchNormGrad",
                     "FusedBatchNormGradV2", "FusedBatchNormGradV3", "TensorArrayV3", "TensorArrayConcatV3",
                     "TensorArrayGradV3", "ConcatOffset", "FusedBatchNorm", "FusedBatchNormV2", "FusedBatchNormV3")


def _split_input(s: str):
    if s.startswith("^"):
        return s[1:], None, True
    name, _, idx = s.partition(":")
    return name, (int(idx) if idx else 0), False


def load_tf(path, inputs, outputs, byteOrder: str = "little_endian", binFile: str | None = None,
            generated_backward: bool = True):
    """Frozen GraphDef → Graph (or DynamicGraph when control-flow ops are present) computing ``outputs`` from
    ``inputs`` (node names). ``binFile`` supplies values for VariableV2 nodes: a ``.npz`` mapping node name → array (the
    reference reads its own dump format written by scripts/export_tf_checkpoint.py)."""
    gd = P.GraphDef()
    if isinstance(path, (bytes, bytearray)):
        gd.ParseFromString(bytes(path))
    else:
        with open(path, "rb") as f:
            gd.ParseFromString(f.read())
    inputs, outputs = list(inputs), list(outputs)
    by_name = {n.name: n for n in gd.node}
    var_values = dict(np.load(binFile)) if binFile else {}
    # nodes needed for the outputs, stopping at the declared inputs
    needed, stack = [], [_split_input(o)[0] for o in outputs]
    seen = set()
    while stack:
        name = stack.pop()
        if name in seen:
            continue
        if name not in by_name:
            raise ValueError(f"node {name} is not in the graph")
        seen.add(name)
        needed.append(name)
        if name in inputs:
            continue
        for inp in by_name[name].input:
            stack.append(_split_input(inp)[0])
    multi_out = set()
    for name in needed:
        for inp in by_name[name].input:
            src, idx, ctrl = _split_input(inp)
            if not ctrl and idx and idx > 0:
                multi_out.add(src)
    trainable_consts = set()
    for name in needed:
        nd = by_name[name]
        for slot in _WEIGHT_SLOTS.get(nd.op, []):
            data_inputs = [i for i in nd.input if not i.startswith("^")]
            if slot < len(data_inputs):
                src = _split_input(data_inputs[slot])[0]
                # look through Identity (the usual "w/read" node)
                while src in by_name and by_name[src].op == "Identity" and by_name[src].input:
                    src = _split_input(by_name[src].input[0])[0]
                if src in by_name and by_name[src].op == "Const":
                    trainable_consts.add(src)
    nodes: dict[str, Node] = {}
    has_control = False
    for name in needed:
        nd = by_name[name]
        if name in inputs:
            m = Identity()
        elif nd.op in ("VariableV2", "Variable", "VarHandleOp"):
            if name not in var_values:
                raise ValueError(f"variable {name} has no value: pass binFile with its contents")
            m = TFO.Variable(torch.from_numpy(np.asarray(var_values[name], dtype=np.float32)))
        elif nd.op == "Const" and name in trainable_consts:
            arr = P.tensor_to_numpy(nd.attr["value"].tensor)
            m = TFO.Variable(torch.from_numpy(arr.astype(np.float32)))
        else:
            loader = _LOADERS.get(nd.op)
            if loader is None:
                raise ValueError(f"unsupported TensorFlow op {nd.op} (node {name})")
            m = loader(nd)
        if isinstance(m, TFO.ControlOps):
            has_control = True
        m.setName(name)
        nodes[name] = Node(m, [])
    for name in needed:
        if name in inputs:
            continue
        nd = by_name[name]
        for inp in nd.input:
            src, idx, ctrl = _split_input(inp)
            if ctrl:
                continue                      # ordering-only edges: data dependencies already order a functional graph
            src_op = by_name[src].op
            multi = src in multi_out or src_op in _MULTI_OUTPUT_OPS
            nodes[src].add(nodes[name], (idx + 1) if multi else None)
    in_nodes = [nodes[i] for i in inputs]
    out_nodes = []
    for o in outputs:
        src, idx, _ = _split_input(o)
        if idx or (src not in inputs and (src in multi_out or by_name[src].op in _MULTI_OUTPUT_OPS)):
            sel = Node(_Fn(lambda *ts, i=idx: ts[i]), [nodes[src]])         # a Table input arrives unpacked
            sel.element.setName(f"{src}_{idx}")
            out_nodes.append(sel)
        else:
            out_nodes.append(nodes[src])
    if has_control:
        from ..nn.dynamic_graph import DynamicGraph

        return DynamicGraph(in_nodes, out_nodes, generate_backward=generated_backward)
    return Graph(in_nodes, out_nodes)


# ------------------------------------------------------------------------------------------------ export
class _GraphBuilder:
    def __init__(self):
        self.gd = P.GraphDef()
        self.gd.versions.producer = 21
        self.count = 0

    def node(self, op_name, name, inputs=(), **attrs):
        n = self.gd.node.add()
        n.op, n.name = op_name, name
        n.input.extend(inputs)
        for k, v in attrs.items():
            a = n.attr[k]
            if isinstance(v, bool):
                a.b = v
            elif isinstance(v, int):
                a.i = v
            elif isinstance(v, float):
                a.f = v
            elif isinstance(v, str):
                a.s = v.encode()
            elif isinstance(v, np.ndarray):
                a.tensor.CopyFrom(P.numpy_to_tensor(v))
            elif isinstance(v, (list, tuple)):
                a.list.i.extend(int(x) for x in v)
            elif isinstance(v, np.dtype) or v in P.DT_OF_NP:
                a.type = P.DT_OF_NP[np.dtype(v)]
        return name

    def const(self, name, arr):
        arr = np.asarray(arr)
        return self.node("Const", name, value=arr, dtype=arr.dtype)

    def fresh(self, base):
        self.count += 1
        return f"{base}_{self.count}"


_CUSTOM_SAVERS: dict = {}


class BigDLToTensorflow:
    """Exporter of one layer type (S/utils/tf/BigDLToTensorflow.scala:30-35): ``toTFDef(module, inputs, byteOrder)``
    gets the NodeDefs feeding the layer and returns the NodeDefs to append — the FIRST one is the layer's output.
    Build them with the ``Tensorflow`` helpers; register with ``TensorflowSaver.register("MyLayer", saver)``."""

    def toTFDef(self, module, inputs, byteOrder):  # pragma: no cover - interface
        raise NotImplementedError


class TensorflowDataFormat:
    """Activation layout of the exported graph (S/utils/tf/Tensorflow.scala:33-85)."""
    NCHW, NHWC = "NCHW", "NHWC"


class Tensorflow:
    """NodeDef builders (S/utils/tf/Tensorflow.scala:88-560). Every function returns a detached ``NodeDef`` whose
    inputs are the NAMES of the NodeDefs passed in."""

    _f32 = np.dtype(np.float32)

    @staticmethod
    def _node(op_name, name, inputs=(), **attrs):
        b = _GraphBuilder()
        b.node(op_name, name, [i.name if hasattr(i, "name") else str(i) for i in inputs], **attrs)
        return b.gd.node[0]

    @staticmethod
    def placeholder(dtype=None, shape=(), name="input"):
        n = Tensorflow._node("Placeholder", name, dtype=Tensorflow._f32 if dtype is None else np.dtype(dtype))
        for s in shape:
            n.attr["shape"].shape.dim.add().size = int(s if s is not None else -1)
        return n

    @staticmethod
    def const(value, name, byteOrder="little_endian"):
        arr = value.detach().cpu().numpy() if isinstance(value, torch.Tensor) else np.asarray(
            getattr(value, "t", value).detach().cpu().numpy() if hasattr(getattr(value, "t", None), "detach") else value)
        return Tensorflow._node("Const", name, value=np.ascontiguousarray(arr), dtype=arr.dtype)

    @staticmethod
    def identity(input, name): return Tensorflow._node("Identity", name, [input], T=Tensorflow._f32)

    @staticmethod
    def matmul(a, b, name, transposeA=False, transposeB=False):
        return Tensorflow._node("MatMul", name, [a, b], T=Tensorflow._f32, transpose_a=bool(transposeA),
                                transpose_b=bool(transposeB))

    @staticmethod
    def transpose(x, perm, name): return Tensorflow._node("Transpose", name, [x, perm], T=Tensorflow._f32)

    @staticmethod
    def biasAdd(value, bias, dataFormat, name):
        return Tensorflow._node("BiasAdd", name, [value, bias], T=Tensorflow._f32, data_format=str(dataFormat))

    @staticmethod
    def expandDims(input, dim, name): return Tensorflow._node("ExpandDims", name, [input, dim], T=Tensorflow._f32)

    @staticmethod
    def relu(features, name): return Tensorflow._node("Relu", name, [features], T=Tensorflow._f32)

    @staticmethod
    def tanh(input, name): return Tensorflow._node("Tanh", name, [input], T=Tensorflow._f32)

    @staticmethod
    def sigmoid(x, name): return Tensorflow._node("Sigmoid", name, [x], T=Tensorflow._f32)

    @staticmethod
    def softmax(logits, name): return Tensorflow._node("Softmax", name, [logits], T=Tensorflow._f32)

    @staticmethod
    def logSoftmax(logits, name): return Tensorflow._node("LogSoftmax", name, [logits], T=Tensorflow._f32)

    @staticmethod
    def rsqrt(x, name): return Tensorflow._node("Rsqrt", name, [x], T=Tensorflow._f32)

    @staticmethod
    def floor(x, name): return Tensorflow._node("Floor", name, [x], T=Tensorflow._f32)

    @staticmethod
    def reshape(tensor, shape, name): return Tensorflow._node("Reshape", name, [tensor, shape], T=Tensorflow._f32)

    @staticmethod
    def squeeze(input, axis, name):
        return Tensorflow._node("Squeeze", name, [input], T=Tensorflow._f32, squeeze_dims=[int(a) for a in axis])

    @staticmethod
    def _spatial(op_name, value, name, k, s, p, dataFormat, **extra):
        nhwc = str(dataFormat).upper() == "NHWC"
        ks = [1, k[0], k[1], 1] if nhwc else [1, 1, k[0], k[1]]
        st = [1, s[0], s[1], 1] if nhwc else [1, 1, s[0], s[1]]
        attrs = dict(T=Tensorflow._f32, strides=st, padding="SAME" if -1 in p else "VALID", data_format=str(dataFormat).upper())
        if op_name != "Conv2D":
            attrs["ksize"] = ks
        return Tensorflow._node(op_name, name, value, **attrs, **extra)

    @staticmethod
    def conv2D(input, filter, sW, sH, kW, kH, pW, pH, dataFormat, name):
        return Tensorflow._spatial("Conv2D", [input, filter], name, (kH, kW), (sH, sW), (pH, pW), dataFormat)

    @staticmethod
    def maxPool(value, kW, kH, pW, pH, sW, sH, dataFormat, name, ceilMode=False):
        return Tensorflow._spatial("MaxPool", [value], name, (kH, kW), (sH, sW), (-1, -1) if ceilMode else (pH, pW),
                                   dataFormat)

    @staticmethod
    def avgPool(value, kW, kH, pW, pH, sW, sH, dataFormat, name, ceilMode=False):
        return Tensorflow._spatial("AvgPool", [value], name, (kH, kW), (sH, sW), (-1, -1) if ceilMode else (pH, pW),
                                   dataFormat)

    @staticmethod
    def multiply(x, y, name): return Tensorflow._node("Mul", name, [x, y], T=Tensorflow._f32)

    @staticmethod
    def add(x, y, name): return Tensorflow._node("Add", name, [x, y], T=Tensorflow._f32)

    @staticmethod
    def subtract(x, y, name): return Tensorflow._node("Sub", name, [x, y], T=Tensorflow._f32)

    @staticmethod
    def realdiv(x, y, name): return Tensorflow._node("RealDiv", name, [x, y], T=Tensorflow._f32)

    @staticmethod
    def shape(input, name, outType=None): return Tensorflow._node("Shape", name, [input], T=Tensorflow._f32)

    @staticmethod
    def randomUniform(shape, name, dtype=None, seed=0):
        return Tensorflow._node("RandomUniform", name, [shape], dtype=Tensorflow._f32, seed=int(seed))

    @staticmethod
    def addN(inputs, name): return Tensorflow._node("AddN", name, list(inputs), T=Tensorflow._f32, N=len(inputs))

    @staticmethod
    def concat(inputs, name): return Tensorflow._node("ConcatV2", name, list(inputs), T=Tensorflow._f32, N=len(inputs) - 1)

    @staticmethod
    def pad(tensor, paddings, name): return Tensorflow._node("Pad", name, [tensor, paddings], T=Tensorflow._f32)

    @staticmethod
    def reduceMean(inputTensor, axis, keepDim, name):
        return Tensorflow._node("Mean", name, [inputTensor, axis], T=Tensorflow._f32, keep_dims=bool(keepDim))

    @staticmethod
    def split(splitDim, value, numSplit, name):
        return Tensorflow._node("Split", name, [splitDim, value], T=Tensorflow._f32, num_split=int(numSplit))

    @staticmethod
    def lrn(input, depthRadius, bias, alpha, beta, name):
        return Tensorflow._node("LRN", name, [input], T=Tensorflow._f32, depth_radius=int(depthRadius), bias=float(bias),
                                alpha=float(alpha), beta=float(beta))


class TensorflowSaver:
    """``saveGraph`` / ``saveGraphWithNodeDef`` / ``register`` (S/utils/tf/TensorflowSaver.scala:40-170)."""

    @staticmethod
    def saveGraph(model, inputs, path, byteOrder="little_endian", dataFormat=TensorflowDataFormat.NHWC):
        return save_tf(model, inputs, path, byteOrder, str(dataFormat).lower())

    @staticmethod
    def saveGraphWithNodeDef(model, inputs, path, byteOrder="little_endian", extraNodes=()):
        """``inputs`` are Placeholder NodeDefs (their names and shapes are used); ``extraNodes`` are appended as is."""
        specs = [(n.name, [d.size for d in n.attr["shape"].shape.dim]) for n in inputs]
        out = save_tf(model, specs, path, byteOrder, "nchw")
        if extraNodes:
            gd = P.GraphDef()
            with open(path, "rb") as f:
                gd.ParseFromString(f.read())
            for n in extraNodes:
                gd.node.add().CopyFrom(n)
            with open(path, "wb") as f:
                f.write(gd.SerializeToString())
        return out

    @staticmethod
    def register(className: str, saver: BigDLToTensorflow):
        _CUSTOM_SAVERS[className.split(".")[-1]] = saver


class TensorflowLoader:
    """``load(graphFile, inputs, outputs, byteOrder, binFile)`` and ``checkpoints(graphFile, binFile, byteOrder)``
    (S/utils/tf/TensorflowLoader.scala:55-110)."""

    @staticmethod
    def load(graphPrototxt, inputs, outputs, byteOrder="little_endian", binFile=None, generatedBackward=True):
        return load_tf(graphPrototxt, inputs, outputs, byteOrder, binFile, generatedBackward)

    @staticmethod
    def checkpoints(graphFile, binFile, byteOrder="little_endian"):
        from .tf_session import Session

        return Session(graphFile, binFile)

    @staticmethod
    def parse(graphProtoTxt: str):
        """The nodes of a binary GraphDef file."""
        gd = P.GraphDef()
        with open(graphProtoTxt, "rb") as f:
            gd.ParseFromString(f.read())
        return list(gd.node)


def save_tf(model, inputs, path: str, byteOrder: str = "little_endian", dataFormat: str = "nhwc"):
    """Export a Sequential / Graph of common layers as a frozen GraphDef (TensorflowSaver.saveGraph,
    BigDLToTensorflow.scala). ``inputs``: [(placeholder name, shape)] — shapes are given in the framework's NCHW
    convention; activations are transposed to NHWC at the boundary when ``dataFormat`` is "nhwc" and rank is 4."""
    from ..nn import layer as L
    from ..nn.module import Sequential

    b = _GraphBuilder()
    nhwc = dataFormat.lower() == "nhwc"
    f32 = np.dtype(np.float32)
    used: set = set()

    def emit(m, ins):
        name = m.getName()                # the generated default name too, so callers can ask loadTF for it
        if name in used:
            name = b.fresh(name)
        used.add(name)
        x = ins[0] if ins else None
        custom = _CUSTOM_SAVERS.get(type(m).__name__) or _CUSTOM_SAVERS.get(f"{type(m).__module__}.{type(m).__name__}")
        if custom is not None:
            by_name = {nd.name: nd for nd in b.gd.node}
            made = custom.toTFDef(m, [by_name[i] for i in ins], byteOrder)
            if not made:
                raise ValueError(f"saveTF: the saver registered for {type(m).__name__} returned no nodes")
            for nd in made:
                b.gd.node.add().CopyFrom(nd)
            return made[0].name           # the FIRST returned node is the layer's output (BigDLToTensorflow contract)
        if isinstance(m, Sequential):
            cur = ins
            for c in m.modules:
                cur = [emit(c, cur)]
            return cur[0]
        if isinstance(m, (L.Identity,)) or type(m).__name__ == "Input":
            return x
        if isinstance(m, L.Linear):
            w = b.const(name + "/weight", m.weight.detach().numpy().T.copy())
            y = b.node("MatMul", name + "/matmul", [x, w], T=f32, transpose_a=False, transpose_b=False)
            if m.with_bias:
                bias = b.const(name + "/bias", m.bias.detach().numpy())
                y = b.node("BiasAdd", name, [y, bias], T=f32, data_format="NHWC")
            return y
        if isinstance(m, L.SpatialConvolution):
            if m.n_group != 1:
                raise ValueError("grouped convolution cannot be exported to a TensorFlow Conv2D")
            w = b.const(name + "/filter", m.weight.detach().permute(2, 3, 1, 0).contiguous().numpy())
            pad = "SAME" if m.pad_w == -1 else "VALID"
            src = x
            if m.pad_w > 0 or m.pad_h > 0:
                p = np.array([[0, 0], [m.pad_h, m.pad_h], [m.pad_w, m.pad_w], [0, 0]] if nhwc else
                             [[0, 0], [0, 0], [m.pad_h, m.pad_h], [m.pad_w, m.pad_w]], dtype=np.int32)
                src = b.node("Pad", name + "/pad", [x, b.const(name + "/paddings", p)], T=f32)
            strides = [1, m.stride_h, m.stride_w, 1] if nhwc else [1, 1, m.stride_h, m.stride_w]
            y = b.node("Conv2D", name + "/conv", [src, w], T=f32, strides=strides, padding=pad,
                       data_format="NHWC" if nhwc else "NCHW")
            if m.with_bias:
                bias = b.const(name + "/bias", m.bias.detach().numpy())
                y = b.node("BiasAdd", name, [y, bias], T=f32, data_format="NHWC" if nhwc else "NCHW")
            return y
        if isinstance(m, (L.SpatialMaxPooling, L.SpatialAveragePooling)):
            if m.pad_w > 0 or m.pad_h > 0:
                raise ValueError("explicitly padded pooling cannot be exported")
            k = [1, m.kh, m.kw, 1] if nhwc else [1, 1, m.kh, m.kw]
            s = [1, m.dh, m.dw, 1] if nhwc else [1, 1, m.dh, m.dw]
            return b.node("MaxPool" if isinstance(m, L.SpatialMaxPooling) else "AvgPool", name, [x], T=f32, ksize=k,
                          strides=s, padding="SAME" if m.pad_w == -1 else "VALID", data_format="NHWC" if nhwc else "NCHW")
        simple = {L.ReLU: "Relu", L.ReLU6: "Relu6", L.Tanh: "Tanh", L.Sigmoid: "Sigmoid", L.SoftMax: "Softmax",
                  L.LogSoftMax: "LogSoftmax", L.SoftPlus: "Softplus", L.SoftSign: "Softsign", L.Abs: "Abs",
                  L.Exp: "Exp", L.Log: "Log", L.Sqrt: "Sqrt", L.Square: "Square", L.ELU: "Elu"}
        for cls, tf_op in simple.items():
            if type(m) is cls:
                return b.node(tf_op, name, [x], T=f32)
        if isinstance(m, L.Dropout):
            return x                                  # inference graph
        if isinstance(m, L.BatchNormalization):
            rm, rv = m._buffers["runningMean"].numpy(), m._buffers["runningVar"].numpy()
            g = m.weight.detach().numpy() if m.affine else np.ones_like(rm)
            be = m.bias.detach().numpy() if m.affine else np.zeros_like(rm)
            outs = b.node("FusedBatchNorm", name, [x, b.const(name + "/gamma", g), b.const(name + "/beta", be),
                                                   b.const(name + "/mean", rm), b.const(name + "/var", rv)], T=f32,
                          epsilon=float(m.eps), is_training=False, data_format="NHWC" if nhwc else "NCHW")
            return outs
        if isinstance(m, (L.Reshape, L.View, L.InferReshape)):
            sizes = list(getattr(m, "size", None) or getattr(m, "sizes", []))
            src = x
            if nhwc and getattr(m, "_tf_from_rank4", True) and len(sizes) <= 2:
                src = b.node("Transpose", name + "/to_nchw", [x, b.const(name + "/perm", np.array([0, 3, 1, 2], np.int32))], T=f32)
            shp = b.const(name + "/shape", np.array([-1] + [int(s) for s in sizes if s != -1], dtype=np.int32))
            return b.node("Reshape", name, [src, shp], T=f32)
        if isinstance(m, L.CAddTable):
            return b.node("AddN", name, ins, T=f32, N=len(ins))
        if isinstance(m, L.CMulTable):
            return b.node("Mul", name, ins, T=f32)
        if isinstance(m, L.JoinTable):
            axis = m.dimension - 1
            if nhwc and axis == 1:
                axis = 3
            return b.node("ConcatV2", name, list(ins) + [b.const(name + "/axis", np.array(axis, np.int32))], T=f32, N=len(ins))
        if isinstance(m, L.SpatialCrossMapLRN):
            if not nhwc:
                raise ValueError("LRN export needs NHWC")
            return b.node("LRN", name, [x], T=f32, depth_radius=(m.size - 1) // 2, bias=float(m.k),
                          alpha=float(m.alpha) / m.size, beta=float(m.beta))
        # the remaining entries of TensorflowSaver.scala:171-194 — these act on non-image tensors, axes are taken as given
        if isinstance(m, L.Squeeze):
            dims = [int(d) - 1 + (1 if m.batch_mode else 0) for d in (m.dims or [])]
            return b.node("Squeeze", name, [x], T=f32, squeeze_dims=dims)
        if isinstance(m, L.Mean):
            axis = b.const(name + "/axis", np.array([m.dimension - 1], np.int32))
            return b.node("Mean", name, [x, axis], T=f32, keep_dims=not m.squeeze)
        if isinstance(m, L.Padding):
            rank = m.n_input_dim + 1                       # exported graphs are batched
            d = abs(m.dim) if m.n_input_dim else m.dim - 1
            p = np.zeros((rank, 2), np.int32)
            p[d, 0 if m.pad < 0 else 1] = abs(m.pad)
            if m.value != 0.0:
                return b.node("PadV2", name, [x, b.const(name + "/paddings", p),
                                              b.const(name + "/value", np.array(m.value, np.float32))], T=f32)
            return b.node("Pad", name, [x, b.const(name + "/paddings", p)], T=f32)
        if isinstance(m, L.Scale):
            ps = [q.detach().numpy() for _, _, q in m.leaf_parameters()]
            w, bias = ps[0], ps[1]
            if nhwc and w.ndim == 4:
                w, bias = w.transpose(0, 2, 3, 1), bias.transpose(0, 2, 3, 1)
            y = b.node("Mul", name + "/mul", [x, b.const(name + "/weight", w.copy())], T=f32)
            return b.node("Add", name, [y, b.const(name + "/bias", bias.copy())], T=f32)
        if isinstance(m, L.TemporalConvolution):
            # (B, T, Cin) → (B, 1, T, Cin) NHWC → Conv2D with a [1, kw, Cin, Cout] filter → back to (B, T', Cout)
            w = m.weight.detach().numpy().reshape(m.out, m.kw, m.inp).transpose(1, 2, 0)[None].copy()
            e = b.node("ExpandDims", name + "/expand", [x, b.const(name + "/expand_axis", np.array(1, np.int32))], T=f32)
            y = b.node("Conv2D", name + "/conv", [e, b.const(name + "/weight", w)], T=f32, strides=[1, 1, m.dw, 1],
                       padding="VALID", data_format="NHWC")
            y = b.node("BiasAdd", name + "/bias_add", [y, b.const(name + "/bias", m.bias.detach().numpy().copy())],
                       T=f32, data_format="NHWC")
            return b.node("Squeeze", name, [y], T=f32, squeeze_dims=[1])
        raise ValueError(f"saveTF: unsupported module {type(m).__name__}")

    placeholders = []
    for pname, shape in inputs:
        n = b.gd.node.add()
        n.op, n.name = "Placeholder", pname
        n.attr["dtype"].type = 1
        shp = list(shape)
        if nhwc and len(shp) == 4:
            shp = [shp[0], shp[2], shp[3], shp[1]]
        for s in shp:
            n.attr["shape"].shape.dim.add().size = int(s if s is not None else -1)
        placeholders.append(pname)
    if isinstance(model, Graph):
        names: dict[int, str] = {}
        it = iter(placeholders)
        for n in model.exec_order:
            if not n.prev:
                names[id(n)] = next(it)
            else:
                names[id(n)] = emit(n.element, [names[id(p)] for p in n.prev])
        out = [names[id(n)] for n in model.output_nodes]
    else:
        out = [emit(model, placeholders[:1])]
    with open(path, "wb") as f:
        f.write(b.gd.SerializeToString())
    return out


class TFUtils:
    """S/utils/tf/TFUtils.scala: ``parseTensor(TensorProto, endian)`` → tensor."""

    @staticmethod
    def parseTensor(tfTensor, endian: str = "little_endian"):
        arr = P.tensor_to_numpy(tfTensor, "big_endian" if str(endian).lower().startswith("big") else "little_endian")
        if arr.dtype == object or arr.dtype.kind in "US":
            return arr
        return torch.from_numpy(np.ascontiguousarray(arr))

    @staticmethod
    def saveToHDFS(localFiles, outputPath: str, overwrite: bool = True):
        """Copy local files into ``outputPath`` (any scheme ``utils.file`` knows: hdfs://, s3://, file)."""
        import os

        from ..utils.file import File

        out = []
        for f in localFiles:
            dst = outputPath.rstrip("/") + "/" + os.path.basename(f)
            with open(f, "rb") as fh:
                File.saveBytes(fh.read(), dst, overwrite)
            out.append(dst)
        return out


# ------------------------------------------------------------------------------------------------ built-in savers by name
class _BuiltinSaver(BigDLToTensorflow):
    """The reference keeps one ``XToTF`` object per exportable layer (BigDLToTensorflow.scala:60-640); here they all
    run the exporter of ``save_tf`` on the single layer and hand back its nodes, output node first."""

    def __init__(self, *layer_names):
        self.layer_names = layer_names

    def toTFDef(self, module, inputs, byteOrder="little_endian"):
        import os
        import tempfile

        from ..nn import layer as L
        from ..nn.module import Graph

        if self.layer_names and type(module).__name__ not in self.layer_names:
            raise TypeError(f"{type(module).__name__} is not exported by this saver ({', '.join(self.layer_names)})")
        ins = [L.Input() for _ in inputs]
        g = Graph(ins, module.inputs(*ins) if len(ins) != 1 else module.inputs(ins[0]))
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, "one.pb")
            out = save_tf(g, [(n.name, []) for n in inputs], path, byteOrder, "nchw")
            nodes = TensorflowLoader.parse(path)
        given = {n.name for n in inputs}
        made = [n for n in nodes if n.name not in given]
        made.sort(key=lambda n: n.name != out[0])
        return made


InputToTF = _BuiltinSaver("Input", "Identity")
ReLUToTF, TanhToTF, SigmoidToTF = _BuiltinSaver("ReLU"), _BuiltinSaver("Tanh"), _BuiltinSaver("Sigmoid")
SoftMaxToTF, LogSoftMaxToTF = _BuiltinSaver("SoftMax"), _BuiltinSaver("LogSoftMax")
LinearToTF = _BuiltinSaver("Linear")
SpatialConvolutionToTF = _BuiltinSaver("SpatialConvolution")
TemporalConvolutionToTF = _BuiltinSaver("TemporalConvolution")
SqueezeToTF, ReshapeToTF, ViewToTF = _BuiltinSaver("Squeeze"), _BuiltinSaver("Reshape", "InferReshape"), _BuiltinSaver("View")
MaxpoolToTF, AvgpoolToTF = _BuiltinSaver("SpatialMaxPooling"), _BuiltinSaver("SpatialAveragePooling")
PaddingToTF, DropoutToTF = _BuiltinSaver("Padding"), _BuiltinSaver("Dropout")
CAddTableToTF, CMultTableToTF, JoinTableToTF = _BuiltinSaver("CAddTable"), _BuiltinSaver("CMulTable"), _BuiltinSaver("JoinTable")
MeanToTF, ScaleToTF, LRNToTF = _BuiltinSaver("Mean"), _BuiltinSaver("Scale"), _BuiltinSaver("SpatialCrossMapLRN")
BatchNorm2DToTF = _BuiltinSaver("SpatialBatchNormalization", "BatchNormalization")
