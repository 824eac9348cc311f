"""TensorFlow-style forward-only operations (S/nn/ops/*.scala, 71 files). An ``Operation`` has no backward
(Operation.scala:34-60: ``updateGradInput`` throws); most take a Table of tensors. They are what the TensorFlow loader
instantiates for graph nodes that are not trainable layers, and what feature-column pipelines are built from
(BucketizedCol, CategoricalColHashBucket, CategoricalColVocaList, CrossCol, IndicatorCol, Kv2Tensor, MkString).
Element-wise math runs as torch ops on whatever device the inputs live on."""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

from ..utils.table import Table
from ..utils.util import HashFunc
from .module import Module


def _ins(x, n=None):
    items = x.to_list() if isinstance(x, Table) else (list(x) if isinstance(x, (list, tuple)) else [x])
    if n is not None and len(items) != n:
        raise ValueError(f"expected {n} inputs, got {len(items)}")
    return items


class Operation(Module):
    def backward(self, input, gradOutput):
        raise RuntimeError("Operation does not support backward() method")

    def updateGradInput(self, input, gradOutput):
        raise RuntimeError("Operation does not support updateGradInput() method")

    def accGradParameters(self, input, gradOutput):
        raise RuntimeError("Operation does not support updateGradInput() method")


class ModuleToOperation(Operation):
    """Wrap any module as a forward-only op (ModuleToOperation.scala)."""

    def __init__(self, module: Module):
        super().__init__()
        self.module = module
        self.modules = [module]

    def updateOutput(self, input):
        return self.module._call(input)


class TensorOp(Operation):
    """Chainable closure over one tensor (TensorOp.scala): ``TensorOp().add(1).pow(2)`` or ``TensorOp(fn)``."""

    def __init__(self, fn=None):
        super().__init__()
        self.fns = [fn] if fn is not None else []

    def _then(self, f):
        self.fns.append(f)
        return self

    def __rshift__(self, other):
        out = TensorOp()
        out.fns = self.fns + other.fns
        return out

    def add(self, v): return self._then(lambda t: t + v)
    def sub(self, v): return self._then(lambda t: t - v)
    def mul(self, v): return self._then(lambda t: t * v)
    def div(self, v): return self._then(lambda t: t / v)
    def pow(self, v): return self._then(lambda t: t.pow(v))
    def sqrt(self): return self._then(torch.sqrt)
    def exp(self): return self._then(torch.exp)
    def log(self): return self._then(torch.log)
    def log1p(self): return self._then(torch.log1p)
    def abs(self): return self._then(torch.abs)
    def tanh(self): return self._then(torch.tanh)
    def sigmoid(self): return self._then(torch.sigmoid)
    def floor(self): return self._then(torch.floor)
    def ceil(self): return self._then(torch.ceil)
    def negative(self): return self._then(torch.neg)
    def inv(self): return self._then(torch.reciprocal)
    def square(self): return self._then(torch.square)
    def ge(self, v): return self._then(lambda t: (t >= v).to(t.dtype))
    def gt(self, v): return self._then(lambda t: (t > v).to(t.dtype))
    def le(self, v): return self._then(lambda t: (t <= v).to(t.dtype))
    def lt(self, v): return self._then(lambda t: (t < v).to(t.dtype))
    def eq(self, v): return self._then(lambda t: (t == v).to(t.dtype))
    def transpose(self, d1, d2): return self._then(lambda t: t.transpose(d1 - 1, d2 - 1))
    def t(self): return self._then(lambda t: t.t())

    def updateOutput(self, input):
        x = input
        for f in self.fns:
            x = f(x)
        return x


def _unary(name, fn, doc=""):
    def updateOutput(self, input):
        return fn(input)
    return type(name, (Operation,), {"updateOutput": updateOutput, "__doc__": doc or f"element-wise {name}"})


def _binary(name, fn, doc=""):
    def updateOutput(self, input):
        a, b = _ins(input, 2)
        return fn(a, b)
    return type(name, (Operation,), {"updateOutput": updateOutput, "__doc__": doc or f"element-wise {name} of two tensors"})


Ceil = _unary("Ceil", torch.ceil)
Floor = _unary("Floor", torch.floor)
Rint = _unary("Rint", torch.round, "round half to even")
Round = _unary("Round", torch.round)
Sign = _unary("Sign", torch.sign)
Exp = _unary("Exp", torch.exp)
Expm1 = _unary("Expm1", torch.expm1)
Inv = _unary("Inv", torch.reciprocal)
Erf = _unary("Erf", torch.erf)
Erfc = _unary("Erfc", torch.erfc)
Lgamma = _unary("Lgamma", torch.lgamma)
Digamma = _unary("Digamma", torch.digamma)
IsFinite = _unary("IsFinite", torch.isfinite)
IsInf = _unary("IsInf", torch.isinf)
IsNan = _unary("IsNan", torch.isnan)
LogicalNot = _unary("LogicalNot", torch.logical_not)
Rank = _unary("Rank", lambda t: torch.tensor(t.dim(), dtype=torch.int32))
L2Loss = _unary("L2Loss", lambda t: (t * t).sum() / 2, "sum(t^2)/2")

Equal = _binary("Equal", torch.eq)
NotEqual = _binary("NotEqual", torch.ne)
Greater = _binary("Greater", torch.gt)
GreaterEqual = _binary("GreaterEqual", torch.ge)
Less = _binary("Less", torch.lt)
LessEqual = _binary("LessEqual", torch.le)
LogicalAnd = _binary("LogicalAnd", torch.logical_and)
LogicalOr = _binary("LogicalOr", torch.logical_or)
Maximum = _binary("Maximum", torch.maximum)
Minimum = _binary("Minimum", torch.minimum)
Pow = _binary("Pow", torch.pow)
SquaredDifference = _binary("SquaredDifference", lambda a, b: (a - b) ** 2)
FloorDiv = _binary("FloorDiv", lambda a, b: torch.floor_divide(a, b))
FloorMod = _binary("FloorMod", lambda a, b: torch.remainder(a, b))
Mod = _binary("Mod", lambda a, b: torch.fmod(a, b))
TruncateDiv = _binary("TruncateDiv", lambda a, b: torch.div(a, b, rounding_mode="trunc"))


class Compare(Operation):
    """Base of the comparison ops in the reference (Compare.scala)."""

    def compareFloat(self, a, b):  # pragma: no cover - abstract
        raise NotImplementedError

    def updateOutput(self, input):
        a, b = _ins(input, 2)
        return self.compareFloat(a, b)


class ApproximateEqual(Operation):
    def __init__(self, tolerance: float = 1e-5):
        super().__init__()
        self.tolerance = tolerance

    def updateOutput(self, input):
        a, b = _ins(input, 2)
        return (a - b).abs() < self.tolerance


class _Reduce(Operation):
    """(tensor, axes[0-based]) reductions with ``keepDims`` (All/Any/Sum/Prod/Max.scala)."""

    fn = None

    def __init__(self, keep_dims: bool = False, start_from_zero: bool = True):
        super().__init__()
        self.keep_dims, self.zero = keep_dims, start_from_zero

    def updateOutput(self, input):
        t, axes = _ins(input, 2)
        ax = [int(a) for a in torch.as_tensor(axes).reshape(-1).tolist()]
        if not self.zero:
            ax = [a - 1 for a in ax]
        out = t
        for a in sorted({a % t.dim() for a in ax}, reverse=True):
            out = type(self).fn(out, a)
        return out

    @staticmethod
    def _wrap(f):
        return f


class All(_Reduce):
    def updateOutput(self, input):
        t, axes = _ins(input, 2)
        return _reduce(t.bool(), axes, self, lambda x, a, k: x.all(a, keepdim=k))


class Any(_Reduce):
    def updateOutput(self, input):
        t, axes = _ins(input, 2)
        return _reduce(t.bool(), axes, self, lambda x, a, k: x.any(a, keepdim=k))


class Sum(_Reduce):
    def updateOutput(self, input):
        t, axes = _ins(input, 2)
        return _reduce(t, axes, self, lambda x, a, k: x.sum(a, keepdim=k))


class Prod(_Reduce):
    def updateOutput(self, input):
        t, axes = _ins(input, 2)
        return _reduce(t, axes, self, lambda x, a, k: x.prod(a, keepdim=k))


class Max(_Reduce):
    def updateOutput(self, input):
        t, axes = _ins(input, 2)
        return _reduce(t, axes, self, lambda x, a, k: x.max(a, keepdim=k)[0])


def _reduce(t, axes, op, fn):
    ax = [int(a) for a in torch.as_tensor(axes).reshape(-1).tolist()]
    if not op.zero:
        ax = [a - 1 for a in ax]
    for a in sorted({a % t.dim() for a in ax}, reverse=True):
        t = fn(t, a, op.keep_dims)
    return t


class ArgMax(Operation):
    def updateOutput(self, input):
        t, axis = _ins(input, 2)
        return t.argmax(int(torch.as_tensor(axis).reshape(-1)[0])) .to(torch.int32)


class BatchMatMul(Operation):
    def __init__(self, adj_x: bool = False, adj_y: bool = False):
        super().__init__()
        self.adj_x, self.adj_y = adj_x, adj_y

    def updateOutput(self, input):
        a, b = _ins(input, 2)
        a = a.transpose(-1, -2) if self.adj_x else a
        b = b.transpose(-1, -2) if self.adj_y else b
        return torch.matmul(a, b)


class Cast(Operation):
    """Convert to ``dtype`` — a torch dtype or the reference's type names ("float", "double", "int", "long", ...)."""

    _NAMES = {"float": torch.float32, "float32": torch.float32, "double": torch.float64, "float64": torch.float64,
              "int": torch.int32, "int32": torch.int32, "long": torch.int64, "int64": torch.int64,
              "short": torch.int16, "char": torch.int8, "byte": torch.uint8, "bool": torch.bool, "boolean": torch.bool,
              "half": torch.float16, "bfloat16": torch.bfloat16}

    def __init__(self, dtype=torch.float32):
        super().__init__()
        self.dtype = self._NAMES[dtype.lower()] if isinstance(dtype, str) else dtype

    def updateOutput(self, input):
        return input.to(self.dtype)


class CrossEntropy(Operation):
    """softmax cross entropy with per-row soft labels: input (logits, labels) → per-sample loss."""

    def updateOutput(self, input):
        logits, labels = _ins(input, 2)
        return -(labels * F.log_softmax(logits, dim=-1)).sum(-1)


class DepthwiseConv2D(Operation):
    def __init__(self, stride_w=1, stride_h=1, pad_w=0, pad_h=0, data_format="NHWC"):
        super().__init__()
        self.s, self.p, self.fmt = (stride_h, stride_w), (pad_h, pad_w), data_format.upper()

    def updateOutput(self, input):
        x, w = _ins(input, 2)                       # w: [kh, kw, in, multiplier] (TensorFlow layout)
        if self.fmt == "NHWC":
            x = x.permute(0, 3, 1, 2)
        kh, kw, cin, mult = w.shape
        wt = w.permute(2, 3, 0, 1).reshape(cin * mult, 1, kh, kw)
        pad = self.p
        if pad[0] == -1 or pad[1] == -1:
            H, W = x.shape[2:]
            th = max((math.ceil(H / self.s[0]) - 1) * self.s[0] + kh - H, 0)
            tw = max((math.ceil(W / self.s[1]) - 1) * self.s[1] + kw - W, 0)
            x = F.pad(x, (tw // 2, tw - tw // 2, th // 2, th - th // 2))
            pad = (0, 0)
        y = F.conv2d(x, wt, None, self.s, pad, 1, cin)
        return y.permute(0, 2, 3, 1) if self.fmt == "NHWC" else y


class Dilation2D(Operation):
    """Grey-scale morphological dilation (Dilation2D.scala): out = max over window of (x + filter)."""

    def __init__(self, strides, rates, padding="VALID"):
        super().__init__()
        self.strides, self.rates, self.padding = strides, rates, padding.upper()

    def updateOutput(self, input):
        x, w = _ins(input, 2)                        # NHWC, filter [kh, kw, C]
        N, H, W, C = x.shape
        kh, kw, _ = w.shape
        sh, sw, rh, rw = self.strides[1], self.strides[2], self.rates[1], self.rates[2]
        xc = x.permute(0, 3, 1, 2)
        if self.padding == "SAME":
            eh, ew = (kh - 1) * rh + 1, (kw - 1) * rw + 1
            th = max((math.ceil(H / sh) - 1) * sh + eh - H, 0)
            tw = max((math.ceil(W / sw) - 1) * sw + ew - W, 0)
            xc = F.pad(xc, (tw // 2, tw - tw // 2, th // 2, th - th // 2), value=float("-inf"))
        cols = F.unfold(xc, (kh, kw), dilation=(rh, rw), stride=(sh, sw))
        L_ = cols.shape[-1]
        cols = cols.reshape(N, C, kh * kw, L_) + w.permute(2, 0, 1).reshape(1, C, kh * kw, 1)
        out = cols.max(2)[0]
        Ho = (xc.shape[2] - ((kh - 1) * rh + 1)) // sh + 1
        return out.reshape(N, C, Ho, -1).permute(0, 2, 3, 1)


class Gather(Operation):
    """(params, 0-based indices) → params[indices] along dim 0."""

    def updateOutput(self, input):
        t, idx = _ins(input, 2)
        return t[idx.long()]


class InTopK(Operation):
    def __init__(self, k: int, start_from_zero: bool = False):
        super().__init__()
        self.k, self.zero = k, start_from_zero

    def updateOutput(self, input):
        pred, tgt = _ins(input, 2)
        tgt = tgt.long() if self.zero else tgt.long() - 1
        top = pred.topk(self.k, dim=1)[1]
        return (top == tgt.reshape(-1, 1)).any(1)


class OneHot(Operation):
    def __init__(self, axis: int = -1):
        super().__init__()
        self.axis = axis

    def updateOutput(self, input):
        idx, depth, on, off = _ins(input, 4)
        d = int(torch.as_tensor(depth).item())
        oh = F.one_hot(idx.long().clamp(min=0), d).to(torch.as_tensor(on).dtype)
        out = oh * on + (1 - oh) * off
        out = torch.where((idx.long() < 0).unsqueeze(-1), torch.as_tensor(off).expand_as(out), out)
        return out if self.axis == -1 else out.movedim(-1, self.axis)


class Pad(Operation):
    def __init__(self, mode: str = "CONSTANT", constant_value: float = 0.0):
        super().__init__()
        self.mode, self.value = mode.upper(), constant_value

    def updateOutput(self, input):
        t, pads = _ins(input, 2)
        p = torch.as_tensor(pads).long().reshape(-1, 2).tolist()
        flat = []
        for lo, hi in reversed(p):
            flat += [lo, hi]
        mode = {"CONSTANT": "constant", "REFLECT": "reflect", "SYMMETRIC": "replicate"}[self.mode]
        return F.pad(t, flat, mode=mode, value=self.value) if mode == "constant" else F.pad(t, flat, mode=mode)


class RandomUniform(Operation):
    def __init__(self, minval=0.0, maxval=1.0, seed=None):
        super().__init__()
        self.lo, self.hi, self.seed = minval, maxval, seed

    def updateOutput(self, input):
        shape = [int(v) for v in torch.as_tensor(input).reshape(-1).tolist()]
        from ..utils.random_generator import RNG

        return RNG.fill_uniform(torch.empty(shape), self.lo, self.hi)


class TruncatedNormal(Operation):
    def __init__(self, mean=0.0, stddev=1.0, seed=0):
        super().__init__()
        self.mean, self.std = mean, stddev

    def updateOutput(self, input):
        shape = [int(v) for v in torch.as_tensor(input).reshape(-1).tolist()]
        t = torch.empty(shape)
        torch.nn.init.trunc_normal_(t, self.mean, self.std, self.mean - 2 * self.std, self.mean + 2 * self.std)
        return t


class RangeOps(Operation):
    def updateOutput(self, input):
        s, e, d = [torch.as_tensor(v).item() for v in _ins(input, 3)]
        return torch.arange(s, e, d)


class ResizeBilinearOps(Operation):
    def __init__(self, align_corners: bool = False):
        super().__init__()
        self.align = align_corners

    def updateOutput(self, input):
        x, size = _ins(input, 2)                     # NHWC
        h, w = [int(v) for v in torch.as_tensor(size).reshape(-1).tolist()]
        y = F.interpolate(x.permute(0, 3, 1, 2), (h, w), mode="bilinear", align_corners=self.align)
        return y.permute(0, 2, 3, 1)


class SegmentSum(Operation):
    def updateOutput(self, input):
        t, seg = _ins(input, 2)
        seg = seg.long()
        out = torch.zeros((int(seg.max()) + 1,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        return out.index_add_(0, seg, t)


class SelectOp(Operation):
    """(condition, then, else) element-wise select (Select.scala in nn/ops)."""

    def updateOutput(self, input):
        c, a, b = _ins(input, 3)
        c = c.bool()
        if c.dim() == 0:
            return a if bool(c) else b
        while c.dim() < a.dim():
            c = c.unsqueeze(-1)
        return torch.where(c, a, b)


class Slice(Operation):
    def __init__(self, begin, size):
        super().__init__()
        self.begin, self.size = list(begin), list(size)

    def updateOutput(self, input):
        t = input
        for d, (b, s) in enumerate(zip(self.begin, self.size)):
            t = t.narrow(d, b, t.shape[d] - b if s == -1 else s)
        return t


class Tile(Operation):
    def updateOutput(self, input):
        t, mult = _ins(input, 2)
        return t.repeat([int(v) for v in torch.as_tensor(mult).reshape(-1).tolist()])


class TopK(Operation):
    def __init__(self, k: int, sorted: bool = True, start_index: int = 1):
        super().__init__()
        self.k, self.start = k, start_index

    def updateOutput(self, input):
        v, i = input.topk(self.k, dim=-1)
        return Table(v, (i + self.start).to(torch.int32))


class Substr(Operation):
    def updateOutput(self, input):
        s, pos, ln = _ins(input, 3)
        p, n = int(torch.as_tensor(pos).item()), int(torch.as_tensor(ln).item())
        if isinstance(s, (bytes, str)):
            return s[p:p + n]
        return np.array([x[p:p + n] for x in np.asarray(s).reshape(-1)], dtype=object).reshape(np.asarray(s).shape)


# ------------------------------------------------------------------------------------------------ feature columns
class BucketizedCol(Operation):
    """Values → bucket index by ``boundaries`` (left-closed): (-inf,b0) → 0, [b0,b1) → 1, …"""

    def __init__(self, boundaries):
        super().__init__()
        if len(boundaries) < 1:
            raise ValueError("the length of boundaries must be more than 1")
        self.boundaries = torch.tensor(sorted(boundaries), dtype=torch.float64)

    def updateOutput(self, input):
        return torch.bucketize(input.double(), self.boundaries, right=True).to(torch.int32)


def _split_strings(col, delimiter):
    return [[] if (s is None or s == "") else str(s).split(delimiter) for s in np.asarray(col, dtype=object).reshape(-1)]


def _sparse_from_rows(rows, dtype=torch.int32):
    idx, vals = [], []
    width = max((len(r) for r in rows), default=0)
    for i, r in enumerate(rows):
        for j, v in enumerate(r):
            idx.append([i, j])
            vals.append(v)
    from ..tensor.sparse import SparseTensor

    return SparseTensor(np.asarray(idx, dtype=np.int64).reshape(-1, 2).T if idx else np.zeros((2, 0), dtype=np.int64),
                        torch.tensor(vals, dtype=dtype), (len(rows), max(width, 1)))


class CategoricalColHashBucket(Operation):
    def __init__(self, hash_bucket_size: int, str_delimiter: str = ",", is_sparse: bool = True):
        super().__init__()
        self.n, self.delim, self.sparse = hash_bucket_size, str_delimiter, is_sparse

    def updateOutput(self, input):
        rows = [[HashFunc.murmurhash3_32(w, 0) % self.n for w in r] for r in _split_strings(input, self.delim)]
        sp = _sparse_from_rows(rows)
        return sp if self.sparse else sp.to_dense().t


class CategoricalColVocaList(Operation):
    def __init__(self, voca_list, str_delimiter: str = ",", is_set_default: bool = False, num_oov_buckets: int = 0):
        super().__init__()
        if len(set(voca_list)) != len(voca_list):
            raise ValueError("The vocabulary list contains duplicate keys")
        if is_set_default and num_oov_buckets:
            raise ValueError("defaultValue and numOovBuckets cannot both be set")
        self.voc = {w: i for i, w in enumerate(voca_list)}
        self.delim, self.default, self.oov = str_delimiter, is_set_default, num_oov_buckets

    def updateOutput(self, input):
        n = len(self.voc)
        rows = []
        for r in _split_strings(input, self.delim):
            out = []
            for w in r:
                if w in self.voc:
                    out.append(self.voc[w])
                elif self.default:
                    out.append(n)
                elif self.oov:
                    out.append(n + HashFunc.murmurhash3_32(w, 0) % self.oov)
            rows.append(out)
        return _sparse_from_rows(rows)


class CrossCol(Operation):
    """Hash-cross of several string columns (CrossCol.scala): every combination of the per-column tokens of a row is
    hashed into ``hash_bucket_size`` buckets."""

    def __init__(self, hash_bucket_size: int, str_delimiter: str = ","):
        super().__init__()
        self.n, self.delim = hash_bucket_size, str_delimiter

    def updateOutput(self, input):
        cols = [_split_strings(c, self.delim) for c in _ins(input)]
        rows = []
        for i in range(len(cols[0])):
            combos = [""]
            for c in cols:
                combos = [a + ("_" if a else "") + b for a in combos for b in c[i]]
            rows.append([HashFunc.murmurhash3_32(s, 0) % self.n for s in combos])
        return _sparse_from_rows(rows)


class IndicatorCol(Operation):
    """Sparse category ids → multi-hot (or count) dense matrix of width ``fea_len``."""

    def __init__(self, fea_len: int, is_count: bool = True):
        super().__init__()
        self.n, self.count = fea_len, is_count

    def updateOutput(self, input):
        sp = input.sp if hasattr(input, "sp") else input.to_sparse()
        rows = sp.indices()[0]
        vals = sp.values().long()
        out = torch.zeros(sp.shape[0], self.n)
        out.index_put_((rows, vals), torch.ones(vals.numel()), accumulate=True)
        return out if self.count else (out > 0).float()


class Kv2Tensor(Operation):
    """"k:v,k:v" strings → dense or sparse [batch, fea_len] (Kv2Tensor.scala)."""

    def __init__(self, kv_delimiter: str = ",", item_delimiter: str = ":", trans_type: int = 0):
        super().__init__()
        self.kvd, self.itd, self.trans = kv_delimiter, item_delimiter, trans_type

    def updateOutput(self, input):
        col, fea_len = _ins(input, 2)
        n = int(torch.as_tensor(fea_len).item())
        rows = np.asarray(col, dtype=object).reshape(-1)
        dense = torch.zeros(len(rows), n)
        for i, s in enumerate(rows):
            for kv in str(s).split(self.kvd):
                if kv:
                    k, v = kv.split(self.itd)
                    dense[i, int(k)] = float(v)
        if self.trans == 0:
            return dense
        from ..tensor.sparse import SparseTensor

        return SparseTensor.from_dense(dense)


class MkString(Operation):
    """Sparse/dense row → delimiter-joined string per row (MkString.scala)."""

    def __init__(self, str_delimiter: str = ","):
        super().__init__()
        self.delim = str_delimiter

    def updateOutput(self, input):
        if hasattr(input, "sp"):
            sp = input.sp
            rows = [[] for _ in range(sp.shape[0])]
            for r, v in zip(sp.indices()[0].tolist(), sp.values().tolist()):
                rows[r].append(v)
        else:
            rows = input.tolist()
        fmt = lambda v: str(int(v)) if float(v).is_integer() else str(v)
        return np.array([self.delim.join(fmt(v) for v in r) for r in rows], dtype=object)


def __getattr__(name):
    """``nn.ops.{Conv3DBackprop*, Dilation2DBackprop*, DepthwiseConv2DBackprop*, MaxPoolGrad, ResizeBilinearGrad,
    SelectTensor, …}`` are defined in ``nn.tf_grad`` next to the other TensorFlow-layout operations."""
    from . import tf_grad

    try:
        return getattr(tf_grad, name)
    except AttributeError:
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}") from None
