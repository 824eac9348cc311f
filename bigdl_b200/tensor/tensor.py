"""Torch7-style tensor API with 1-based indexing on top of ``torch.Tensor``.

Reference: S/tensor/Tensor.scala (trait + companion factories), S/tensor/TensorMath.scala (add/cmul/addmm/…),
S/tensor/DenseTensor*.scala (strided implementation, BLAS dispatch), S/tensor/Storage.scala / ArrayStorage.scala.
The reference needs its own strided tensor because the JVM has none; here the storage, strides, dtype conversion and the
math kernels are torch's (CPU or CUDA), and this class supplies the reference's *semantics*: 1-based ``narrow / select /
valueAt / setValue / storageOffset``, in-place-by-default math that returns ``self``, the ``(value, tensor)`` overloads of
``add/mul``, ``max(dim)`` returning 1-based indices, ``resize`` reusing storage, and ``apply1/map/zipWith`` closures.
Layers do not go through this class (they use torch tensors directly); it exists for user code and loaders written
against the reference's Tensor API.
"""
from __future__ import annotations

import math
from typing import Callable, Sequence

import numpy as np
import torch

_DTYPES = {"float": torch.float32, "double": torch.float64, "int": torch.int32, "long": torch.int64,
           "short": torch.int16, "char": torch.int8, "bool": torch.bool, "bfloat16": torch.bfloat16}


def _sizes(args) -> list[int]:
    if len(args) == 1 and isinstance(args[0], (list, tuple, torch.Size)):
        return [int(v) for v in args[0]]
    return [int(v) for v in args]


def _raw(x):
    return x.t if isinstance(x, Tensor) else x


class Storage:
    """Flat view of a tensor's backing memory (S/tensor/Storage.scala: ``array()``, ``length``, ``apply/update``)."""

    def __init__(self, data):
        if isinstance(data, torch.Tensor):
            self.data = data.reshape(-1) if data.is_contiguous() else data.contiguous().reshape(-1)
        elif isinstance(data, int):
            self.data = torch.zeros(data)
        else:
            self.data = torch.as_tensor(np.asarray(data))

    def length(self) -> int:
        return self.data.numel()

    __len__ = length

    def array(self) -> torch.Tensor:
        return self.data

    def __getitem__(self, i: int):
        return self.data[i].item()

    def __setitem__(self, i: int, v):
        self.data[i] = v

    def fill(self, v, offset: int = 1, length: int | None = None):
        n = self.length() - offset + 1 if length is None else length
        self.data[offset - 1:offset - 1 + n] = v
        return self

    def copy(self, other: "Storage"):
        self.data.copy_(other.data)
        return self


def _native_math(*ts):
    """The native fp32/fp64 BLAS + VML module when every operand qualifies (see tensor/mkl.py), else None."""
    from . import mkl
    t0 = ts[0]
    for t in ts:
        if not mkl.enabled_for(t) or t.dtype != t0.dtype or t.device != t0.device:
            return None
    return mkl


class Tensor:
    __slots__ = ("t",)

    # ------------------------------------------------------------------ construction
    def __init__(self, *args, dtype: str | torch.dtype = "float", device=None):
        dt = _DTYPES.get(dtype, dtype) if isinstance(dtype, str) else dtype
        if len(args) == 0:
            self.t = torch.empty(0, dtype=dt, device=device)
        elif isinstance(args[0], torch.Tensor):
            self.t = args[0]
        elif isinstance(args[0], Tensor):
            self.t = args[0].t
        elif isinstance(args[0], np.ndarray):
            self.t = torch.from_numpy(np.ascontiguousarray(args[0]))
        elif isinstance(args[0], Storage):
            st = args[0]
            off = int(args[1]) if len(args) > 1 else 1
            sizes = _sizes(args[2:3]) if len(args) > 2 else [st.length() - off + 1]
            strides = _sizes(args[3:4]) if len(args) > 3 else None
            if strides is None:
                strides, acc = [], 1
                for s in reversed(sizes):
                    strides.insert(0, acc)
                    acc *= s
            self.t = torch.as_strided(st.data, sizes, strides, off - 1)
        elif isinstance(args[0], (list, tuple)) and args[0] and not isinstance(args[0][0], int):
            self.t = torch.tensor(args[0], dtype=dt, device=device)
        else:
            self.t = torch.zeros(_sizes(args), dtype=dt, device=device)

    @staticmethod
    def scalar(v, dtype="float") -> "Tensor":
        return Tensor(torch.tensor(v, dtype=_DTYPES.get(dtype, dtype)))

    @staticmethod
    def ones(*sizes) -> "Tensor":
        return Tensor(torch.ones(_sizes(sizes)))

    @staticmethod
    def zeros(*sizes) -> "Tensor":
        return Tensor(torch.zeros(_sizes(sizes)))

    @staticmethod
    def range(xmin, xmax, step=1) -> "Tensor":
        n = int(math.floor((xmax - xmin) / step)) + 1
        return Tensor(torch.arange(n, dtype=torch.float32) * step + xmin)

    @staticmethod
    def randperm(n: int) -> "Tensor":
        from ..utils.random_generator import RNG

        return Tensor(torch.as_tensor(RNG.randperm(n) + 1, dtype=torch.float32))     # 1-based like Torch

    @staticmethod
    def sparse(indices, values, shape) -> "SparseTensor":
        from .sparse import SparseTensor

        return SparseTensor(indices, values, shape)

    @staticmethod
    def dense(sp) -> "Tensor":
        return sp.to_dense()

    # ------------------------------------------------------------------ shape / storage
    def nDimension(self) -> int:
        return 0 if self.t.numel() == 0 and self.t.dim() == 1 else self.t.dim()

    dim = nDimension

    def size(self, d: int | None = None):
        if d is None:
            return list(self.t.shape)
        return self.t.shape[d - 1]

    def stride(self, d: int | None = None):
        if d is None:
            return list(self.t.stride())
        return self.t.stride(d - 1)

    def nElement(self) -> int:
        return self.t.numel()

    def isEmpty(self) -> bool:
        return self.t.numel() == 0

    def isScalar(self) -> bool:
        return self.t.dim() == 0

    def isContiguous(self) -> bool:
        return self.t.is_contiguous()

    def isSameSizeAs(self, other) -> bool:
        return tuple(self.t.shape) == tuple(_raw(other).shape)

    def storage(self) -> Storage:
        base = self.t._base if self.t._base is not None else self.t
        s = Storage.__new__(Storage)
        s.data = torch.as_strided(base, (base.untyped_storage().nbytes() // base.element_size(),), (1,), 0)
        return s

    def storageOffset(self) -> int:
        return self.t.storage_offset() + 1

    def contiguous(self) -> "Tensor":
        return Tensor(self.t.contiguous())

    def clone(self) -> "Tensor":
        return Tensor(self.t.clone())

    def getType(self) -> torch.dtype:
        return self.t.dtype

    def toTensor(self):
        return self

    # ---- Activity protocol (S/nn/abstractnn/Activity.scala): a Tensor is an activity that is not a Table
    def isTensor(self) -> bool:
        return True

    def isTable(self) -> bool:
        return False

    def toTable(self):
        raise TypeError("Tensor cannot be cast to Table")

    def unary_(self) -> "Tensor":
        """``-tensor`` (TensorMath.unary_-); the operator form is ``-t``."""
        return Tensor(-self.t)

    def numpy(self) -> np.ndarray:
        return self.t.detach().cpu().numpy()

    def value(self):
        assert self.t.numel() == 1
        return self.t.item()

    # ------------------------------------------------------------------ views (1-based)
    def narrow(self, dim: int, index: int, size: int) -> "Tensor":
        return Tensor(self.t.narrow(dim - 1, index - 1, size))

    def select(self, dim: int, index: int) -> "Tensor":
        return Tensor(self.t.select(dim - 1, index - 1))

    def __call__(self, *idx):
        """``t(i)`` selects along dim 1 (1-based); ``t(i, j, …)`` with as many indices as dims returns the value."""
        if len(idx) == self.t.dim():
            return self.valueAt(*idx)
        out = self
        for i in idx:
            out = out.select(1, i)
        return out

    def valueAt(self, *idx):
        return self.t[tuple(i - 1 for i in idx)].item()

    def setValue(self, *args):
        *idx, v = args
        self.t[tuple(i - 1 for i in idx)] = v
        return self

    def view(self, *sizes) -> "Tensor":
        return Tensor(self.t.view(_sizes(sizes)))

    def reshape(self, *sizes) -> "Tensor":
        return Tensor(self.t.reshape(_sizes(sizes)).clone())

    def transpose(self, d1: int, d2: int) -> "Tensor":
        return Tensor(self.t.transpose(d1 - 1, d2 - 1))

    def t_(self) -> "Tensor":
        assert self.t.dim() == 2
        return Tensor(self.t.t())

    def squeeze(self, dim: int | None = None) -> "Tensor":
        self.t = self.t.squeeze() if dim is None else self.t.squeeze(dim - 1)
        return self

    def squeezeNewTensor(self) -> "Tensor":
        return Tensor(self.t.squeeze())

    def addSingletonDimension(self, dim: int = 1) -> "Tensor":
        self.t = self.t.unsqueeze(dim - 1)
        return self

    def unfold(self, dim: int, size: int, step: int) -> "Tensor":
        return Tensor(self.t.unfold(dim - 1, size, step))

    def expand(self, *sizes) -> "Tensor":
        return Tensor(self.t.expand(_sizes(sizes)))

    def expandAs(self, other) -> "Tensor":
        return Tensor(self.t.expand_as(_raw(other)))

    def repeatTensor(self, *sizes) -> "Tensor":
        return Tensor(self.t.repeat(_sizes(sizes)))

    def split(self, size: int, dim: int = 1) -> list["Tensor"]:
        return [Tensor(p) for p in self.t.split(size, dim - 1)]

    def index(self, dim: int, index) -> "Tensor":
        idx = (_raw(index).long() - 1)
        return Tensor(self.t.index_select(dim - 1, idx))

    def indexAdd(self, dim: int, index, src):
        self.t.index_add_(dim - 1, _raw(index).long() - 1, _raw(src))
        return self

    def resize(self, *sizes) -> "Tensor":
        """Reuse the storage when it is large enough (contents preserved in storage order), else reallocate."""
        s = _sizes(sizes)
        n = int(np.prod(s)) if s else 1
        if self.t.is_contiguous() and self.t.numel() == n:
            self.t = self.t.view(s)
        else:
            new = torch.zeros(s, dtype=self.t.dtype, device=self.t.device)
            k = min(n, self.t.numel())
            if k:
                new.view(-1)[:k] = self.t.contiguous().view(-1)[:k]
            self.t = new
        return self

    def resizeAs(self, other) -> "Tensor":
        return self.resize(*_raw(other).shape)

    def set(self, other=None, storage_offset: int = 1, sizes=None, strides=None) -> "Tensor":
        if other is None:
            self.t = torch.empty(0, dtype=self.t.dtype)
        elif isinstance(other, Storage):
            self.t = Tensor(other, storage_offset, sizes or [other.length() - storage_offset + 1], strides).t
        else:
            self.t = _raw(other)
        return self

    # ------------------------------------------------------------------ fill / copy / closures
    def fill(self, v) -> "Tensor":
        self.t.fill_(v)
        return self

    def zero(self) -> "Tensor":
        self.t.zero_()
        return self

    def copy(self, other) -> "Tensor":
        src = _raw(other)
        if src.shape != self.t.shape:
            src = src.reshape(self.t.shape)        # Torch copies by element count
        self.t.copy_(src)
        return self

    def randn(self, mean: float = 0.0, stdv: float = 1.0) -> "Tensor":
        from ..utils.random_generator import RNG

        RNG.fill_normal(self.t, mean, stdv)
        return self

    def rand(self, lower: float = 0.0, upper: float = 1.0) -> "Tensor":
        from ..utils.random_generator import RNG

        RNG.fill_uniform(self.t, lower, upper)
        return self

    def bernoulli(self, p: float) -> "Tensor":
        from ..utils.random_generator import RNG

        RNG.fill_bernoulli(self.t, p)
        return self

    def apply1(self, fn: Callable) -> "Tensor":
        flat = self.t.reshape(-1)
        vals = [fn(v) for v in flat.tolist()]
        self.t.copy_(torch.tensor(vals, dtype=self.t.dtype).reshape(self.t.shape))
        return self

    def map(self, other, fn: Callable) -> "Tensor":
        a, b = self.t.reshape(-1).tolist(), _raw(other).reshape(-1).tolist()
        self.t.copy_(torch.tensor([fn(x, y) for x, y in zip(a, b)], dtype=self.t.dtype).reshape(self.t.shape))
        return self

    def zipWith(self, t1, t2, fn: Callable) -> "Tensor":
        a, b = _raw(t1).reshape(-1).tolist(), _raw(t2).reshape(-1).tolist()
        self.t = torch.tensor([fn(x, y) for x, y in zip(a, b)], dtype=self.t.dtype).reshape(_raw(t1).shape)
        return self

    # ------------------------------------------------------------------ TensorMath (in place, returns self)
    def add(self, *a) -> "Tensor":
        """add(value) | add(tensor) | add(value, tensor) | add(x, value, y) → self = x + value*y."""
        if len(a) == 1:
            self.t.add_(_raw(a[0]))
        elif len(a) == 2 and not isinstance(a[0], (Tensor, torch.Tensor)):
            self.t.add_(_raw(a[1]), alpha=a[0])
        elif len(a) == 2:
            self._assign(_raw(a[0]) + _raw(a[1]))
        else:
            self._assign(torch.add(_raw(a[0]), _raw(a[2]), alpha=a[1]))
        return self

    def sub(self, *a) -> "Tensor":
        if len(a) == 1:
            self.t.sub_(_raw(a[0]))
        elif len(a) == 2 and not isinstance(a[0], (Tensor, torch.Tensor)):
            self.t.sub_(_raw(a[1]), alpha=a[0])
        elif len(a) == 2:
            self._assign(_raw(a[0]) - _raw(a[1]))
        else:
            self._assign(torch.sub(_raw(a[0]), _raw(a[2]), alpha=a[1]))
        return self

    def mul(self, *a) -> "Tensor":
        if len(a) == 1:
            self.t.mul_(_raw(a[0]))
        else:
            self._assign(_raw(a[0]) * a[1])
        return self

    def div(self, *a) -> "Tensor":
        if len(a) == 1:
            self.t.div_(_raw(a[0]))
        else:
            self._assign(_raw(a[0]) / _raw(a[1]))
        return self

    def cmul(self, *a) -> "Tensor":
        if len(a) == 1:
            self.t.mul_(_raw(a[0]))
        else:
            self._assign(_raw(a[0]) * _raw(a[1]))
        return self

    def cdiv(self, *a) -> "Tensor":
        if len(a) == 1:
            self.t.div_(_raw(a[0]))
        else:
            self._assign(_raw(a[0]) / _raw(a[1]))
        return self

    def cmax(self, *a) -> "Tensor":
        if len(a) == 1:
            o = _raw(a[0])
            self._assign(torch.clamp(self.t, min=o) if not isinstance(o, torch.Tensor) else torch.maximum(self.t, o))
        else:
            self._assign(torch.maximum(_raw(a[0]), _raw(a[1])))
        return self

    def cmin(self, *a) -> "Tensor":
        if len(a) == 1:
            o = _raw(a[0])
            self._assign(torch.clamp(self.t, max=o) if not isinstance(o, torch.Tensor) else torch.minimum(self.t, o))
        else:
            self._assign(torch.minimum(_raw(a[0]), _raw(a[1])))
        return self

    def addcmul(self, *a) -> "Tensor":
        """addcmul(value, t1, t2) | addcmul(t1, t2)."""
        v, t1, t2 = (a if len(a) == 3 else (1.0, *a))
        self.t.addcmul_(_raw(t1), _raw(t2), value=v)
        return self

    def addcdiv(self, v, t1, t2) -> "Tensor":
        self.t.addcdiv_(_raw(t1), _raw(t2), value=v)
        return self

    def addmm(self, *a) -> "Tensor":
        """addmm(m1, m2) | addmm(v2, m1, m2) | addmm(v1, M, v2, m1, m2): self = v1*M + v2*m1@m2."""
        if len(a) in (2, 3):
            alpha, m1, m2 = (1.0, *a) if len(a) == 2 else a
            m1, m2 = _raw(m1), _raw(m2)
            nm = _native_math(self.t, m1, m2)
            if nm is not None and self.t.dim() == 2 and self.t.stride(1) == 1 and m1.dim() == 2 and m2.dim() == 2:
                nm.mm(self.t, m1, m2, alpha=alpha, beta=1.0)
            else:
                self.t.addmm_(m1, m2, alpha=alpha)
        elif len(a) == 4:
            self._assign(torch.addmm(_raw(a[0]) * a[1] if False else _raw(a[0]), _raw(a[2]), _raw(a[3]), beta=1, alpha=a[1]))
        else:
            v1, M, v2, m1, m2 = a
            self._assign(torch.addmm(_raw(M), _raw(m1), _raw(m2), beta=v1, alpha=v2))
        return self

    def addmv(self, *a) -> "Tensor":
        if len(a) in (2, 3):
            alpha, m, x = (1.0, *a) if len(a) == 2 else a
            m, x = _raw(m), _raw(x)
            nm = _native_math(self.t, m, x)
            if nm is not None and self.t.dim() == 1 and m.dim() == 2 and x.dim() == 1:
                nm.mv(self.t, m, x, alpha=alpha, beta=1.0)
            else:
                self.t.addmv_(m, x, alpha=alpha)
        else:
            beta, v, alpha, m, x = a
            self._assign(torch.addmv(_raw(v), _raw(m), _raw(x), beta=beta, alpha=alpha))
        return self

    def addr(self, *a) -> "Tensor":
        if len(a) in (2, 3):
            alpha, x, y = (1.0, *a) if len(a) == 2 else a
            x, y = _raw(x), _raw(y)
            nm = _native_math(self.t, x, y)
            if nm is not None and self.t.dim() == 2 and self.t.stride(1) == 1 and x.dim() == 1 and y.dim() == 1:
                nm.ger(self.t, x, y, alpha=alpha)
            else:
                self.t.addr_(x, y, alpha=alpha)
        else:
            v1, M, v2, x, y = a
            self._assign(torch.addr(_raw(M), _raw(x), _raw(y), beta=v1, alpha=v2))
        return self

    def mm(self, m1, m2) -> "Tensor":
        a, b = _raw(m1), _raw(m2)
        nm = _native_math(a, b)
        if nm is not None and a.dim() == 2 and b.dim() == 2:
            out = self._result_buffer((a.shape[0], b.shape[1]), a)
            nm.mm(out, a, b)                  # sgemm / dgemm of csrc/blas_vml.cu (N1)
            return self
        self._assign(a @ b)
        return self

    def mv(self, m, v) -> "Tensor":
        a, x = _raw(m), _raw(v)
        nm = _native_math(a, x)
        if nm is not None and a.dim() == 2 and x.dim() == 1:
            out = self._result_buffer((a.shape[0],), a)
            nm.mv(out, a, x)
            return self
        self._assign(a @ x)
        return self

    def _result_buffer(self, shape, like: torch.Tensor) -> torch.Tensor:
        """self.t resized to ``shape`` (contiguous, dtype/device of ``like``) for a native kernel to write into."""
        if (tuple(self.t.shape) != tuple(shape) or self.t.dtype != like.dtype or self.t.device != like.device
                or not self.t.is_contiguous()):
            self.t = torch.empty(shape, dtype=like.dtype, device=like.device)
        return self.t

    def bmm(self, b1, b2) -> "Tensor":
        self._assign(torch.bmm(_raw(b1), _raw(b2)))
        return self

    def baddbmm(self, beta, M, alpha, b1, b2) -> "Tensor":
        self._assign(torch.baddbmm(_raw(M), _raw(b1), _raw(b2), beta=beta, alpha=alpha))
        return self

    def dot(self, other) -> float:
        o = _raw(other)
        nm = _native_math(self.t, o)
        if nm is not None and self.t.numel() == o.numel():
            return nm.dot(self.t, o)
        return float((self.t.reshape(-1) * o.reshape(-1)).sum())

    def _assign(self, val: torch.Tensor):
        if tuple(val.shape) == tuple(self.t.shape) and val.dtype == self.t.dtype:
            self.t.copy_(val)
        else:
            self.t = val.to(self.t.dtype) if self.t.numel() else val

    _VML = {torch.sqrt: "Sqrt", torch.log: "Ln", torch.log1p: "Log1p", torch.exp: "Exp", torch.tanh: "Tanh",
            torch.abs: "Abs"}

    def _unary(self, fn, src=None) -> "Tensor":
        op = Tensor._VML.get(fn)
        if op is not None:
            s = self.t if src is None else _raw(src)
            nm = _native_math(s)
            if nm is not None and s.is_contiguous() and s.numel():
                out = self.t if src is None else self._result_buffer(tuple(s.shape), s)
                if out.is_contiguous():
                    nm.unary(op, out, s)          # vsSqrt / vdLn ... of the native vector-math library (N1)
                    return self
        if src is None:
            self.t.copy_(fn(self.t))
        else:
            self._assign(fn(_raw(src)))
        return self

    def pow(self, *a) -> "Tensor":
        return self._unary(lambda x: x.pow(a[-1]), a[0] if len(a) == 2 else None)

    def sqrt(self, src=None): return self._unary(torch.sqrt, src)
    def log(self, src=None): return self._unary(torch.log, src)
    def log1p(self, src=None): return self._unary(torch.log1p, src)
    def exp(self, src=None): return self._unary(torch.exp, src)
    def tanh(self, src=None): return self._unary(torch.tanh, src)
    def sigmoid(self, src=None): return self._unary(torch.sigmoid, src)
    def abs(self, src=None): return self._unary(torch.abs, src)
    def sign(self, src=None): return self._unary(torch.sign, src)
    def floor(self, src=None): return self._unary(torch.floor, src)
    def ceil(self, src=None): return self._unary(torch.ceil, src)
    def negative(self, src=None): return self._unary(torch.neg, src)
    def inv(self, src=None): return self._unary(torch.reciprocal, src)
    def square(self, src=None): return self._unary(torch.square, src)

    def clamp(self, lo, hi) -> "Tensor":
        self.t.clamp_(lo, hi)
        return self

    # ------------------------------------------------------------------ reductions
    def sum(self, dim: int | None = None):
        if dim is None:
            return self.t.sum().item()
        return Tensor(self.t.sum(dim - 1, keepdim=True))

    def mean(self, dim: int | None = None):
        if dim is None:
            return self.t.float().mean().item()
        return Tensor(self.t.mean(dim - 1, keepdim=True))

    def prod(self, dim: int | None = None):
        if dim is None:
            return self.t.prod().item()
        return Tensor(self.t.prod(dim - 1, keepdim=True))

    def max(self, dim: int | None = None):
        """max() → value; max(dim) → (values, 1-based indices), both keeping the reduced dim."""
        if dim is None:
            return self.t.max().item()
        v, i = self.t.max(dim - 1, keepdim=True)
        return Tensor(v), Tensor((i + 1).to(self.t.dtype))

    def min(self, dim: int | None = None):
        if dim is None:
            return self.t.min().item()
        v, i = self.t.min(dim - 1, keepdim=True)
        return Tensor(v), Tensor((i + 1).to(self.t.dtype))

    def topk(self, k: int, dim: int = -1, increase: bool = True):
        d = self.t.dim() - 1 if dim == -1 else dim - 1
        v, i = self.t.topk(k, d, largest=not increase)
        return Tensor(v), Tensor((i + 1).to(self.t.dtype))

    def sort(self, dim: int = -1, descending: bool = False):
        d = self.t.dim() - 1 if dim == -1 else dim - 1
        v, i = self.t.sort(d, descending=descending)
        return Tensor(v), Tensor((i + 1).to(self.t.dtype))

    def norm(self, p: float = 2) -> float:
        return float(self.t.double().norm(p))

    def sumSquare(self) -> float:
        return float(self.t.double().pow(2).sum())

    def std(self, dim: int | None = None):
        return self.t.std().item() if dim is None else Tensor(self.t.std(dim - 1, keepdim=True))

    def var(self, dim: int | None = None):
        return self.t.var().item() if dim is None else Tensor(self.t.var(dim - 1, keepdim=True))

    def dist(self, other, p: float = 2) -> float:
        return float(torch.dist(self.t, _raw(other), p))

    def numNonZeroByRow(self) -> list[int]:
        return (self.t != 0).sum(1).tolist()

    # ------------------------------------------------------------------ comparison / masks / gather
    def gt(self, a, b=None): return self._cmp(torch.gt, a, b)
    def ge(self, a, b=None): return self._cmp(torch.ge, a, b)
    def lt(self, a, b=None): return self._cmp(torch.lt, a, b)
    def le(self, a, b=None): return self._cmp(torch.le, a, b)
    def eq(self, a, b=None): return self._cmp(torch.eq, a, b)

    def _cmp(self, fn, a, b):
        if b is None:
            return Tensor(fn(self.t, _raw(a)).to(self.t.dtype))
        self._assign(fn(_raw(a), _raw(b)).to(self.t.dtype))
        return self

    def almostEqual(self, other, delta: float) -> bool:
        o = _raw(other)
        return tuple(o.shape) == tuple(self.t.shape) and bool(((self.t - o).abs() <= delta).all())

    def __eq__(self, other):
        if not isinstance(other, Tensor):
            return NotImplemented
        return tuple(self.t.shape) == tuple(other.t.shape) and bool(torch.equal(self.t, other.t))

    __hash__ = None

    def maskedFill(self, mask, v) -> "Tensor":
        self.t.masked_fill_(_raw(mask).bool(), v)
        return self

    def maskedSelect(self, mask, res: "Tensor | None" = None) -> "Tensor":
        out = self.t.masked_select(_raw(mask).bool())
        if res is not None:
            res.t = out
            return res
        return Tensor(out)

    def maskedCopy(self, mask, src) -> "Tensor":
        self.t.masked_scatter_(_raw(mask).bool(), _raw(src))
        return self

    def gather(self, dim: int, index, src) -> "Tensor":
        self._assign(torch.gather(_raw(src), dim - 1, _raw(index).long() - 1))
        return self

    def scatter(self, dim: int, index, src) -> "Tensor":
        self.t.scatter_(dim - 1, _raw(index).long() - 1, _raw(src))
        return self

    def conv2(self, kernel, vf: str = "V") -> "Tensor":
        """2-D 'valid' or 'full' convolution of a matrix (S/tensor/DenseTensorConv.scala)."""
        k = _raw(kernel).flip(0, 1)[None, None]
        pad = (k.shape[2] - 1, k.shape[3] - 1) if vf.upper() == "F" else (0, 0)
        return Tensor(torch.nn.functional.conv2d(self.t[None, None], k, padding=pad)[0, 0])

    def xcorr2(self, kernel, vf: str = "V") -> "Tensor":
        k = _raw(kernel)[None, None]
        pad = (k.shape[2] - 1, k.shape[3] - 1) if vf.upper() == "F" else (0, 0)
        return Tensor(torch.nn.functional.conv2d(self.t[None, None], k, padding=pad)[0, 0])

    # ------------------------------------------------------------------ operators (out of place, as in the Scala DSL)
    def __add__(self, o): return Tensor(self.t + _raw(o))
    def __sub__(self, o): return Tensor(self.t - _raw(o))
    def __mul__(self, o):
        r = _raw(o)
        if isinstance(r, torch.Tensor) and r.dim() == 2 and self.t.dim() == 2:
            return Tensor(self.t @ r)                  # Scala `*` on matrices is a matrix product
        if isinstance(r, torch.Tensor) and r.dim() == 1 and self.t.dim() == 2:
            return Tensor(self.t @ r)
        return Tensor(self.t * r)
    def __truediv__(self, o): return Tensor(self.t / _raw(o))
    def __neg__(self): return Tensor(-self.t)

    # ------------------------------------------------------------------ rest of the Tensor / TensorMath surface
    def _subset(self, spec):
        """View selected by a Table / sequence of per-dimension entries (S/tensor/Tensor.scala:196-234): an int selects
        (1-based; negative counts from the end), a 2-element entry ``(start, end)`` narrows inclusively, an empty entry
        keeps the dimension."""
        from ..utils.table import Table
        items = spec.to_list() if isinstance(spec, Table) else list(spec)
        out, dim = self.t, 0
        for it in items:
            if isinstance(it, Table):
                it = it.to_list()
            if isinstance(it, (list, tuple)):
                if len(it) == 0:
                    dim += 1
                    continue
                n = out.shape[dim]
                a = it[0] if it[0] > 0 else n + it[0] + 1
                b = (it[1] if it[1] > 0 else n + it[1] + 1) if len(it) > 1 else n
                out = out.narrow(dim, a - 1, b - a + 1)
                dim += 1
            else:
                i = int(it)
                i = i if i > 0 else out.shape[dim] + i + 1
                out = out.select(dim, i - 1)
        return out

    def apply(self, index):
        """``apply(i)`` = select along dim 1; ``apply([i, j, ...])`` with one index per dim = the value;
        ``apply(Table(...))`` = sub-tensor view."""
        from ..utils.table import Table
        if isinstance(index, int):
            return self.select(1, index if index > 0 else self.t.shape[0] + index + 1) if self.t.dim() > 1 \
                else Tensor(self.t[index - 1 if index > 0 else index])
        if isinstance(index, Table) or any(isinstance(i, (list, tuple, Table)) for i in index):
            return Tensor(self._subset(index))
        return self.valueAt(*index)

    def update(self, index, value):
        """``update(i, v | tensor)``, ``update([i, j, ...], v)``, ``update(Table(...), v | tensor)``,
        ``update(predicate, v)`` (S/tensor/Tensor.scala:236-319)."""
        from ..utils.table import Table
        if callable(index):
            flat = self.t.reshape(-1)
            mask = torch.tensor([bool(index(x)) for x in flat.tolist()], dtype=torch.bool, device=flat.device)
            self.t.masked_fill_(mask.reshape(self.t.shape), value)
            return self
        if isinstance(index, int):
            target = self.t.select(0, index - 1 if index > 0 else index)
        elif isinstance(index, Table) or any(isinstance(i, (list, tuple, Table)) for i in index):
            target = self._subset(index)
        else:
            self.setValue(*index, value)
            return self
        if isinstance(value, (Tensor, torch.Tensor)):
            target.copy_(_raw(value))
        else:
            target.fill_(value)
        return self

    def applyFun(self, other, func) -> "Tensor":
        """self[i] = func(other[i]) over same-sized tensors (possibly of another numeric type)."""
        src = _raw(other).reshape(-1).tolist()
        self.t.copy_(torch.tensor([func(x) for x in src], dtype=self.t.dtype).reshape(self.t.shape))
        return self

    def cast(self, castTensor: "Tensor") -> "Tensor":
        """Copy this tensor's values into ``castTensor`` converted to its type; returns ``castTensor``."""
        castTensor.t = self.t.to(castTensor.t.dtype).clone() if castTensor.t.shape != self.t.shape else \
            castTensor.t.copy_(self.t.to(castTensor.t.dtype))
        return castTensor

    def diff(self, other, count: int = 1, reverse: bool = False) -> bool:
        """Print up to ``count`` differing elements; True when the tensors differ (debug helper)."""
        o = _raw(other)
        if tuple(o.shape) != tuple(self.t.shape):
            print(f"size mismatch: {tuple(self.t.shape)} vs {tuple(o.shape)}")
            return True
        idx = (self.t != o).reshape(-1).nonzero().reshape(-1).tolist()
        for i in (reversed(idx) if reverse else idx)[:count] if not reverse else list(reversed(idx))[:count]:
            print(f"Find difference at offset {i + 1}: {self.t.reshape(-1)[i].item()} vs {o.reshape(-1)[i].item()}")
        return bool(idx)

    def erf(self): self.t.copy_(torch.erf(self.t)); return self
    def erfc(self): self.t.copy_(torch.erfc(self.t)); return self
    def logGamma(self): self.t.copy_(torch.lgamma(self.t)); return self
    def digamma(self): self.t.copy_(torch.digamma(self.t)); return self

    def emptyInstance(self) -> "Tensor":
        return Tensor(torch.empty(0, dtype=self.t.dtype, device=self.t.device))

    def forceFill(self, v) -> "Tensor":
        self.t.fill_(v)
        return self

    def shallowClone(self) -> "Tensor":
        """A new Tensor object over the same storage."""
        return Tensor(self.t.view(self.t.shape))

    def getTensorType(self) -> str:
        return "DenseType"

    def getTensorNumeric(self):
        return self.t.dtype

    def notEqualValue(self, value) -> bool:
        return bool((self.t != value).any())

    def reduce(self, dim: int, result: "Tensor", reducer) -> "Tensor":
        """Fold ``reducer(acc, x)`` along ``dim`` (1-based) into ``result`` (that dimension becomes size 1)."""
        d = dim - 1
        slices = self.t.unbind(d)
        acc = slices[0].clone()
        for sl in slices[1:]:
            acc = torch.tensor([reducer(a, b) for a, b in zip(acc.reshape(-1).tolist(), sl.reshape(-1).tolist())],
                               dtype=self.t.dtype, device=self.t.device).reshape(acc.shape)
        result._assign(acc.unsqueeze(d))
        return result

    def toArray(self) -> list:
        if self.t.dim() != 1:
            raise ValueError("toArray needs a 1-D tensor")
        return self.t.tolist()

    def toBreezeVector(self):
        if self.t.dim() != 1:
            raise ValueError("tensor is not a vector")
        return self.t.detach().cpu().numpy().copy()

    def toBreezeMatrix(self):
        if self.t.dim() != 2:
            raise ValueError("tensor is not a matrix")
        return self.t.detach().cpu().numpy().copy()

    toMLlibVector, toMLlibMatrix = toBreezeVector, toBreezeMatrix

    def uniform(self, *a) -> float:
        """A uniform random number from the shared generator: ``uniform()`` in [0,1), ``uniform(lo, hi)``."""
        from ..utils.random_generator import RNG
        lo, hi = (0.0, 1.0) if not a else a
        return float(RNG.uniform(lo, hi))

    def unique(self):
        """(distinct values in order of first appearance, for every element the 0-based index of its value) — TensorMath
        .unique for 1-D tensors."""
        vals = self.t.reshape(-1).tolist()
        seen, order, idx = {}, [], []
        for v in vals:
            if v not in seen:
                seen[v] = len(order)
                order.append(v)
            idx.append(seen[v])
        return (Tensor(torch.tensor(order, dtype=self.t.dtype)), Tensor(torch.tensor(idx, dtype=torch.int32)))

    def addMultiDimension(self, t: "Tensor", dims) -> "Tensor":
        """self += t summed over ``dims`` (1-based) — the broadcasting-bias gradient helper."""
        r = _raw(t)
        for d in sorted((d - 1 for d in dims), reverse=True):
            r = r.sum(d, keepdim=True)
        self.t.add_(r.reshape(self.t.shape))
        return self

    def save(self, path: str, overWrite: bool = False) -> "Tensor":
        from ..utils.file import File
        File.save(self.t.detach().cpu(), path, overWrite)
        return self

    @staticmethod
    def load(path: str) -> "Tensor":
        from ..utils.file import File
        return Tensor(File.load(path))

    @staticmethod
    def gaussian1D(size: int = 3, sigma: float = 0.25, amplitude: int = 1, normalize: bool = False,
                   mean: float = 0.5, tensor: "Tensor | None" = None) -> "Tensor":
        """1-D Gaussian kernel (S/tensor/Tensor.scala:1208-1233): value_i = amplitude * exp(-((i/size - mean)/sigma)^2 / 2)
        for i = 1..size, centred per ``mean`` in units of the size."""
        i = torch.arange(1, size + 1, dtype=torch.float32)
        center = mean * size + 0.5
        g = amplitude * torch.exp(-(((i - center) / (sigma * size)) ** 2) / 2)
        if normalize:
            g = g / g.sum()
        if tensor is not None:
            tensor._assign(g)
            return tensor
        return Tensor(g)

    @staticmethod
    def getDims(t) -> int:
        return _raw(t).dim()

    # ------------------------------------------------------------------ conversions
    def toQuantizedTensor(self):
        from .quantized import QuantizedTensor

        return QuantizedTensor.from_float(self.t)

    def cuda(self):
        return Tensor(self.t.cuda())

    def cpu(self):
        return Tensor(self.t.cpu())

    def __repr__(self):
        return f"{self.t}\n[bigdl_b200.tensor.Tensor of size {'x'.join(map(str, self.t.shape))}]"


def cat(tensors: Sequence[Tensor], dim: int = 1) -> Tensor:
    return Tensor(torch.cat([_raw(t) for t in tensors], dim - 1))
