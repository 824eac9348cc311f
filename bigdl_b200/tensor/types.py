"""The type tags and small helper objects of ``com.intel.analytics.bigdl.tensor`` (S/tensor/Tensor.scala:823-851,
Convertable.scala, TensorFunc.scala, DenseTensor{BLAS,Conv,Apply,DimApply,Math}.scala).

A tensor here is a ``torch.Tensor`` wrapped by ``tensor.Tensor``; the element type is a ``torch.dtype`` and storage is
torch's. These names give reference code its vocabulary: ``FloatType`` compares equal to ``torch.float32``,
``DenseTensorBLAS.gemm`` is the column-major GEMM the reference calls MKL for (routed to the library's cuBLAS /
OpenBLAS entry points by ``tensor.mkl``), ``DenseTensorConv.conv2Dmul`` the 2-D convolution behind
``Tensor.conv2`` / ``xcorr2``."""
from __future__ import annotations

import torch


class TensorDataType:
    """Element type tag; equal to the matching ``torch.dtype`` (or to ``str`` for ``StringType``)."""

    def __init__(self, name: str, dtype):
        self.name, self.dtype = name, dtype

    def __eq__(self, o):
        return o is self or (isinstance(o, TensorDataType) and o.name == self.name) or o == self.dtype

    def __hash__(self):
        return hash(self.name)

    def __repr__(self):
        return self.name

    @staticmethod
    def of(dtype) -> "TensorDataType":
        for t in _ALL_TYPES:
            if t.dtype == dtype:
                return t
        raise TypeError(f"no TensorDataType for {dtype}")


BooleanType = TensorDataType("BooleanType", torch.bool)
CharType = TensorDataType("CharType", torch.uint8)        # no 16-bit char in torch; code points fit callers' use
ByteType = TensorDataType("ByteType", torch.int8)
StringType = TensorDataType("StringType", str)
IntType = TensorDataType("IntType", torch.int32)
ShortType = TensorDataType("ShortType", torch.int16)
LongType = TensorDataType("LongType", torch.int64)
FloatType = TensorDataType("FloatType", torch.float32)
DoubleType = TensorDataType("DoubleType", torch.float64)
_ALL_TYPES = (BooleanType, CharType, ByteType, StringType, IntType, ShortType, LongType, FloatType, DoubleType)


class TensorType:
    """Storage kind tag; equal to the string ``getTensorType()`` returns."""

    def __init__(self, name: str):
        self.name = name

    def __eq__(self, o):
        return o is self or (isinstance(o, TensorType) and o.name == self.name) or o == self.name

    def __hash__(self):
        return hash(self.name)

    def __repr__(self):
        return self.name


DenseType, SparseType, QuantizedType, MklDnnType = (TensorType("DenseType"), TensorType("SparseType"),
                                                    TensorType("QuantizedType"), TensorType("MklDnnType"))


# ------------------------------------------------------------------------------------------------ converters
class ConvertableTo:
    """``ConvertableTo[A].fromX(v)``: numeric conversion INTO a fixed element type (Convertable.scala)."""

    def __init__(self, cast, name):
        self._cast, self.name = cast, name

    def _from(self, v):
        return self._cast(v)

    fromFloat = fromDouble = fromInt = fromShort = fromLong = fromBoolean = fromChar = fromByte = _from

    def fromString(self, s: str):
        return self._cast(float(s))

    def __repr__(self):
        return f"ConvertableTo{self.name}"


def _to_int(bits):
    lo, hi = -(1 << (bits - 1)), (1 << (bits - 1)) - 1

    def cast(v):
        i = int(v)                               # truncation toward zero, like the JVM's narrowing from floating point
        if isinstance(v, float):
            return max(lo, min(hi, i))
        return ((i - lo) % (1 << bits)) + lo     # integer narrowing wraps
    return cast


ConvertableToFloat = ConvertableTo(lambda v: float(torch.tensor(float(v), dtype=torch.float32)), "Float")
ConvertableToDouble = ConvertableTo(float, "Double")
ConvertableToInt = ConvertableTo(_to_int(32), "Int")
ConvertableToShort = ConvertableTo(_to_int(16), "Short")
ConvertableToLong = ConvertableTo(_to_int(64), "Long")


class ConvertableFrom:
    """``ConvertableFrom[A].toX(a)``: conversion OUT of a fixed element type."""

    def __init__(self, name):
        self.name = name

    toFloat = staticmethod(lambda a: ConvertableToFloat._from(a))
    toDouble = staticmethod(lambda a: float(a))
    toInt = staticmethod(lambda a: ConvertableToInt._from(a))
    toShort = staticmethod(lambda a: ConvertableToShort._from(a))
    toLong = staticmethod(lambda a: ConvertableToLong._from(a))
    toByte = staticmethod(lambda a: _to_int(8)(a))
    toBoolean = staticmethod(lambda a: bool(a))
    toChar = staticmethod(lambda a: chr(int(a)))
    toString = staticmethod(lambda a: str(a))

    def __repr__(self):
        return f"ConvertableFrom{self.name}"


(ConvertableFromFloat, ConvertableFromDouble, ConvertableFromInt, ConvertableFromShort, ConvertableFromLong,
 ConvertableFromBoolean, ConvertableFromChar, ConvertableFromByte) = (ConvertableFrom(n) for n in (
     "Float", "Double", "Int", "Short", "Long", "Boolean", "Char", "Byte"))


class _FromString(ConvertableFrom):
    toFloat = staticmethod(lambda a: ConvertableToFloat._from(float(a)))
    toDouble = staticmethod(lambda a: float(a))
    toInt = staticmethod(lambda a: int(float(a)))
    toLong = staticmethod(lambda a: int(float(a)))
    toShort = staticmethod(lambda a: ConvertableToShort._from(int(float(a))))
    toBoolean = staticmethod(lambda a: str(a).strip().lower() == "true")


ConvertableFromString = _FromString("String")


# ------------------------------------------------------------------------------------------------ function types
class TensorFunc2:
    """``apply(data, offset)``: visit one element in place. ``Tensor.apply1``-style callers pass plain callables
    here; subclasses or ``TensorFunc2(fn)`` both work."""

    def __init__(self, fn=None):
        self.fn = fn

    def apply(self, *a):
        return self.fn(*a)

    __call__ = apply


class TensorFunc4(TensorFunc2):
    """``apply(data1, offset1, data2, offset2)``"""


class TensorFunc6(TensorFunc2):
    """``apply(data1, offset1, data2, offset2, data3, offset3)``"""


TensorDiffTypeFunc4 = TensorFunc4
TensorDiffTypeFunc6 = TensorFunc6


# ------------------------------------------------------------------------------------------------ dense helpers
def _raw(t):
    return t.t if hasattr(t, "t") and isinstance(getattr(t, "t"), torch.Tensor) else t


class DenseTensorBLAS:
    """Column-major BLAS on tensors (DenseTensorBLAS.scala:70-240)."""

    @staticmethod
    def gemm(transa: str, transb: str, m: int, n: int, k: int, alpha, a, lda: int, b, ldb: int, beta, c, ldc: int):
        """C[m×n] = alpha · op(A) · op(B) + beta · C on flat column-major buffers with leading dimensions."""
        A, B, C = _raw(a).reshape(-1), _raw(b).reshape(-1), _raw(c).reshape(-1)

        def view(buf, rows, cols, ld, trans):
            r, cdim = (cols, rows) if trans else (rows, cols)
            mat = torch.as_strided(buf, (r, cdim), (1, ld))
            return mat.t() if trans else mat
        opa = view(A, m, k, lda, transa.upper() == "T")
        opb = view(B, k, n, ldb, transb.upper() == "T")
        cm = torch.as_strided(C, (m, n), (1, ldc))
        cm.copy_(alpha * (opa @ opb) + (beta * cm if beta != 0 else 0))
        return c

    @staticmethod
    def gemv(*args):
        """``gemv(alpha, matrix, vector, beta, r)`` → r = alpha · M v + beta · r."""
        alpha, mat, vec, beta, r = args
        M, v, out = _raw(mat), _raw(vec), _raw(r)
        out.copy_(alpha * (M @ v) + (beta * out if beta != 0 else 0))
        return r

    @staticmethod
    def ger(alpha, x, y, a):
        """A += alpha · x yᵀ"""
        _raw(a).add_(alpha * torch.outer(_raw(x), _raw(y)))
        return a


class DenseTensorConv:
    """2-D convolution / cross-correlation of matrices (DenseTensorConv.scala:24-262)."""

    @staticmethod
    def convSize(x: int, k: int, s: int, vf: str) -> int:
        if vf not in ("F", "V"):
            raise ValueError("type of convolution can be 'V' or 'F'")
        return (x - k) // s + 1 if vf == "V" else (x - 1) * s + k

    @staticmethod
    def conv2Dmul(alpha, t, k, srow: int = 1, scol: int = 1, vf: str = "V", xc: str = "C"):
        """alpha · (t ⊛ k): ``vf`` 'V'alid or 'F'ull, ``xc`` 'C'onvolution (kernel flipped) or 'X' cross-correlation."""
        import torch.nn.functional as F

        from .tensor import Tensor

        x, w = _raw(t), _raw(k)
        if x.dim() != 2 or w.dim() != 2:
            raise ValueError("input and kernel: 2D Tensor expected")
        if srow < 1 or scol < 1:
            raise ValueError("Stride should be a positive integer")
        if vf == "V" and (x.shape[0] < w.shape[0] or x.shape[1] < w.shape[1]):
            raise ValueError("conv2Dmul : Input image is smaller than kernel")
        kern = w.flip(0, 1) if xc == "C" else w
        xi, ki = x[None, None].to(torch.float64), kern[None, None].to(torch.float64)
        if vf == "V":
            out = F.conv2d(xi, ki, stride=(srow, scol))
        else:
            # full mode is the adjoint of the valid cross-correlation with the flipped kernel
            out = F.conv_transpose2d(xi, ki.flip(2, 3), stride=(srow, scol))
        return Tensor((alpha * out[0, 0]).to(x.dtype))


class DenseTensorApply:
    """Element-wise visitors (DenseTensorApply.scala): ``apply1(t, fn)`` maps in place, ``apply2(t1, t2, fn)`` writes
    ``fn(a, b)`` into ``t1``, ``apply3`` writes ``fn(b, c)`` into ``t1``."""

    @staticmethod
    def apply1(tensor, func):
        t = _raw(tensor)
        flat = t.reshape(-1).tolist()
        t.copy_(torch.tensor([func(v) for v in flat], dtype=t.dtype).reshape(t.shape))
        return tensor

    @staticmethod
    def apply2(tensor1, tensor2, func):
        a, b = _raw(tensor1), _raw(tensor2)
        if a.numel() != b.numel():
            raise ValueError("inconsistent tensor size")
        vals = [func(x, y) for x, y in zip(a.reshape(-1).tolist(), b.reshape(-1).tolist())]
        a.copy_(torch.tensor(vals, dtype=a.dtype).reshape(a.shape))
        return tensor1

    @staticmethod
    def apply3(tensor1, tensor2, tensor3, func):
        a, b, c = _raw(tensor1), _raw(tensor2), _raw(tensor3)
        if not a.numel() == b.numel() == c.numel():
            raise ValueError("inconsistent tensor size")
        vals = [func(y, z) for y, z in zip(b.reshape(-1).tolist(), c.reshape(-1).tolist())]
        a.copy_(torch.tensor(vals, dtype=a.dtype).reshape(a.shape))
        return tensor1


class DenseTensorDimApply:
    """Visit every 1-D slice along a (0-based) dimension (DenseTensorDimApply.scala): ``func`` gets the slices of each
    tensor as lists and returns the new values of the FIRST tensor's slice (or None to leave it)."""

    @staticmethod
    def dimApply2(tensor1, tensor2, _dim: int, func):
        a, b = _raw(tensor1), _raw(tensor2)
        am, bm = a.movedim(_dim, -1), b.movedim(_dim, -1)
        af, bf = am.reshape(-1, am.shape[-1]), bm.reshape(-1, bm.shape[-1])
        out = af.clone()
        for i in range(af.shape[0]):
            r = func(af[i].tolist(), bf[i].tolist())
            if r is not None:
                out[i] = torch.tensor(r, dtype=a.dtype)
        a.copy_(out.reshape(am.shape).movedim(-1, _dim))
        return tensor1

    @staticmethod
    def dimApply3(tensor1, tensor2, tensor3, dim: int, func):
        a, b, c = _raw(tensor1), _raw(tensor2), _raw(tensor3)
        am, bm, cm = a.movedim(dim, -1), b.movedim(dim, -1), c.movedim(dim, -1)
        af, bf, cf = (m.reshape(-1, m.shape[-1]) for m in (am, bm, cm))
        out = af.clone()
        for i in range(af.shape[0]):
            r = func(af[i].tolist(), bf[i].tolist(), cf[i].tolist())
            if r is not None:
                out[i] = torch.tensor(r, dtype=a.dtype)
        a.copy_(out.reshape(am.shape).movedim(-1, dim))
        return tensor1


class QuantizedTensorUnsupported:
    """Operations a quantised tensor refuses (QuantizedTensorUnsupported.scala): any attribute access explains why."""

    def __getattr__(self, name):
        raise NotImplementedError(f"QuantizedTensor doesn't support the operation `{name}`")
