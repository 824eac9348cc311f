"""fp32 / fp64 BLAS + vector math behind ``Tensor[Float]`` / ``Tensor[Double]`` — the native surface of N1.

The reference's ``TensorNumeric`` calls one JNI class for all dense math (``com.intel.analytics.bigdl.mkl.MKL``,
core/mkl2017-xeon-blas/src/main/c/mkl.c; call sites S/tensor/TensorNumeric.scala:253-420, S/tensor/DenseTensorBLAS.scala).
``MKL`` below keeps those static-method names and argument orders (storage, offset, leading dimension; column-major),
operating on torch tensors used as flat storages: CUDA storages go to ``csrc/blas_vml.cu`` (sm_100a SIMT kernels in the
tensor's own precision), host storages to ``csrc/host/blas_vml.cpp``. ``mm`` / ``mv`` / ``ger`` / ``dot`` / ``unary`` are
the row-major conveniences ``tensor.Tensor`` uses.
"""
from __future__ import annotations

import ctypes

import torch

from ..ops import _lib
from ..utils import config

_OPS = {"Add": 0, "Sub": 1, "Mul": 2, "Div": 3, "Powx": 4, "Ln": 5, "Exp": 6, "Sqrt": 7, "Tanh": 8, "Abs": 9, "Log1p": 10}
_P, _I, _L = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
_declared = {"cuda": False, "host": False}


def _suffix(t: torch.Tensor) -> str:
    if t.dtype == torch.float32:
        return "f32"
    if t.dtype == torch.float64:
        return "f64"
    raise TypeError(f"native BLAS/VML handles float32 and float64 storages, not {t.dtype}")


def _real(suf):
    return ctypes.c_float if suf == "f32" else ctypes.c_double


def _declare(lib, prefix: str, with_stream: bool):
    tail = [_P] if with_stream else []
    for suf in ("f32", "f64"):
        R = _real(suf)
        sigs = {
            f"{prefix}vml_{suf}": ([_I, _L, _P, _P, R, _P], _I),
            f"{prefix}blas_axpy_{suf}": ([_L, R, _P, _L, _P, _L], None),
            f"{prefix}blas_scal_{suf}": ([_L, R, _P, _L], None),
            f"{prefix}blas_ger_{suf}": ([_I, _I, R, _P, _L, _P, _L, _P, _L], None),
            f"{prefix}blas_gemv_{suf}": ([_I, _I, _I, R, _P, _L, _P, _L, R, _P, _L], None),
            f"{prefix}blas_gemm_{suf}": ([_I, _I, _I, _I, _I, R, _P, _L, _P, _L, R, _P, _L], None),
        }
        for name, (args, res) in sigs.items():
            fn = getattr(lib, name)
            fn.argtypes = args + tail
            fn.restype = _I if (with_stream or res is _I) else None
        dot = getattr(lib, f"{prefix}blas_dot_{suf}")
        if with_stream:
            dot.argtypes, dot.restype = [_L, _P, _L, _P, _L, _P, _P], _I
        else:
            dot.argtypes, dot.restype = [_L, _P, _L, _P, _L], R


def _cuda():
    lib = _lib.cuda_lib()
    if lib is None:
        raise _lib.NativeError("the CUDA kernel library is not available in this process")
    if not _declared["cuda"]:
        _declare(lib, "bdl_", True)
        _declared["cuda"] = True
    return lib


def _host():
    lib = _lib.host_lib()
    if not _declared["host"]:
        _declare(lib, "bdlh_", False)
        _declared["host"] = True
    return lib


def _at(t: torch.Tensor, off: int) -> int:
    return t.data_ptr() + off * t.element_size()


def _call(t: torch.Tensor, name: str, *args):
    """Dispatch ``name`` (without prefix / suffix) on the library that owns ``t``'s memory."""
    suf = _suffix(t)
    if t.is_cuda:
        with torch.cuda.device(t.device):
            _lib.check(getattr(_cuda(), f"bdl_{name}_{suf}")(*args, _lib.stream_ptr()), f"bdl_{name}_{suf}")
            _lib.count_launch()
        return None
    return getattr(_host(), f"bdlh_{name}_{suf}")(*args)


def enabled_for(t: torch.Tensor) -> bool:
    """Does Tensor math on ``t`` go through the native entry points? ``bigdl.tensor.nativeMath``: cuda | all | off."""
    if t.dtype not in (torch.float32, torch.float64):
        return False
    mode = str(config.get("bigdl.tensor.nativeMath")).lower()
    if mode == "off":
        return False
    if t.is_cuda:
        return _lib.cuda_lib() is not None
    return mode == "all"


class MKL:
    """Static methods with the reference's JNI names. ``a``/``b``/``y`` are flat torch tensors (storages); offsets are
    0-based element offsets like the JNI's (the Scala callers pass ``storageOffset() - 1``)."""

    @staticmethod
    def isMKLLoaded() -> bool:
        return True

    @staticmethod
    def getNumThreads() -> int:
        return torch.get_num_threads()

    @staticmethod
    def setNumThreads(n: int) -> None:
        torch.set_num_threads(int(n))

    @staticmethod
    def _v2(op, n, a, aOffset, b, bOffset, y, yOffset):
        rc = _call(y, "vml", _OPS[op], int(n), _at(a, aOffset), _at(b, bOffset), 0.0, _at(y, yOffset))
        if rc:
            raise _lib.NativeError(f"v?{op} failed")

    @staticmethod
    def _v1(op, n, a, aOffset, y, yOffset, scalar=0.0):
        rc = _call(y, "vml", _OPS[op], int(n), _at(a, aOffset), None, float(scalar), _at(y, yOffset))
        if rc:
            raise _lib.NativeError(f"v?{op} failed")

    @staticmethod
    def vAdd(n, a, aOffset, b, bOffset, y, yOffset): MKL._v2("Add", n, a, aOffset, b, bOffset, y, yOffset)
    @staticmethod
    def vSub(n, a, aOffset, b, bOffset, y, yOffset): MKL._v2("Sub", n, a, aOffset, b, bOffset, y, yOffset)
    @staticmethod
    def vMul(n, a, aOffset, b, bOffset, y, yOffset): MKL._v2("Mul", n, a, aOffset, b, bOffset, y, yOffset)
    @staticmethod
    def vDiv(n, a, aOffset, b, bOffset, y, yOffset): MKL._v2("Div", n, a, aOffset, b, bOffset, y, yOffset)
    @staticmethod
    def vPowx(n, a, aOffset, b, y, yOffset): MKL._v1("Powx", n, a, aOffset, y, yOffset, b)
    @staticmethod
    def vLn(n, a, aOffset, y, yOffset): MKL._v1("Ln", n, a, aOffset, y, yOffset)
    @staticmethod
    def vExp(n, a, aOffset, y, yOffset): MKL._v1("Exp", n, a, aOffset, y, yOffset)
    @staticmethod
    def vSqrt(n, a, aOffset, y, yOffset): MKL._v1("Sqrt", n, a, aOffset, y, yOffset)
    @staticmethod
    def vTanh(n, a, aOffset, y, yOffset): MKL._v1("Tanh", n, a, aOffset, y, yOffset)
    @staticmethod
    def vAbs(n, a, aOffset, y, yOffset): MKL._v1("Abs", n, a, aOffset, y, yOffset)
    @staticmethod
    def vLog1p(n, a, aOffset, y, yOffset): MKL._v1("Log1p", n, a, aOffset, y, yOffset)

    @staticmethod
    def vaxpy(n, da, dx, dxOffset, incx, dy, dyOffset, incy):
        _call(dy, "blas_axpy", int(n), float(da), _at(dx, dxOffset), int(incx), _at(dy, dyOffset), int(incy))

    @staticmethod
    def vscal(n, sa, sx, offset, incx):
        _call(sx, "blas_scal", int(n), float(sa), _at(sx, offset), int(incx))

    @staticmethod
    def vdot(n, dx, dxOffset, incx, dy, dyOffset, incy) -> float:
        if dx.is_cuda:
            out = torch.empty(1, dtype=dx.dtype, device=dx.device)
            _call(dx, "blas_dot", int(n), _at(dx, dxOffset), int(incx), _at(dy, dyOffset), int(incy), out.data_ptr())
            return float(out)
        return float(_call(dx, "blas_dot", int(n), _at(dx, dxOffset), int(incx), _at(dy, dyOffset), int(incy)))

    @staticmethod
    def vger(m, n, alpha, x, xOffset, incx, y, yOffset, incy, a, aOffset, lda):
        _call(a, "blas_ger", int(m), int(n), float(alpha), _at(x, xOffset), int(incx), _at(y, yOffset), int(incy),
              _at(a, aOffset), int(lda))

    @staticmethod
    def vgemv(trans, m, n, alpha, a, aOffset, lda, x, xOffset, incx, beta, y, yOffset, incy):
        _call(y, "blas_gemv", int(str(trans).upper().startswith("T")), int(m), int(n), float(alpha), _at(a, aOffset),
              int(lda), _at(x, xOffset), int(incx), float(beta), _at(y, yOffset), int(incy))

    @staticmethod
    def vgemm(transa, transb, m, n, k, alpha, a, aOffset, lda, b, bOffset, ldb, beta, c, cOffset, ldc):
        _call(c, "blas_gemm", int(str(transa).upper().startswith("T")), int(str(transb).upper().startswith("T")), int(m),
              int(n), int(k), float(alpha), _at(a, aOffset), int(lda), _at(b, bOffset), int(ldb), float(beta),
              _at(c, cOffset), int(ldc))


# the reference has one entry per precision (vsAdd / vdAdd ...); the storage dtype selects the kernel here
for _n in ("Add", "Sub", "Mul", "Div", "Powx", "Ln", "Exp", "Sqrt", "Tanh", "Abs", "Log1p", "axpy", "scal", "dot", "ger",
           "gemv", "gemm"):
    for _p in ("s", "d"):
        setattr(MKL, f"v{_p}{_n}", getattr(MKL, f"v{_n}"))


# ------------------------------------------------------------------------------------------ row-major conveniences
def _as_colmajor(t: torch.Tensor):
    """(tensor to take the pointer of, transposed?, ld) describing the 2-D row-major ``t`` as a column-major operand:
    transposed=False means the memory IS column-major t^T with leading dimension ld."""
    if t.stride(1) == 1 and t.stride(0) >= max(1, t.shape[1]):
        return t, False, t.stride(0)
    if t.stride(0) == 1 and t.stride(1) >= max(1, t.shape[0]):
        return t, True, t.stride(1)
    t = t.contiguous()
    return t, False, t.stride(0)


def mm(out: torch.Tensor, m1: torch.Tensor, m2: torch.Tensor, alpha: float = 1.0, beta: float = 0.0) -> torch.Tensor:
    """out[m,n] = alpha * m1[m,k] @ m2[k,n] + beta * out (row-major tensors; out has unit inner stride)."""
    m, k = m1.shape
    k2, n = m2.shape
    if k != k2 or tuple(out.shape) != (m, n) or out.stride(1) != 1:
        raise ValueError("mm: shape mismatch")
    a, ta, lda = _as_colmajor(m1)
    b, tb, ldb = _as_colmajor(m2)
    # row-major out = column-major out^T (n x m) = op(m2^T) op(m1^T)
    MKL.vgemm("T" if tb else "N", "T" if ta else "N", n, m, k, alpha, b, 0, ldb, a, 0, lda, beta, out, 0, out.stride(0))
    return out


def mv(out: torch.Tensor, mat: torch.Tensor, vec: torch.Tensor, alpha: float = 1.0, beta: float = 0.0) -> torch.Tensor:
    m, n = mat.shape
    a, t, lda = _as_colmajor(mat)
    # memory is column-major mat^T (n x m) when not transposed: y = (mat^T)^T x
    if t:
        MKL.vgemv("N", m, n, alpha, a, 0, lda, vec, 0, vec.stride(0), beta, out, 0, out.stride(0))
    else:
        MKL.vgemv("T", n, m, alpha, a, 0, lda, vec, 0, vec.stride(0), beta, out, 0, out.stride(0))
    return out


def ger(out: torch.Tensor, x: torch.Tensor, y: torch.Tensor, alpha: float = 1.0) -> torch.Tensor:
    """out[m,n] += alpha * x y^T (row-major out with unit inner stride = column-major out^T += alpha * y x^T)."""
    m, n = out.shape
    MKL.vger(n, m, alpha, y, 0, y.stride(0), x, 0, x.stride(0), out, 0, out.stride(0))
    return out


def dot(x: torch.Tensor, y: torch.Tensor) -> float:
    x, y = x.reshape(-1), y.reshape(-1)
    return MKL.vdot(x.numel(), x, 0, x.stride(0) if x.numel() > 1 else 1, y, 0, y.stride(0) if y.numel() > 1 else 1)


def unary(op: str, out: torch.Tensor, src: torch.Tensor, scalar: float = 0.0) -> torch.Tensor:
    """out = op(src) for contiguous tensors of one dtype (op: Ln, Exp, Sqrt, Tanh, Abs, Log1p, Powx)."""
    MKL._v1(op, src.numel(), src, 0, out, 0, scalar)
    return out


def binary(op: str, out: torch.Tensor, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    MKL._v2(op, a.numel(), a, 0, b, 0, out, 0)
    return out
