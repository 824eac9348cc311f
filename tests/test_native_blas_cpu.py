"""fp32/fp64 BLAS + VML entry points (N1): host library against torch, and Tensor math routed through them.
Reference behaviour: core/mkl2017-xeon-blas/src/main/c/mkl.c via S/tensor/TensorNumeric.scala:253-420."""
import pytest
import torch

from bigdl_b200.tensor import mkl
from bigdl_b200.tensor.mkl import MKL
from bigdl_b200.tensor.tensor import Tensor
from bigdl_b200.utils import config


@pytest.fixture(autouse=True)
def _native_everywhere():
    old = config.get("bigdl.tensor.nativeMath")
    config.set("bigdl.tensor.nativeMath", "all")
    yield
    config.set("bigdl.tensor.nativeMath", old)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_vml_matches_torch(dtype):
    torch.manual_seed(0)
    a = torch.rand(1003, dtype=dtype) + 0.5
    b = torch.rand(1003, dtype=dtype) + 0.5
    y = torch.empty_like(a)
    tol = 1e-6 if dtype == torch.float32 else 1e-13
    for name, ref in (("Add", a + b), ("Sub", a - b), ("Mul", a * b), ("Div", a / b)):
        getattr(MKL, "v" + name)(a.numel(), a, 0, b, 0, y, 0)
        assert torch.allclose(y, ref, rtol=tol, atol=tol), name
    for name, ref in (("Ln", a.log()), ("Exp", a.exp()), ("Sqrt", a.sqrt()), ("Tanh", a.tanh()), ("Abs", (-a).abs()),
                      ("Log1p", a.log1p())):
        src = -a if name == "Abs" else a
        getattr(MKL, "v" + name)(a.numel(), src, 0, y, 0)
        assert torch.allclose(y, ref, rtol=10 * tol, atol=tol), name
    MKL.vPowx(a.numel(), a, 0, 2.5, y, 0)
    assert torch.allclose(y, a.pow(2.5), rtol=10 * tol, atol=tol)
    # offsets address into the storage like the JNI's (storageOffset - 1)
    MKL.vsAdd(10, a, 5, b, 7, y, 100) if dtype == torch.float32 else MKL.vdAdd(10, a, 5, b, 7, y, 100)
    assert torch.allclose(y[100:110], a[5:15] + b[7:17])


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_level1_level2(dtype):
    torch.manual_seed(1)
    x, y = torch.randn(777, dtype=dtype), torch.randn(777, dtype=dtype)
    ref = y + 0.3 * x
    MKL.vaxpy(777, 0.3, x, 0, 1, y, 0, 1)
    assert torch.allclose(y, ref)
    assert abs(MKL.vdot(777, x, 0, 1, y, 0, 1) - float(x @ y)) < 1e-3
    assert abs(MKL.vdot(100, x, 3, 2, y, 1, 3) - float(x[3:203:2] @ y[1:301:3])) < 1e-4
    z = x.clone()
    MKL.vscal(777, -2.0, z, 0, 1)
    assert torch.allclose(z, -2 * x)
    A = torch.randn(5, 7, dtype=dtype)            # row-major [5,7] == column-major 7 x 5
    u, v = torch.randn(7, dtype=dtype), torch.randn(5, dtype=dtype)
    ref = A + 0.5 * torch.outer(v, u)
    MKL.vger(7, 5, 0.5, u, 0, 1, v, 0, 1, A, 0, 7)
    assert torch.allclose(A, ref, atol=1e-6)
    M = torch.randn(6, 4, dtype=dtype)
    xv, yv = torch.randn(4, dtype=dtype), torch.randn(6, dtype=dtype)
    ref = 2.0 * (M @ xv) + 0.5 * yv
    mkl.mv(yv, M, xv, alpha=2.0, beta=0.5)
    assert torch.allclose(yv, ref, atol=1e-5)
    yt = torch.randn(4, dtype=dtype)
    ref = M.t() @ torch.ones(6, dtype=dtype)
    mkl.mv(yt, M.t(), torch.ones(6, dtype=dtype))
    assert torch.allclose(yt, ref, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
def test_gemm_all_transposes(dtype, ta, tb):
    torch.manual_seed(2)
    m, n, k = 37, 29, 53
    A = torch.randn(k, m, dtype=dtype).t() if ta else torch.randn(m, k, dtype=dtype)
    B = torch.randn(n, k, dtype=dtype).t() if tb else torch.randn(k, n, dtype=dtype)
    C = torch.randn(m, n, dtype=dtype)
    ref = 1.5 * (A @ B) - 0.25 * C
    mkl.mm(C, A, B, alpha=1.5, beta=-0.25)
    assert torch.allclose(C, ref, atol=1e-4 if dtype == torch.float32 else 1e-11)


def test_tensor_math_uses_native_path(monkeypatch):
    calls = []
    real = mkl._call

    def spy(t, name, *a):
        calls.append(name)
        return real(t, name, *a)
    monkeypatch.setattr(mkl, "_call", spy)
    a, b = Tensor(torch.randn(8, 5)), Tensor(torch.randn(5, 3))
    c = Tensor().mm(a, b)
    assert torch.allclose(c.t, a.t @ b.t, atol=1e-5)
    v = Tensor(torch.randn(5))
    w = Tensor().mv(a, v)
    assert torch.allclose(w.t, a.t @ v.t, atol=1e-5)
    acc = Tensor(torch.zeros(8, 3))
    acc.addmm(2.0, a, b)
    assert torch.allclose(acc.t, 2 * (a.t @ b.t), atol=1e-5)
    r = Tensor(torch.zeros(8, 5))
    r.addr(1.0, Tensor(torch.ones(8)), v)
    assert torch.allclose(r.t, torch.outer(torch.ones(8), v.t), atol=1e-6)
    assert abs(v.dot(v) - float(v.t @ v.t)) < 1e-5
    e = Tensor(torch.rand(11) + 1).log()
    assert e.t.min() >= 0
    assert {"blas_gemm", "blas_gemv", "blas_ger", "blas_dot", "vml"} <= set(calls)
