"""csrc/blas_vml.cu (fp32/fp64 gemm, gemv, ger, dot, axpy, scal, vector math) against torch in fp64 on the GPU."""
import pytest
import torch

from bigdl_b200.tensor import mkl
from bigdl_b200.tensor.mkl import MKL
from bigdl_b200.tensor.tensor import Tensor

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_vml_and_level1(dtype):
    torch.manual_seed(0)
    n = 100003
    a = torch.rand(n, dtype=dtype, device=DEV) + 0.5
    b = torch.rand(n, dtype=dtype, device=DEV) + 0.5
    y = torch.empty_like(a)
    tol = 2e-6 if dtype == torch.float32 else 1e-13
    for name, ref in (("Add", a + b), ("Sub", a - b), ("Mul", a * b), ("Div", a / b)):
        getattr(MKL, "v" + name)(n, a, 0, b, 0, y, 0)
        assert torch.allclose(y, ref, rtol=tol, atol=tol), name
    for name, ref in (("Ln", a.log()), ("Exp", a.exp()), ("Sqrt", a.sqrt()), ("Tanh", a.tanh()), ("Log1p", a.log1p())):
        getattr(MKL, "v" + name)(n, a, 0, y, 0)
        assert torch.allclose(y, ref, rtol=10 * tol, atol=tol), name
    MKL.vPowx(n, a, 0, 1.7, y, 0)
    assert torch.allclose(y, a.pow(1.7), rtol=20 * tol, atol=tol)
    MKL.vAbs(n - 3, -a, 3, y, 1)               # unaligned offsets take the scalar path
    assert torch.equal(y[1:n - 2], a[3:])
    ref = b + 0.25 * a
    MKL.vaxpy(n, 0.25, a, 0, 1, b, 0, 1)
    assert torch.allclose(b, ref)
    d = MKL.vdot(n, a, 0, 1, b, 0, 1)
    assert abs(d - float(a.double() @ b.double())) <= 1e-5 * abs(d)
    assert MKL.vdot(n, a, 0, 1, b, 0, 1) == d      # fixed summation order: bit-identical on repeat
    MKL.vscal(n, 3.0, a, 0, 1)
    assert float(a.min()) >= 1.5


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("m,n,k", [(37, 29, 53), (512, 384, 300), (2048, 2304, 257)])
@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
def test_gemm(dtype, m, n, k, ta, tb):
    torch.manual_seed(m + n)
    A = torch.randn(k, m, dtype=dtype, device=DEV).t() if ta else torch.randn(m, k, dtype=dtype, device=DEV)
    B = torch.randn(n, k, dtype=dtype, device=DEV).t() if tb else torch.randn(k, n, dtype=dtype, device=DEV)
    C = torch.randn(m, n, dtype=dtype, device=DEV)
    ref = (1.5 * (A.double() @ B.double()) - 0.25 * C.double())
    mkl.mm(C, A, B, alpha=1.5, beta=-0.25)
    err = float((C.double() - ref).abs().max() / ref.abs().max())
    assert err < (2e-6 if dtype == torch.float32 else 1e-14), err


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_gemv_ger_and_tensor_routing(dtype):
    torch.manual_seed(3)
    M = torch.randn(300, 170, dtype=dtype, device=DEV)
    x, y = torch.randn(170, dtype=dtype, device=DEV), torch.randn(300, dtype=dtype, device=DEV)
    ref = 2.0 * (M.double() @ x.double()) + 0.5 * y.double()
    mkl.mv(y, M, x, alpha=2.0, beta=0.5)
    assert torch.allclose(y.double(), ref, atol=1e-4)
    yt = torch.empty(170, dtype=dtype, device=DEV)
    mkl.mv(yt, M.t(), torch.ones(300, dtype=dtype, device=DEV))
    assert torch.allclose(yt.double(), M.double().sum(0), atol=1e-4)
    A = torch.zeros(300, 170, dtype=dtype, device=DEV)
    mkl.ger(A, y, x, alpha=0.5)
    assert torch.allclose(A, 0.5 * torch.outer(y, x), atol=1e-5)
    c = Tensor().mm(Tensor(M), Tensor(M.t().contiguous()))
    assert torch.allclose(c.t.double(), M.double() @ M.double().t(), atol=1e-3)
