"""Torch7-style Tensor API (mirrors T/tensor/DenseTensorSpec.scala, DenseTensorMathSpec.scala, SparseTensorSpec.scala,
QuantizedTensorSpec.scala)."""
import numpy as np
import pytest
import torch

from bigdl_b200.tensor import QuantizedTensor, SparseTensor, Storage, Tensor, cat


def test_one_based_views_share_storage():
    t = Tensor(3, 4).fill(0)
    assert t.nDimension() == 2 and t.size() == [3, 4] and t.size(2) == 4 and t.stride() == [4, 1]
    row = t.select(1, 2)
    row.fill(7)
    assert t.valueAt(2, 1) == 7 and t.valueAt(1, 1) == 0
    sub = t.narrow(2, 2, 2)
    assert sub.size() == [3, 2] and sub.storageOffset() == 2
    sub.setValue(1, 1, 5.0)
    assert t.valueAt(1, 2) == 5.0
    assert t(2).size() == [4] and t(2, 3) == 7
    s = Storage(torch.arange(12.0))
    v = Tensor(s, 3, [2, 3])
    assert v.valueAt(1, 1) == 2.0 and v.valueAt(2, 3) == 7.0


def test_math_overloads_are_in_place_and_return_self():
    a = Tensor([[1.0, 2.0], [3.0, 4.0]])
    b = Tensor([[1.0, 1.0], [1.0, 1.0]])
    assert a.add(2.0, b) is a and a.t.tolist() == [[3.0, 4.0], [5.0, 6.0]]
    a.add(b, -1.0, b)
    assert a.t.tolist() == [[0.0, 0.0], [0.0, 0.0]]
    c = Tensor(2, 2).fill(1)
    c.addmm(0.5, Tensor([[1.0, 2.0], [3.0, 4.0]]), Tensor([[1.0, 0.0], [0.0, 1.0]]))
    assert c.t.tolist() == [[1.5, 2.0], [2.5, 3.0]]
    d = Tensor(2).fill(0)
    d.addmv(1.0, Tensor([[1.0, 2.0], [3.0, 4.0]]), Tensor([1.0, 1.0]))
    assert d.t.tolist() == [3.0, 7.0]
    e = Tensor(2, 2).fill(0).addr(1.0, Tensor([1.0, 2.0]), Tensor([3.0, 4.0]))
    assert e.t.tolist() == [[3.0, 4.0], [6.0, 8.0]]
    f = Tensor([1.0, 4.0, 9.0]).sqrt()
    assert f.t.tolist() == [1.0, 2.0, 3.0]
    g = Tensor(3).fill(1).addcmul(2.0, Tensor([1.0, 2.0, 3.0]), Tensor([1.0, 1.0, 2.0]))
    assert g.t.tolist() == [3.0, 5.0, 13.0]
    assert (Tensor([[1.0, 2.0]]) * Tensor([[1.0], [1.0]])).t.tolist() == [[3.0]]      # matrix product like Scala `*`
    assert Tensor([3.0, 4.0]).norm(2) == 5.0 and Tensor([3.0, 4.0]).sumSquare() == 25.0
    assert Tensor([1.0, 2.0]).dot(Tensor([3.0, 4.0])) == 11.0


def test_reductions_use_one_based_indices():
    t = Tensor([[1.0, 9.0, 3.0], [7.0, 2.0, 8.0]])
    v, i = t.max(2)
    assert v.t.tolist() == [[9.0], [8.0]] and i.t.tolist() == [[2.0], [3.0]]
    v, i = t.min(1)
    assert i.t.tolist() == [[1.0, 2.0, 1.0]]
    v, i = t.topk(2, 2, increase=False)
    assert v.t.tolist() == [[9.0, 3.0], [8.0, 7.0]] and i.t.tolist() == [[2.0, 3.0], [3.0, 1.0]]
    assert t.sum() == 30.0 and t.sum(1).t.tolist() == [[8.0, 11.0, 11.0]] and abs(t.mean() - 5.0) < 1e-6
    idx = Tensor([[1.0, 1.0, 1.0]])
    assert Tensor(1, 3).gather(1, idx, t).t.tolist() == [[1.0, 9.0, 3.0]]
    assert t.index(2, Tensor([3.0, 1.0])).t.tolist() == [[3.0, 1.0], [8.0, 7.0]]


def test_resize_copy_closures_and_masks():
    t = Tensor(2, 3).fill(1)
    t.resize(3, 2)
    assert t.size() == [3, 2] and t.sum() == 6.0
    t.resize(2, 2)
    assert t.size() == [2, 2]
    t.copy(Tensor([1.0, 2.0, 3.0, 4.0]))
    assert t.t.tolist() == [[1.0, 2.0], [3.0, 4.0]]
    t.apply1(lambda v: v * 2)
    assert t.t.tolist() == [[2.0, 4.0], [6.0, 8.0]]
    t.map(Tensor(2, 2).fill(1), lambda a, b: a - b)
    assert t.t.tolist() == [[1.0, 3.0], [5.0, 7.0]]
    m = t.gt(4.0)
    assert m.t.tolist() == [[0.0, 0.0], [1.0, 1.0]]
    assert t.maskedSelect(m).t.tolist() == [5.0, 7.0]
    t.maskedFill(m, 0.0)
    assert t.sum() == 4.0
    assert Tensor.range(1, 5, 2).t.tolist() == [1.0, 3.0, 5.0]
    assert cat([Tensor(1, 2).fill(1), Tensor(1, 2).fill(2)], 1).size() == [2, 2]
    assert Tensor([1.0]).almostEqual(Tensor([1.0005]), 1e-3) and Tensor([1.0]) == Tensor([1.0])
    k = Tensor([[1.0, 0.0], [0.0, -1.0]])
    x = Tensor(torch.arange(9.0).reshape(3, 3))
    assert x.xcorr2(k).t.tolist() == [[-4.0, -4.0], [-4.0, -4.0]]
    assert x.conv2(k).t.tolist() == [[4.0, 4.0], [4.0, 4.0]]


def test_sparse_tensor():
    sp = Tensor.sparse([[1, 1, 3], [1, 3, 2]], [1.0, 2.0, 3.0], (3, 3))          # 0-based like torch? no: explicit below
    sp = SparseTensor([[1, 1, 3], [1, 3, 2]], [1.0, 2.0, 3.0], (3, 3), one_based=True)
    d = sp.to_dense()
    assert d.t.tolist() == [[1.0, 0.0, 2.0], [0.0, 0.0, 0.0], [0.0, 3.0, 0.0]]
    assert sp.numNonZeroByRow() == [2, 0, 1] and sp.nElement() == 3
    w = torch.arange(6.0).reshape(3, 2)
    assert torch.equal(sp.mm(w), d.t @ w)
    assert sp.narrow(1, 3, 1).to_dense().t.tolist() == [[0.0, 3.0, 0.0]]
    j = SparseTensor.concat(2, [sp, sp])
    assert j.size() == [3, 6] and j.nElement() == 6
    assert SparseTensor.from_dense(d).to_dense() == d


def test_quantized_tensor_matches_reference_rule():
    w = torch.tensor([[0.5, -1.0, 0.25], [2.0, 1.0, -0.5]])
    q = Tensor(w).toQuantizedTensor()
    assert isinstance(q, QuantizedTensor)
    assert q.q.tolist() == [[64, -127, 32], [127, 64, -32]]       # round(x/absmax*127)
    assert q.maxOfRow.tolist() == [0.5, 2.0] and q.minOfRow.tolist() == [-1.0, -0.5]
    assert (q.dequantize() - w).abs().max() < 0.01
    b, am = q.to_e4m3()
    assert b.dtype == torch.uint8 and b.shape == (2, 16) and torch.allclose(am, torch.tensor([1.0, 2.0]), atol=0.01)


def test_remaining_tensor_math_and_shape_methods():
    """The TensorMath / Tensor methods no other test calls (S/tensor/TensorMath.scala, Tensor.scala), 1-based dims."""
    import torch
    from bigdl_b200.tensor import Tensor

    a = Tensor(torch.tensor([[1., 2., 3.], [4., 5., 6.]]))
    b = Tensor(torch.tensor([[2., 2., 2.], [0.5, 0.5, 0.5]]))
    assert torch.equal(Tensor(2, 3).cmul(a, b).t, a.t * b.t) and torch.equal(Tensor(2, 3).cdiv(a, b).t, a.t / b.t)
    assert torch.equal(Tensor(2, 3).mul(a, 3.0).t, a.t * 3) and torch.equal(Tensor(2, 3).div(a, b).t, a.t / b.t)
    assert torch.equal(a.clone().cmul(b).t, a.t * b.t) and torch.equal(a.clone().mul(2.0).t, a.t * 2)
    assert torch.equal(Tensor(2, 3).cmax(a, b).t, torch.maximum(a.t, b.t))
    assert torch.equal(Tensor(2, 3).cmin(a, b).t, torch.minimum(a.t, b.t))
    assert torch.equal(a.clone().cmax(3.5).t, a.t.clamp(min=3.5))
    assert torch.allclose(a.clone().addcdiv(2.0, a, b).t, a.t + 2.0 * a.t / b.t)
    m, v = Tensor(torch.randn(3, 4)), Tensor(torch.randn(4))
    assert torch.allclose(Tensor(3).mv(m, v).t, m.t @ v.t)
    b1, b2 = Tensor(torch.randn(2, 3, 4)), Tensor(torch.randn(2, 4, 5))
    assert torch.allclose(Tensor(2, 3, 5).bmm(b1, b2).t, torch.bmm(b1.t, b2.t))
    M = Tensor(torch.randn(2, 3, 5))
    assert torch.allclose(Tensor(2, 3, 5).baddbmm(0.5, M, 2.0, b1, b2).t, 0.5 * M.t + 2.0 * torch.bmm(b1.t, b2.t))
    assert torch.equal(a.eq(a).t, torch.ones(2, 3)) and float(a.ge(4.0).t.sum()) == 3 and float(a.le(2.0).t.sum()) == 2
    assert abs(a.prod() - 720.0) < 1e-6 if not isinstance(a.prod(), Tensor) else abs(float(a.prod().t) - 720.0) < 1e-6
    assert torch.allclose(Tensor(2, 3).log1p(a).t, torch.log1p(a.t)) and torch.allclose(Tensor(2, 3).inv(a).t, 1 / a.t)
    vals, idx = a.sort(2, True)
    assert torch.equal(vals.t[0], torch.tensor([3., 2., 1.])) and int(idx.t[0, 0]) == 3          # 1-based indices
    assert a.clone().addSingletonDimension(1).size() == [1, 2, 3]      # in place on the receiver, like the reference
    assert Tensor(1, 3, 1).squeeze().size() == [3] and Tensor(1, 3, 1).squeeze(1).size() == [3, 1]
    assert Tensor(1, 3).squeezeNewTensor().size() == [3]
    c = Tensor(torch.arange(6.).reshape(2, 3))
    assert c.t_().size() == [3, 2]
    assert Tensor(torch.ones(1, 3)).expandAs(Tensor(4, 3)).size() == [4, 3]
    assert Tensor(torch.ones(2)).repeatTensor(3, 2).size() == [3, 4]
    assert Tensor(5).resizeAs(Tensor(2, 3)).size() == [2, 3]
    assert Tensor(torch.arange(6.)).unfold(1, 2, 2).size() == [3, 2]
    assert Tensor.scalar(3.5).isScalar() and abs(float(Tensor.scalar(3.5).t) - 3.5) < 1e-6
    assert Tensor().isEmpty() and not a.isEmpty() and a.isContiguous() and a.isSameSizeAs(b)
    assert a.getType() == torch.float32 and a.toTensor() is a
    z = Tensor(2, 3).zipWith(a, b, lambda p, q: p + 2 * q)
    assert torch.allclose(z.t, a.t + 2 * b.t)
    acc = Tensor(torch.zeros(3, 2))
    acc.indexAdd(1, torch.tensor([1., 3., 1.]), torch.ones(3, 2))
    assert torch.equal(acc.t, torch.tensor([[2., 2.], [0., 0.], [1., 1.]]))
    dst = Tensor(torch.zeros(2, 3))
    dst.maskedCopy(torch.tensor([[1, 0, 1], [0, 1, 0]]), torch.tensor([7., 8., 9.]))
    assert torch.equal(dst.t, torch.tensor([[7., 0., 8.], [0., 9., 0.]]))
    bern = Tensor(1000).bernoulli(0.3)
    assert 0.2 < float(bern.t.mean()) < 0.4 and set(bern.t.unique().tolist()) <= {0.0, 1.0}


def test_remaining_tensor_surface():
    """apply / update with index tables, applyFun, cast, reduce, unique, erf..., conversions, gaussian1D
    (S/tensor/Tensor.scala:188-319, TensorMath.scala; reference checks in DenseTensorSpec / DenseTensorMathSpec)."""
    import math
    from bigdl_b200.utils.table import T as Tb
    t = Tensor(torch.arange(1., 25.).reshape(2, 3, 4))
    assert t.apply([2, 3, 4]) == 24.0
    assert t.apply(2).size() == [3, 4] and t.apply(-1).valueAt(1, 1) == 13.0
    sub = t.apply(Tb(2, Tb(2, 3)))                       # select 2 in dim 1, narrow rows 2..3
    assert sub.size() == [2, 4] and sub.valueAt(1, 1) == 17.0
    assert t.apply(Tb(Tb(), Tb(), 4)).size() == [2, 3]   # empty entries keep the dimension
    t.update(1, 0.0)
    assert float(t.t[0].abs().sum()) == 0.0
    t.update([2, 1, 1], -5.0)
    assert t.valueAt(2, 1, 1) == -5.0
    t.update(Tb(2, Tb(2, 3)), Tensor(torch.ones(2, 4)))
    assert float(t.t[1, 1:].sum()) == 8.0
    t.update(lambda v: v < 0, 7.0)
    assert t.valueAt(2, 1, 1) == 7.0
    a = Tensor(torch.zeros(3))
    a.applyFun(Tensor(torch.tensor([1, 2, 3], dtype=torch.int32)), lambda v: v * 0.5)
    assert a.toArray() == [0.5, 1.0, 1.5]
    d = Tensor(torch.zeros(3, dtype=torch.float64))
    assert a.cast(d).t.dtype == torch.float64 and d.toArray() == [0.5, 1.0, 1.5]
    m = Tensor(torch.tensor([[1., 5.], [4., 2.]]))
    r = Tensor()
    m.reduce(1, r, lambda x, y: max(x, y))
    assert r.t.tolist() == [[4.0, 5.0]]
    vals, idx = Tensor(torch.tensor([3., 1., 3., 2., 1.])).unique()
    assert vals.toArray() == [3.0, 1.0, 2.0] and idx.t.tolist() == [0, 1, 0, 2, 1]
    e = Tensor(torch.tensor([0.5])).erf()
    assert abs(e.valueAt(1) - math.erf(0.5)) < 1e-6
    assert abs(Tensor(torch.tensor([0.5])).erfc().valueAt(1) - math.erfc(0.5)) < 1e-6
    assert abs(Tensor(torch.tensor([4.0])).logGamma().valueAt(1) - math.lgamma(4.0)) < 1e-5
    assert abs(Tensor(torch.tensor([1.0])).digamma().valueAt(1) + 0.5772157) < 1e-5
    assert m.notEqualValue(1.0) and not Tensor(torch.ones(2)).notEqualValue(1.0)
    sc = m.shallowClone()
    sc.setValue(1, 1, 9.0)
    assert m.valueAt(1, 1) == 9.0                        # same storage
    assert m.toBreezeMatrix().shape == (2, 2) and Tensor(torch.ones(3)).toMLlibVector().tolist() == [1, 1, 1]
    assert m.emptyInstance().nElement() == 0 and m.getTensorType() == "DenseType"
    assert m.diff(m.clone()) is False and m.diff(Tensor(torch.zeros(2, 2)), 1) is True
    bias_grad = Tensor(torch.zeros(1, 3, 1))
    bias_grad.addMultiDimension(Tensor(torch.ones(2, 3, 4)), [1, 3])
    assert bias_grad.t.reshape(-1).tolist() == [8.0, 8.0, 8.0]
    g = Tensor.gaussian1D(5, 0.25, 1, False, 0.5)
    want = [math.exp(-(((i - 3.0) / (0.25 * 5)) ** 2) / 2) for i in range(1, 6)]
    assert all(abs(x - y) < 1e-6 for x, y in zip(g.toArray(), want))
    assert abs(Tensor.gaussian1D(7, normalize=True).sum() - 1.0) < 1e-6
    assert 0.0 <= Tensor().uniform() < 1.0 and 2.0 <= Tensor().uniform(2.0, 3.0) < 3.0
    assert Tensor.getDims(m) == 2


def test_tensor_save_load(tmp_path):
    t = Tensor(torch.arange(6.).reshape(2, 3))
    p = str(tmp_path / "t.obj")
    t.save(p)
    assert torch.equal(Tensor.load(p).t, t.t)
    with pytest.raises(FileExistsError):
        t.save(p)
    t.save(p, True)


def test_sparse_tensor_blas_and_math():
    """S/tensor/SparseTensorBLAS.scala coomv / coomm / vdot and SparseTensorMath addmv / addmm against dense math."""
    from bigdl_b200.tensor.sparse import SparseTensor, SparseTensorBLAS, SparseTensorMath
    torch.manual_seed(0)
    d = torch.randn(5, 7) * (torch.rand(5, 7) < 0.4)
    sp = SparseTensor.from_dense(d)
    assert sp.getTensorType() == "SparseType" and sp.nElement() == int((d != 0).sum())
    nz = d.nonzero()[0]
    assert abs(sp.apply([int(nz[0]) + 1, int(nz[1]) + 1]) - float(d[nz[0], nz[1]])) < 1e-7
    x, y = torch.randn(7), torch.randn(5)
    want = 0.5 * (d @ x) + 2.0 * y
    SparseTensorBLAS.coomv(0.5, sp, x, 2.0, y)
    assert torch.allclose(y, want, atol=1e-6)
    B, C = torch.randn(7, 3), torch.randn(5, 3)
    want = 1.5 * (d @ B) - C
    SparseTensorBLAS.coomm(1.5, sp, B, -1.0, C)
    assert torch.allclose(C, want, atol=1e-5)
    A2, C2 = torch.randn(4, 5), torch.zeros(4, 7)
    SparseTensorBLAS.coomm(1.0, A2, sp, 0.0, C2)                 # dense x sparse
    assert torch.allclose(C2, A2 @ d, atol=1e-5)
    v = torch.randn(7) * (torch.rand(7) < 0.5)
    sv = SparseTensor.from_dense(v)
    w = torch.randn(7)
    assert abs(SparseTensorBLAS.vdot(w, sv) - float(w @ v)) < 1e-5 and abs(sv.dot(w) - float(w @ v)) < 1e-5
    r, t = torch.zeros(5), torch.randn(5)
    SparseTensorMath.addmv(r, 0.5, t, 2.0, sp, x)
    assert torch.allclose(r, 0.5 * t + 2.0 * (d @ x), atol=1e-5)
    r2, t2 = torch.zeros(5, 3), torch.randn(5, 3)
    SparseTensorMath.addmm(r2, 1.0, t2, 1.0, sp, B)
    assert torch.allclose(r2, t2 + d @ B, atol=1e-5)
    assert abs(sp.sum() - float(d.sum())) < 1e-5
    small = sp.clone().resize([3, 7])
    assert torch.equal(small.to_dense().t, d[:3])


def test_dnn_storage_native_pool():
    """DnnStorage / DnnTensor on the host: aligned native blocks, explicit release, leak accounting, reuse
    (S/tensor/DnnStorage.scala:99-125, DnnTensorSpec)."""
    from bigdl_b200.tensor.dnn import DnnStorage, DnnTensor
    base = DnnStorage.stats()
    a, b = DnnTensor(4, 5, device="cpu"), DnnTensor([4, 5], device="cpu")
    assert a.storageAddress() % 64 == 0 and a.size() == [4, 5] and a.nElement() == 20
    a.copy(torch.ones(4, 5))
    b.copy(a).axpby(2.0, 0.5, a)                         # b = 2a + 0.5b
    assert float(b.dense().t.sum()) == 50.0
    assert DnnStorage.stats()["live_blocks"] == base["live_blocks"] + 2
    a.release()
    a.release()                                          # idempotent
    with pytest.raises(RuntimeError):
        a.zero()
    b.release()
    mid = DnnStorage.stats()
    assert mid["live_blocks"] == base["live_blocks"]     # nothing leaked
    c = DnnTensor(20, device="cpu")                      # same size class: served from the cache
    assert DnnStorage.stats()["reused"] == mid["reused"] + 1
    c.release()
    s = DnnStorage(8)
    s.copy(torch.arange(8.))
    assert s.tensor().tolist() == list(range(8))
    with pytest.raises(NotImplementedError):
        s.fill(1.0, 0, 8)
    s.release()
    with pytest.raises(RuntimeError):
        s.tensor()


def test_type_tags_converters_and_dense_helpers():
    import torch.nn.functional as F

    from bigdl_b200 import tensor as TT
    from bigdl_b200.tensor import types as ty

    assert TT.FloatType == torch.float32 and TT.Tensor(torch.zeros(2)).getType() == TT.FloatType
    assert ty.TensorDataType.of(torch.int64) is TT.LongType and repr(TT.DoubleType) == "DoubleType"
    assert TT.Tensor(torch.zeros(2)).getTensorType() == TT.DenseType and TT.DenseType != TT.SparseType
    assert ty.ConvertableToInt.fromFloat(3.9) == 3 and ty.ConvertableToInt.fromDouble(-3.9) == -3
    assert ty.ConvertableToShort.fromInt(40000) == 40000 - 65536 and ty.ConvertableToInt.fromDouble(1e12) == 2 ** 31 - 1
    assert ty.ConvertableToFloat.fromDouble(0.1) == float(torch.tensor(0.1)) and ty.ConvertableToDouble.fromString("2.5") == 2.5
    assert ty.ConvertableFromDouble.toInt(7.7) == 7 and ty.ConvertableFromString.toBoolean("True") and ty.ConvertableFromInt.toChar(65) == "A"
    assert ty.TensorFunc2(lambda d, o: d[o] + 1).apply([1, 2], 1) == 3

    a, b = torch.randn(3, 4, dtype=torch.float64), torch.randn(4, 5, dtype=torch.float64)
    c = torch.zeros(15, dtype=torch.float64)
    ty.DenseTensorBLAS.gemm("N", "N", 3, 5, 4, 1.0, a.t().contiguous().reshape(-1), 3, b.t().contiguous().reshape(-1), 4, 0.0, c, 3)
    assert torch.allclose(c.reshape(5, 3).t(), a @ b)
    c2 = torch.ones(15, dtype=torch.float64)
    ty.DenseTensorBLAS.gemm("T", "N", 3, 5, 4, 2.0, a.contiguous().reshape(-1), 4, b.t().contiguous().reshape(-1), 4, 1.0, c2, 3)
    assert torch.allclose(c2.reshape(5, 3).t(), 2 * (a @ b) + 1)
    r = torch.ones(3, dtype=torch.float64)
    ty.DenseTensorBLAS.gemv(2.0, a, b[:, 0].clone(), 1.0, r)
    assert torch.allclose(r, 2 * (a @ b[:, 0]) + 1)
    m = torch.zeros(3, 5, dtype=torch.float64)
    ty.DenseTensorBLAS.ger(0.5, a[:, 0].clone(), b[0].clone(), m)
    assert torch.allclose(m, 0.5 * torch.outer(a[:, 0], b[0]))

    x, k = torch.randn(6, 7), torch.randn(3, 2)
    valid_x = ty.DenseTensorConv.conv2Dmul(1.0, x, k, 1, 1, "V", "X").t
    assert torch.allclose(valid_x, F.conv2d(x[None, None], k[None, None])[0, 0], atol=1e-5)
    valid_c = ty.DenseTensorConv.conv2Dmul(2.0, x, k, 1, 1, "V", "C").t
    assert torch.allclose(valid_c, 2 * F.conv2d(x[None, None], k.flip(0, 1)[None, None])[0, 0], atol=1e-5)
    full_c = ty.DenseTensorConv.conv2Dmul(1.0, x, k, 1, 1, "F", "C").t
    assert tuple(full_c.shape) == (8, 8)
    assert torch.allclose(full_c, F.conv2d(F.pad(x, (1, 1, 2, 2))[None, None], k.flip(0, 1)[None, None])[0, 0], atol=1e-5)
    assert torch.allclose(TT.Tensor(x).conv2(TT.Tensor(k), "F").t, full_c, atol=1e-5)
    assert ty.DenseTensorConv.convSize(7, 3, 2, "V") == 3 and ty.DenseTensorConv.convSize(3, 3, 2, "F") == 7

    t1 = torch.zeros(2, 3)
    ty.DenseTensorApply.apply2(t1, torch.arange(6.0).reshape(2, 3), lambda p, q: p + 2 * q)
    assert t1.tolist() == [[0, 2, 4], [6, 8, 10]]
    ty.DenseTensorApply.apply1(t1, lambda v: v / 2)
    assert t1.tolist() == [[0, 1, 2], [3, 4, 5]]
    ty.DenseTensorDimApply.dimApply2(t1, torch.ones(2, 3), 1, lambda row, other: [sum(row)] * len(row))
    assert t1.tolist() == [[3, 3, 3], [12, 12, 12]]
    with pytest.raises(NotImplementedError):
        ty.QuantizedTensorUnsupported().addmm
