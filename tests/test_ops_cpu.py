"""nn/ops forward-only operations (mirrors T/nn/ops/*Spec.scala)."""
import numpy as np
import pytest
import torch

from bigdl_b200.nn import ops as O
from bigdl_b200.utils.table import Table


def T(*a):
    return Table(*a)


def test_operations_have_no_backward():
    op = O.Ceil()
    assert op.forward(torch.tensor([1.2, -1.2])).tolist() == [2.0, -1.0]
    with pytest.raises(RuntimeError):
        op.backward(torch.zeros(1), torch.zeros(1))


def test_elementwise_and_compare():
    a, b = torch.tensor([1.0, 5.0, 3.0]), torch.tensor([2.0, 5.0, 1.0])
    assert O.Greater().forward(T(a, b)).tolist() == [False, False, True]
    assert O.LessEqual().forward(T(a, b)).tolist() == [True, True, False]
    assert O.Maximum().forward(T(a, b)).tolist() == [2.0, 5.0, 3.0]
    assert O.SquaredDifference().forward(T(a, b)).tolist() == [1.0, 0.0, 4.0]
    assert O.FloorDiv().forward(T(torch.tensor([7.0, -7.0]), torch.tensor([2.0, 2.0]))).tolist() == [3.0, -4.0]
    assert O.FloorMod().forward(T(torch.tensor([7.0, -7.0]), torch.tensor([3.0, 3.0]))).tolist() == [1.0, 2.0]
    assert O.TruncateDiv().forward(T(torch.tensor([-7.0]), torch.tensor([2.0]))).tolist() == [-3.0]
    assert O.ApproximateEqual(0.1).forward(T(a, a + 0.05)).all()
    assert O.IsNan().forward(torch.tensor([float("nan"), 1.0])).tolist() == [True, False]
    assert abs(float(O.L2Loss().forward(torch.tensor([3.0, 4.0]))) - 12.5) < 1e-6
    assert torch.allclose(O.Erf().forward(torch.tensor([0.5])), torch.erf(torch.tensor([0.5])))
    op = O.TensorOp().add(1.0).pow(2.0) >> O.TensorOp().sqrt()
    assert op.forward(torch.tensor([3.0])).tolist() == [4.0]


def test_reductions_shape_ops():
    t = torch.arange(24.0).reshape(2, 3, 4)
    assert tuple(O.Sum().forward(T(t, torch.tensor([0, 2]))).shape) == (3,)
    assert tuple(O.Prod(keep_dims=True).forward(T(t + 1, torch.tensor([1]))).shape) == (2, 1, 4)
    assert O.All().forward(T(torch.tensor([[1, 0], [1, 1]]), torch.tensor([1]))).tolist() == [False, True]
    assert O.Max().forward(T(t, torch.tensor([2]))).tolist() == t.max(2)[0].tolist()
    assert O.ArgMax().forward(T(torch.tensor([[1.0, 9.0], [7.0, 2.0]]), torch.tensor(1))).tolist() == [1, 0]
    assert O.Slice([0, 1, 0], [2, 2, -1]).forward(t).shape == (2, 2, 4)
    assert O.Tile().forward(T(torch.ones(2, 1), torch.tensor([1, 3]))).shape == (2, 3)
    assert O.Pad().forward(T(torch.ones(2, 2), torch.tensor([[1, 0], [0, 2]]))).shape == (3, 4)
    assert O.RangeOps().forward(T(torch.tensor(0), torch.tensor(10), torch.tensor(3))).tolist() == [0, 3, 6, 9]
    assert O.Gather().forward(T(t, torch.tensor([1, 0]))).shape == (2, 3, 4)
    v, i = O.TopK(2).forward(torch.tensor([[1.0, 5.0, 3.0]])).to_list()
    assert v.tolist() == [[5.0, 3.0]] and i.tolist() == [[2, 3]]
    oh = O.OneHot().forward(T(torch.tensor([0, 2, -1]), torch.tensor(3), torch.tensor(1.0), torch.tensor(0.0)))
    assert oh.tolist() == [[1, 0, 0], [0, 0, 1], [0, 0, 0]]
    assert O.InTopK(1).forward(T(torch.tensor([[0.1, 0.9], [0.8, 0.2]]), torch.tensor([2, 2]))).tolist() == [True, False]
    assert O.SegmentSum().forward(T(torch.ones(4, 2), torch.tensor([0, 0, 1, 1]))).tolist() == [[2, 2], [2, 2]]
    assert O.SelectOp().forward(T(torch.tensor([True, False]), torch.tensor([1, 2]), torch.tensor([3, 4]))).tolist() == [1, 4]
    assert O.BatchMatMul(adj_y=True).forward(T(torch.ones(2, 3, 4), torch.ones(2, 5, 4))).shape == (2, 3, 5)


def test_conv_like_ops_match_torch():
    x = torch.randn(1, 6, 6, 2)
    w = torch.randn(3, 3, 2, 2)
    y = O.DepthwiseConv2D().forward(T(x, w))
    ref = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), w.permute(2, 3, 0, 1).reshape(4, 1, 3, 3), groups=2)
    assert torch.allclose(y, ref.permute(0, 2, 3, 1), atol=1e-5)
    d = O.Dilation2D([1, 1, 1, 1], [1, 1, 1, 1]).forward(T(x, torch.zeros(2, 2, 2)))
    assert torch.allclose(d, torch.nn.functional.max_pool2d(x.permute(0, 3, 1, 2), 2, 1).permute(0, 2, 3, 1))
    r = O.ResizeBilinearOps().forward(T(x, torch.tensor([12, 12])))
    assert r.shape == (1, 12, 12, 2)
    ce = O.CrossEntropy().forward(T(torch.tensor([[2.0, 1.0]]), torch.tensor([[1.0, 0.0]])))
    assert torch.allclose(ce, -torch.log_softmax(torch.tensor([[2.0, 1.0]]), 1)[:, 0])


def test_feature_column_ops():
    assert O.BucketizedCol([0.0, 10.0, 100.0]).forward(torch.tensor([[-1.0, 1.0], [101.0, 10.0]])).tolist() == [[0, 1], [3, 2]]
    col = np.array(["a,b", "c", ""], dtype=object)
    hb = O.CategoricalColHashBucket(100).forward(col)
    assert hb.size() == [3, 2] and hb.nElement() == 3 and int(hb.values().max()) < 100
    voc = O.CategoricalColVocaList(["a", "b", "c"], num_oov_buckets=2).forward(np.array(["a,z", "c,b"], dtype=object))
    d = voc.to_dense().t
    assert d[1].tolist() == [2, 1] and d[0, 0] == 0 and d[0, 1] in (3, 4)
    with pytest.raises(ValueError):
        O.CategoricalColVocaList(["a", "a"])
    cr = O.CrossCol(50).forward(T(np.array(["a,b"], dtype=object), np.array(["x"], dtype=object)))
    assert cr.nElement() == 2
    ind = O.IndicatorCol(5).forward(voc)
    assert ind.shape == (2, 5) and ind[1].tolist() == [0, 1, 1, 0, 0]
    kv = O.Kv2Tensor().forward(T(np.array(["0:0.5,2:1.5", "1:2"], dtype=object), torch.tensor(3)))
    assert kv.tolist() == [[0.5, 0, 1.5], [0, 2, 0]]
    assert O.MkString().forward(voc).tolist()[1] == "2,1"


def test_dynamic_graph_cond_and_while_loop():
    """Switch/Merge conditional and an Enter/Merge/Switch/NextIteration/Exit loop (T/nn/DynamicGraphSpec.scala)."""
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn import tf as TF
    from bigdl_b200.nn.dynamic_graph import DynamicGraph
    from bigdl_b200.nn.module import Node

    # cond: y = x * 2 if pred else x + 10
    x, pred = L.Input("x"), L.Input("pred")
    sw = TF.ControlNodes.switch(x, pred)
    t_branch = Node(L.MulConstant(2.0), [])
    sw.add(t_branch, 2)
    f_branch = Node(L.AddConstant(10.0), [])
    sw.add(f_branch, 1)
    out = TF.ControlNodes.merge(t_branch, f_branch)
    g = DynamicGraph([x, pred], [out])
    assert g.forward([torch.tensor([1.0, 2.0]), torch.tensor(True)]).tolist() == [2.0, 4.0]
    assert g.forward([torch.tensor([1.0, 2.0]), torch.tensor(False)]).tolist() == [11.0, 12.0]

    # while i < 5: i += 1; acc = acc * 2
    i0, a0 = L.Input("i"), L.Input("acc")
    limit = Node(TF.Const(torch.tensor(5.0)), [])

    def cond(vars_):
        lim = Node(TF.Enter("w", is_constant=True), [limit])
        return Node(O.Less(), [vars_[0], lim])

    def body(vars_):
        return [Node(L.AddConstant(1.0), [vars_[0]]), Node(L.MulConstant(2.0), [vars_[1]])]

    exits = TF.ControlNodes.whileLoop(cond, body, [i0, a0], "w")
    loop = DynamicGraph([i0, a0], exits)
    i_out, a_out = loop.forward([torch.tensor(0.0), torch.tensor(1.0)]).to_list()
    assert float(i_out) == 5.0 and float(a_out) == 32.0
    i_out, a_out = loop.forward([torch.tensor(7.0), torch.tensor(3.0)]).to_list()      # zero iterations
    assert float(i_out) == 7.0 and float(a_out) == 3.0
    # gradients flow through the executed iterations
    a = torch.tensor(1.5, requires_grad=True)
    with torch.enable_grad():
        res = loop._call(Table(torch.tensor(2.0), a))
    res[2].backward()
    assert float(a.grad) == 8.0            # three doublings


def test_tf_helper_ops():
    from bigdl_b200.nn import tf as TF

    t = torch.arange(24.0).reshape(2, 3, 4)
    assert TF.StridedSlice([(2, 1, 3, 1), (3, 1, 4, 2)]).forward(t).shape == (2, 2, 2)
    assert TF.SplitAndSelect(3, 2, 2).forward(t).tolist() == t[:, :, 2:].tolist()
    assert TF.Shape().forward(t).tolist() == [2, 3, 4]
    assert TF.Fill().forward(T(torch.tensor([2, 2]), torch.tensor(7.0))).tolist() == [[7, 7], [7, 7]]
    assert TF.InvertPermutation().forward(torch.tensor([2, 0, 1])).tolist() == [1, 2, 0]
    ra, rb = TF.BroadcastGradientArgs().forward(T(torch.tensor([2, 1, 4]), torch.tensor([4]))).to_list()
    assert ra.tolist() == [1] and rb.tolist() == [0, 1]
    with pytest.raises(AssertionError):
        TF.Assert().forward(T(torch.tensor(False), "boom"))
    h, flow = TF.TensorArrayCreator().forward(torch.tensor(2)).to_list()
    TF.TensorArrayWrite().forward(T(h, torch.tensor(0), torch.ones(2), flow))
    TF.TensorArrayWrite().forward(T(h, torch.tensor(1), torch.zeros(2), flow))
    assert TF.TensorArrayGather().forward(T(h, torch.tensor([1, 0]), flow)).tolist() == [[0, 0], [1, 1]]
    assert int(TF.TensorArraySize().forward(T(h, flow))) == 2
    # tf.train.Example: {"x": FloatList[1.5, 2.5], "id": Int64List[7]}
    import struct
    def ld(field, payload): return bytes([(field << 3) | 2, len(payload)]) + payload
    fx = ld(2, ld(1, struct.pack("<2f", 1.5, 2.5)))
    fid = ld(3, bytes([0x08, 7]))
    feats = ld(1, ld(1, b"x") + ld(2, fx)) + ld(1, ld(1, b"id") + ld(2, fid))
    ex = ld(1, feats)
    parsed = TF.parse_example(ex)
    assert parsed["x"] == ("float", [1.5, 2.5]) and parsed["id"] == ("int64", [7])
    outs = TF.ParseExample(2, ["float", "int64"], [[2], [1]]).forward(
        T(np.array([ex, ex], dtype=object), np.array(["", ""], dtype=object), "x", "id"))
    assert outs[1].tolist() == [[1.5, 2.5], [1.5, 2.5]] and outs[2].tolist() == [[7], [7]]
    import cv2
    ok, png = cv2.imencode(".png", np.arange(48, dtype=np.uint8).reshape(4, 4, 3))
    img = TF.DecodePng(3).forward(np.array([png.tobytes()], dtype=object))
    assert img.shape == (4, 4, 3) and img[0, 0].tolist() == [2, 1, 0]          # RGB order
    raw = TF.DecodeRaw(torch.int16).forward(np.array([struct.pack("<3h", 1, -2, 3)], dtype=object))
    assert raw.tolist() == [1, -2, 3]


def test_fanout_fn_sums_branch_gradients():
    """FanOutFn is an identity whose backward is the sum of its consumers' gradients (torch path on CPU)."""
    import torch
    from bigdl_b200.ops.functions import FanOutFn

    x = torch.randn(2, 8, 4, 4, requires_grad=True)
    a, b, c = FanOutFn.apply(x, 3)
    (a * 2 + b.pow(2) + c.sin()).sum().backward()
    ref = 2 + 2 * x.detach() + x.detach().cos()
    assert torch.allclose(x.grad, ref, atol=1e-6)
    y = torch.randn(3, 5, requires_grad=True)
    (only,) = FanOutFn.apply(y, 1)
    only.sum().backward()
    assert torch.equal(y.grad, torch.ones_like(y))


def test_elementwise_and_logical_ops_against_torch():
    """The TF-style forward-only ops no other test constructs (S/nn/ops/*.scala), each against its torch equivalent."""
    import torch
    import bigdl_b200.nn.ops as O
    from bigdl_b200.utils.table import T

    x = torch.tensor([[-1.5, 0.4, 2.5], [3.0, -0.2, 0.5]])
    y = torch.tensor([[1.0, 0.4, -2.5], [2.0, 5.0, 0.5]])
    p = x.abs() + 0.5
    special = torch.tensor([1.0, float("inf"), float("nan")])
    cases = [
        (O.Floor(), x, torch.floor(x)), (O.Rint(), x, torch.round(x)), (O.Sign(), x, torch.sign(x)),
        (O.Expm1(), x, torch.expm1(x)), (O.Inv(), x, 1 / x), (O.Erfc(), x, torch.erfc(x)),
        (O.Lgamma(), p, torch.lgamma(p)), (O.Digamma(), p, torch.digamma(p)),
        (O.IsFinite(), special, torch.tensor([True, False, False])),
        (O.IsInf(), special, torch.tensor([False, True, False])),
        (O.LogicalNot(), torch.tensor([True, False]), torch.tensor([False, True])),
        (O.Equal(), T(x, y), x == y), (O.NotEqual(), T(x, y), x != y), (O.GreaterEqual(), T(x, y), x >= y),
        (O.LogicalAnd(), T(x > 0, y > 0), (x > 0) & (y > 0)), (O.LogicalOr(), T(x > 0, y > 0), (x > 0) | (y > 0)),
        (O.Minimum(), T(x, y), torch.minimum(x, y)), (O.Pow(), T(x.abs(), y), x.abs() ** y),
    ]
    for op, inp, ref in cases:
        out = op.forward(inp)
        assert out.shape == ref.shape and torch.allclose(out.float(), ref.float(), atol=1e-5), type(op).__name__
    assert int(O.Rank().forward(x)) == 2
    assert O.Round().forward(x).shape == x.shape and O.Mod().forward(T(x, y)).shape == x.shape
    assert O.Cast("int").forward(x).dtype == torch.int32 and O.Cast(torch.float64).forward(x).dtype == torch.float64
    u = O.RandomUniform(0.0, 1.0).forward(torch.tensor([200, 3]))
    assert tuple(u.shape) == (200, 3) and 0.0 <= float(u.min()) and float(u.max()) <= 1.0
    tn = O.TruncatedNormal(0.0, 1.0).forward(torch.tensor([500, 2]))
    assert tuple(tn.shape) == (500, 2) and float(tn.abs().max()) <= 2.0 + 1e-5
