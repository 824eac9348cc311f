"""The backward-half TensorFlow op loaders (interop/tensorflow_grad.py) against torch autograd / closed forms, built from
NodeDefs the way a training GraphDef carries them (reference: S/utils/tf/loaders/*Grad.scala and their specs)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from bigdl_b200.interop import tf_pb as P
from bigdl_b200.interop.tensorflow import SUPPORTED_OPS, _LOADERS
from bigdl_b200.utils.table import Table


def _node(op, **attrs):
    n = P.NodeDef()
    n.name, n.op = op.lower(), op
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
        elif isinstance(v, (list, tuple)):
            a.list.i.extend(int(i) for i in v)
    return n


def _run(op, inputs, **attrs):
    m = _LOADERS[op](_node(op, **attrs))
    return m.forward(Table(*inputs) if len(inputs) > 1 else inputs[0])


def test_every_reference_loader_has_a_counterpart():
    import os
    d = "/root/reference/spark/dl/src/main/scala/com/intel/analytics/bigdl/utils/tf/loaders"
    if not os.path.isdir(d):
        pytest.skip("reference tree not available")
    skip = {"Utils", "TensorflowOpsLoader", "Adapter", "ArrayOps", "ControlFlowOps", "DataFlowOps", "DependencyNode"}
    ref = sorted(f[:-6] for f in os.listdir(d) if f.endswith(".scala"))
    # queue ops are consumed by interop/tf_session.py (the input pipeline), not turned into modules
    queue = {"QueueDequeueManyV2", "QueueDequeueV2", "QueueEnqueueManyV2", "QueueEnqueueV2", "ReaderReadV2"}
    missing = [o for o in ref if o not in skip and o not in queue and o not in SUPPORTED_OPS]
    assert not missing, missing


@pytest.mark.parametrize("op,fwd", [("ReluGrad", torch.relu), ("Relu6Grad", F.relu6), ("SoftplusGrad", F.softplus),
                                    ("SoftsignGrad", F.softsign)])
def test_feature_based_activation_grads(op, fwd):
    torch.manual_seed(0)
    x = (torch.randn(5, 7) * 3).requires_grad_(True)
    g = torch.randn(5, 7)
    fwd(x).backward(g)
    assert torch.allclose(_run(op, [g, x.detach()]), x.grad, atol=1e-6)


@pytest.mark.parametrize("op,fwd", [("SigmoidGrad", torch.sigmoid), ("TanhGrad", torch.tanh), ("SqrtGrad", torch.sqrt),
                                    ("RsqrtGrad", torch.rsqrt), ("ReciprocalGrad", torch.reciprocal), ("InvGrad", torch.reciprocal)])
def test_output_based_grads(op, fwd):
    torch.manual_seed(1)
    x = (torch.rand(4, 6) + 0.5).requires_grad_(True)
    y = fwd(x)
    g = torch.randn(4, 6)
    y.backward(g)
    assert torch.allclose(_run(op, [y.detach(), g]), x.grad, rtol=1e-5, atol=1e-6)


def test_elu_grad_bias_add_grad_and_broadcast_args():
    x = torch.randn(3, 5, requires_grad=True)
    y = F.elu(x)
    g = torch.randn(3, 5)
    y.backward(g)
    assert torch.allclose(_run("EluGrad", [g, y.detach()]), x.grad, atol=1e-6)
    gb = torch.randn(2, 4, 4, 3)
    assert torch.allclose(_run("BiasAddGrad", [gb]), gb.sum((0, 1, 2)))
    assert torch.allclose(_run("BiasAddGrad", [gb.permute(0, 3, 1, 2)], data_format="NCHW"), gb.sum((0, 1, 2)))
    r = _run("BroadcastGradientArgs", [torch.tensor([2, 3, 5]), torch.tensor([5])])
    assert r[1].tolist() == [] and r[2].tolist() == [0, 1]
    r = _run("BroadcastGradientArgs", [torch.tensor([4, 1]), torch.tensor([1, 6])])
    assert r[1].tolist() == [1] and r[2].tolist() == [0]


@pytest.mark.parametrize("pad", ["VALID", "SAME"])
def test_pool_grads(pad):
    torch.manual_seed(2)
    x = torch.randn(2, 7, 7, 3)
    xr = x.permute(0, 3, 1, 2).clone().requires_grad_(True)
    if pad == "VALID":
        y = F.max_pool2d(xr, 3, 2)
        ya = F.avg_pool2d(xr, 3, 2)
    else:
        y = F.max_pool2d(F.pad(xr, (1, 1, 1, 1), value=float("-inf")), 3, 2)
        ya = None
    g = torch.randn_like(y)
    y.backward(g)
    got = _run("MaxPoolGrad", [x, y.detach().permute(0, 2, 3, 1), g.permute(0, 2, 3, 1)], ksize=[1, 3, 3, 1],
               strides=[1, 2, 2, 1], padding=pad)
    assert torch.allclose(got, xr.grad.permute(0, 2, 3, 1), atol=1e-6)
    if ya is not None:
        xr.grad = None
        ga = torch.randn_like(ya)
        ya.backward(ga)
        got = _run("AvgPoolGrad", [torch.tensor([2, 7, 7, 3]), ga.permute(0, 2, 3, 1)], ksize=[1, 3, 3, 1],
                   strides=[1, 2, 2, 1], padding=pad)
        assert torch.allclose(got, xr.grad.permute(0, 2, 3, 1), atol=1e-6)


def test_conv_backprop_filter_and_conv3d_family():
    torch.manual_seed(3)
    x = torch.randn(2, 6, 6, 4)
    w = torch.randn(3, 3, 4, 5, requires_grad=True)
    y = F.conv2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), None, 1, 1)
    g = torch.randn_like(y)
    y.backward(g)
    got = _run("Conv2DBackpropFilter", [x, torch.tensor([3, 3, 4, 5]), g.permute(0, 2, 3, 1)], strides=[1, 1, 1, 1],
               padding="SAME")
    assert torch.allclose(got, w.grad, atol=1e-4)
    # 3-D
    x3 = torch.randn(1, 5, 5, 5, 2, requires_grad=True)
    w3 = torch.randn(2, 2, 2, 2, 3, requires_grad=True)
    ref = F.conv3d(x3.permute(0, 4, 1, 2, 3), w3.permute(4, 3, 0, 1, 2))
    out = _run("Conv3D", [x3.detach(), w3.detach()], strides=[1, 1, 1, 1, 1], padding="VALID")
    assert torch.allclose(out, ref.permute(0, 2, 3, 4, 1).detach(), atol=1e-5)
    g3 = torch.randn_like(ref)
    ref.backward(g3)
    gi = _run("Conv3DBackpropInputV2", [torch.tensor(list(x3.shape)), w3.detach(), g3.permute(0, 2, 3, 4, 1)],
              strides=[1, 1, 1, 1, 1], padding="VALID")
    gf = _run("Conv3DBackpropFilterV2", [x3.detach(), torch.tensor(list(w3.shape)), g3.permute(0, 2, 3, 4, 1)],
              strides=[1, 1, 1, 1, 1], padding="VALID")
    assert torch.allclose(gi, x3.grad, atol=1e-4) and torch.allclose(gf, w3.grad, atol=1e-4)


def test_depthwise_backprops_and_dilation2d():
    torch.manual_seed(4)
    x = torch.randn(2, 6, 6, 3, requires_grad=True)
    w = torch.randn(3, 3, 3, 2, requires_grad=True)
    wt = w.permute(2, 3, 0, 1).reshape(6, 1, 3, 3)
    y = F.conv2d(x.permute(0, 3, 1, 2), wt, None, 1, 0, 1, 3)
    g = torch.randn_like(y)
    y.backward(g)
    gi = _run("DepthwiseConv2dNativeBackpropInput", [torch.tensor(list(x.shape)), w.detach(), g.permute(0, 2, 3, 1)],
              strides=[1, 1, 1, 1], padding="VALID")
    gf = _run("DepthwiseConv2dNativeBackpropFilter", [x.detach(), torch.tensor(list(w.shape)), g.permute(0, 2, 3, 1)],
              strides=[1, 1, 1, 1], padding="VALID")
    assert torch.allclose(gi, x.grad, atol=1e-4) and torch.allclose(gf, w.grad, atol=1e-4)
    # dilation: every output picks one (input + filter) pair, so input and filter gradients carry the same mass
    xd, wd = torch.randn(1, 5, 5, 2), torch.randn(2, 2, 2)
    gd = torch.rand(1, 4, 4, 2)
    a = _run("Dilation2DBackpropInput", [xd, wd, gd], strides=[1, 1, 1, 1], rates=[1, 1, 1, 1], padding="VALID")
    b = _run("Dilation2DBackpropFilter", [xd, wd, gd], strides=[1, 1, 1, 1], rates=[1, 1, 1, 1], padding="VALID")
    assert a.shape == xd.shape and b.shape == wd.shape
    assert abs(float(a.sum()) - float(gd.sum())) < 1e-4 and abs(float(b.sum()) - float(gd.sum())) < 1e-4


def test_fused_batch_norm_grad_lrn_grad_resize_grad_xent():
    torch.manual_seed(5)
    x = torch.randn(4, 3, 3, 5, requires_grad=True)
    scale = torch.rand(5, requires_grad=True) + 0.5
    scale.retain_grad()
    mu, var = x.mean((0, 1, 2)), x.var((0, 1, 2), unbiased=False)
    y = (x - mu) * torch.rsqrt(var + 1e-3) * scale
    g = torch.randn_like(y)
    y.backward(g)
    out = _run("FusedBatchNormGrad", [g, x.detach(), scale.detach(), mu.detach(), var.detach()], epsilon=1e-3, is_training=True)
    assert torch.allclose(out[1], x.grad, atol=1e-4) and torch.allclose(out[2], scale.grad, atol=1e-4)
    assert torch.allclose(out[3], g.sum((0, 1, 2)), atol=1e-5)
    # LRN
    xl = torch.rand(2, 4, 4, 8, requires_grad=True)
    yl = F.local_response_norm(xl.permute(0, 3, 1, 2), 5, 1e-2 * 5, 0.75, 2.0).permute(0, 2, 3, 1)
    gl = torch.randn_like(yl)
    yl.backward(gl)
    got = _run("LRNGrad", [gl, xl.detach(), yl.detach()], depth_radius=2, bias=2.0, alpha=1e-2, beta=0.75)
    assert torch.allclose(got, xl.grad, atol=1e-5)
    # resize
    xr = torch.randn(1, 3, 3, 2, requires_grad=True)
    yr = F.interpolate(xr.permute(0, 3, 1, 2), size=(6, 6), mode="bilinear", align_corners=False).permute(0, 2, 3, 1)
    gr = torch.randn_like(yr)
    yr.backward(gr)
    assert torch.allclose(_run("ResizeBilinearGrad", [gr, xr.detach()], align_corners=False), xr.grad, atol=1e-5)
    # softmax cross entropy: (loss, backprop)
    lg, lb = torch.randn(6, 10), F.one_hot(torch.randint(0, 10, (6,)), 10).float()
    out = _run("SoftmaxCrossEntropyWithLogits", [lg, lb])
    assert torch.allclose(out[1], -(lb * F.log_softmax(lg, -1)).sum(-1), atol=1e-5)
    assert torch.allclose(out[2], F.softmax(lg, -1) - lb, atol=1e-6)


def _const(name, arr):
    n = P.NodeDef()
    n.name, n.op = name, "Const"
    n.attr["value"].tensor.CopyFrom(P.numpy_to_tensor(np.asarray(arr)))
    n.attr["dtype"].type = n.attr["value"].tensor.dtype
    return n


def _n(name, op, inputs, **attrs):
    n = _node(op, **attrs)
    n.name = name
    n.input.extend(inputs)
    return n


def test_session_trains_from_the_graphs_own_queue_pipeline(tmp_path):
    """Session.train without a DataSet (S/utils/tf/Session.scala:112-420): the data comes from QueueEnqueueMany ->
    FIFOQueue -> QueueDequeueMany inside the GraphDef; and from a TFRecord reader + ParseExample pipeline."""
    from bigdl_b200.interop.tf_session import Session
    from bigdl_b200.interop.tfrecord import TFRecordWriter, encode_example
    from bigdl_b200.nn.criterion import MSECriterion
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.optim.trigger import MaxEpoch

    rng = np.random.RandomState(0)
    X = rng.randn(64, 4).astype(np.float32)
    Wt = rng.randn(4, 1).astype(np.float32)
    Y = X @ Wt
    g = P.GraphDef()
    q = _n("queue", "FIFOQueueV2", [])
    q.attr["component_types"].list.type.extend([1, 1])
    g.node.extend([_const("xdata", X), _const("ydata", Y), q,
                   _n("enq", "QueueEnqueueManyV2", ["queue", "xdata", "ydata"]),
                   _const("n", np.int32(8)), _n("deq", "QueueDequeueManyV2", ["queue", "n"]),
                   _const("w", np.zeros((4, 1), np.float32)), _n("w/read", "Identity", ["w"]),
                   _n("output", "MatMul", ["deq", "w/read"])])
    s = Session(g)
    names, rows = s.queue_dataset(["output"])
    assert names == ["deq", "deq:1"] and len(rows) == 64 and tuple(rows[0][0].shape) == (4,)
    model = s.train(["output"], None, SGD(0.1), MSECriterion(), MaxEpoch(30), batch_size=16)
    w = [p for p in model.torch_parameters()][0].detach().reshape(4, 1).numpy()
    assert np.abs(w - Wt).max() < 0.05, w.ravel()

    # reader pipeline: string_input_producer -> TFRecordReaderV2 -> ParseExample -> queue
    path = str(tmp_path / "d.tfrecord")
    wr = TFRecordWriter(path)
    for i in range(32):
        wr.write(encode_example({"x": X[i].tolist(), "y": Y[i].tolist()}))
    wr.close()
    g2 = P.GraphDef()
    fq = _n("fq", "FIFOQueueV2", [])
    fq.attr["component_types"].list.type.extend([7])
    q2 = _n("queue", "FIFOQueueV2", [])
    q2.attr["component_types"].list.type.extend([1, 1])
    pe = _n("parse", "ParseExample", ["read:1", "names", "kx", "ky", "dx", "dy"], Ndense=2)
    pe.attr["Tdense"].list.type.extend([1, 1])
    for dims in ([4], [1]):
        sh = pe.attr["dense_shapes"].list.shape.add()
        for d in dims:
            sh.dim.add().size = d
    g2.node.extend([_const("files", np.array([path.encode()], dtype=object)), fq,
                    _n("fenq", "QueueEnqueueManyV2", ["fq", "files"]), _n("reader", "TFRecordReaderV2", []),
                    _n("read", "ReaderReadV2", ["reader", "fq"]),
                    _const("names", np.array([], dtype=object)), _const("kx", np.array(b"x", dtype=object)),
                    _const("ky", np.array(b"y", dtype=object)), _const("dx", np.zeros(0, np.float32)),
                    _const("dy", np.zeros(0, np.float32)), pe, q2,
                    _n("enq", "QueueEnqueueManyV2", ["queue", "parse", "parse:1"]),
                    _const("n", np.int32(8)), _n("deq", "QueueDequeueManyV2", ["queue", "n"]),
                    _const("w", np.zeros((4, 1), np.float32)), _n("w/read", "Identity", ["w"]),
                    _n("output", "MatMul", ["deq", "w/read"])])
    names, rows = Session(g2).queue_dataset(["output"])
    assert len(rows) == 32
    got = np.stack([np.asarray(r[0]) for r in rows])
    assert np.allclose(got, X[:32], atol=1e-6) and np.allclose(np.stack([np.asarray(r[1]) for r in rows]), Y[:32], atol=1e-6)


# ------------------------------------------------------------------------------------------------ resources
def _graph(nodes):
    """nodes: [(op, name, inputs, attrs)] → serialized GraphDef file contents."""
    from bigdl_b200.interop.tensorflow import _GraphBuilder

    b = _GraphBuilder()
    for op_name, name, ins, attrs in nodes:
        if op_name == "Const":
            b.const(name, attrs["value"])
        else:
            b.node(op_name, name, ins, **attrs)
    return b.gd.SerializeToString()


def test_tensor_array_and_stack_graph(tmp_path):
    from bigdl_b200.interop.tensorflow import load_tf

    f32 = np.dtype(np.float32)
    nodes = [
        ("Placeholder", "x", [], {"dtype": f32}),
        ("Const", "size", [], {"value": np.array(2, np.int32)}),
        ("Const", "i0", [], {"value": np.array(0, np.int32)}),
        ("Const", "i1", [], {"value": np.array(1, np.int32)}),
        ("Const", "two", [], {"value": np.array(2.0, np.float32)}),
        ("Const", "idx", [], {"value": np.array([1, 0], np.int32)}),
        ("TensorArrayV3", "ta", ["size"], {"dtype": f32, "dynamic_size": False}),
        ("Mul", "x2", ["x", "two"], {"T": f32}),
        ("TensorArrayWriteV3", "w0", ["ta", "i0", "x", "ta:1"], {"T": f32}),
        ("TensorArrayWriteV3", "w1", ["ta", "i1", "x2", "w0"], {"T": f32}),
        ("TensorArrayGatherV3", "gathered", ["ta", "idx", "w1"], {"dtype": f32}),
        ("TensorArraySizeV3", "n", ["ta", "w1"], {}),
        ("TensorArrayReadV3", "r1", ["ta", "i1", "w1"], {"dtype": f32}),
        ("TensorArrayConcatV3", "cat", ["ta", "w1"], {"dtype": f32}),
    ]
    p = tmp_path / "ta.pb"
    p.write_bytes(_graph(nodes))
    x = torch.tensor([1.0, 2.0, 3.0])
    g = load_tf(str(p), ["x"], ["gathered", "n", "r1", "cat:0", "cat:1"])
    out = g.forward(x)
    gathered, n, r1, cat, lens = [out[i] for i in range(1, 6)]
    assert torch.equal(gathered, torch.stack([2 * x, x])) and int(n) == 2 and torch.equal(r1, 2 * x)
    assert torch.equal(cat, torch.cat([x, 2 * x])) and lens.tolist() == [3, 3]

    nodes = [
        ("Placeholder", "x", [], {"dtype": f32}),
        ("Const", "max", [], {"value": np.array(-1, np.int32)}),
        ("Const", "two", [], {"value": np.array(2.0, np.float32)}),
        ("StackV2", "st", ["max"], {"elem_type": f32}),
        ("StackPushV2", "p0", ["st", "x"], {"T": f32}),
        ("Mul", "x2", ["p0", "two"], {"T": f32}),
        ("StackPushV2", "p1", ["st", "x2"], {"T": f32}),
        ("Identity", "after", ["p1"], {"T": f32}),
        ("StackPopV2", "top", ["st", "^after"], {"elem_type": f32}),
    ]
    p2 = tmp_path / "st.pb"
    p2.write_bytes(_graph(nodes))
    g = load_tf(str(p2), ["x"], ["after", "top"])
    out = g.forward(x)
    assert torch.equal(out[1], 2 * x) and torch.equal(out[2], 2 * x)


def test_tensor_array_grad_accumulates_and_close():
    from bigdl_b200.nn import tf as TFO

    handle, flow = TFO.TensorArrayCreator(False).forward(torch.tensor(2)).to_list()
    TFO.TensorArrayWrite().forward(Table(handle, torch.tensor(0), torch.ones(2), flow))
    gh, gflow = _LOADERS["TensorArrayGradV3"](_node("TensorArrayGradV3", source="grad1")).forward(Table(handle, flow)).to_list()
    assert gh is not handle and len(gh.items) == 2 and not handle.dynamic
    w = TFO.TensorArrayWrite()
    w.forward(Table(gh, torch.tensor(1), torch.ones(3), gflow))
    w.forward(Table(gh, torch.tensor(1), 2 * torch.ones(3), gflow))          # second write to one slot is summed
    assert gh.items[1].tolist() == [3.0, 3.0, 3.0]
    again = TFO.TensorArrayGrad("grad1").forward(Table(handle, flow))[1]
    assert again is gh and TFO.TensorArrayGrad("other").forward(Table(handle, flow))[1] is not gh
    with pytest.raises(IndexError):
        TFO.TensorArrayWrite().forward(Table(handle, torch.tensor(5), torch.ones(1), flow))
    _LOADERS["TensorArrayCloseV3"](_node("TensorArrayCloseV3")).forward(handle)
    assert handle.items == []
    perm = _LOADERS["InvertPermutation"](_node("InvertPermutation")).forward(torch.tensor([2, 0, 1]))
    assert perm.tolist() == [1, 2, 0]
    offs = _LOADERS["ConcatOffset"](_node("ConcatOffset")).forward(
        Table(torch.tensor(1), torch.tensor([2, 3]), torch.tensor([2, 5])))
    assert offs[1].tolist() == [0, 0] and offs[2].tolist() == [0, 3]
    for name in ("RefEnter", "RefSwitch", "RefMerge", "RefExit", "RefNextIteration", "ControlTrigger", "Stack", "StackPush",
                 "StackPop"):
        assert name in _LOADERS


# ------------------------------------------------------------------------------------------------ nn.tf / nn.ops classes
def test_grad_ops_as_layers():
    from bigdl_b200.nn import ops as O
    from bigdl_b200.nn import tf as TFO
    from bigdl_b200.nn import tf_grad as G

    torch.manual_seed(0)
    x = torch.randn(2, 6, 6, 3, requires_grad=True)
    g = torch.randn(2, 6, 6, 3)
    torch.tanh(x).backward(g)
    assert torch.allclose(TFO.TanhGrad().forward(Table(torch.tanh(x.detach()), g)), x.grad, atol=1e-6)
    assert TFO.ReluGrad is G.ReluGrad and O.MaxPoolGrad is G.MaxPoolGrad and O.SelectTensor is G.SelectTensor
    with pytest.raises(AttributeError):
        TFO.NoSuchOp

    w = torch.randn(3, 3, 3, 4)
    y = G.Conv2D(1, 1, -1, -1, "NHWC").forward(Table(x.detach(), w))
    ref = F.conv2d(x.detach().permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), padding=1)
    assert torch.allclose(y, ref.permute(0, 2, 3, 1), atol=1e-5)
    gy = torch.randn_like(y)
    xr = x.detach().permute(0, 3, 1, 2).requires_grad_(True)
    wr = w.permute(3, 2, 0, 1).detach().requires_grad_(True)
    F.conv2d(xr, wr, padding=1).backward(gy.permute(0, 3, 1, 2))
    dx = G.Conv2DTranspose(1, 1, -1, -1, "NHWC").forward(Table(torch.tensor([2, 6, 6, 3]), w, gy))
    dw = G.Conv2DBackFilter(1, 1, -1, -1, "NHWC").forward(Table(x.detach(), torch.tensor([3, 3, 3, 4]), gy))
    assert torch.allclose(dx, xr.grad.permute(0, 2, 3, 1), atol=1e-4)
    assert torch.allclose(dw, wr.grad.permute(2, 3, 1, 0), atol=1e-4)

    p = G.MaxPool([1, 2, 2, 1], [1, 2, 2, 1], "VALID").forward(x.detach())
    xp = x.detach().permute(0, 3, 1, 2).requires_grad_(True)
    F.max_pool2d(xp, 2, 2).backward(torch.ones(2, 3, 3, 3))
    mg = G.MaxPoolGrad(2, 2, 2, 2, 0, 0, "NHWC").forward(Table(x.detach(), p, torch.ones_like(p)))
    assert torch.allclose(mg, xp.grad.permute(0, 2, 3, 1))
    ag = G.AvgPoolGrad(2, 2, 2, 2, 0, 0, "NHWC").forward(Table(torch.tensor([2, 6, 6, 3]), torch.ones_like(p)))
    assert torch.allclose(ag, torch.full((2, 6, 6, 3), 0.25))
    assert torch.allclose(G.BiasAddGrad("NHWC").forward(g), g.sum((0, 1, 2)))

    out = G.FusedBatchNorm(1e-3, True).forward(Table(x.detach(), torch.ones(3), torch.zeros(3), torch.zeros(0), torch.zeros(0)))
    yb, bm, bv, sm, sv = out.to_list()
    flat = x.detach().reshape(-1, 3)
    assert torch.allclose(bm, flat.mean(0), atol=1e-6) and torch.allclose(sv, flat.var(0, unbiased=False), atol=1e-6)
    assert torch.allclose(bv, flat.var(0, unbiased=True), atol=1e-6) and torch.allclose(yb.reshape(-1, 3).mean(0), torch.zeros(3), atol=1e-5)
    dxb, dscale, dbias = G.FusedBatchNormGrad(1e-3, "NHWC", True).forward(Table(g, x.detach(), torch.ones(3), sm, sv)).to_list()[:3]
    xb = x.detach().permute(0, 3, 1, 2).requires_grad_(True)
    sc = torch.ones(3, requires_grad=True)
    F.batch_norm(xb, None, None, sc, torch.zeros(3), True, 0.0, 1e-3).backward(g.permute(0, 3, 1, 2))
    assert torch.allclose(dxb, xb.grad.permute(0, 2, 3, 1), atol=1e-4) and torch.allclose(dscale, sc.grad, atol=1e-4)
    assert torch.allclose(dbias, g.sum((0, 1, 2)), atol=1e-5)

    assert G.Conv3D(1, 2, 2, -1, -1, -1, "NHWC").attrs == {"strides": [1, 1, 2, 2, 1], "padding": "SAME", "data_format": "NDHWC"}
    assert G.LRNGrad(2, 1.0, 1e-4, 0.75).attrs["depth_radius"] == 2 and G.ResizeBilinearGrad(True).attrs["align_corners"]
    for cls in (G.Conv3DBackpropInput, G.Conv3DBackpropInputV2, G.Conv3DBackpropFilter, G.Conv3DBackpropFilterV2):
        cls(1, 1, 1, 0, 0, 0, "NHWC")
    G.Dilation2DBackpropInput([1, 1, 1, 1], [1, 1, 1, 1], "SAME")
    G.Dilation2DBackpropFilter([1, 1, 1, 1], [1, 1, 1, 1], "SAME")
    G.DepthwiseConv2DBackpropInput(1, 1, -1, -1)
    G.DepthwiseConv2DBackpropFilter(1, 1, -1, -1)
    with pytest.raises(ValueError):
        G.DecodeJpeg(3, 2)


def test_small_state_and_wrapper_ops():
    import cv2

    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn import tf_grad as G
    from bigdl_b200.utils.table import T

    buf = torch.zeros(2, 2)
    assert G.AssignGrad(buf).forward(torch.arange(4.0)).tolist() == [0, 1, 2, 3] and buf.tolist() == [[0, 1], [2, 3]]
    tab = T()
    tab["a"], tab["b"] = torch.ones(2), torch.zeros(2)
    assert G.SelectTensor("b").forward(tab).tolist() == [0, 0]
    assert G.SelectTensor("a", L.MulConstant(3.0)).forward(tab).tolist() == [3, 3]
    with pytest.raises(KeyError):
        G.SelectTensor("zz").forward(tab)
    assert G.TensorModuleWrapper(L.ReLU()).forward(torch.tensor([-1.0, 2.0])).tolist() == [0, 2]
    img = (np.arange(5 * 4 * 3) % 255).astype(np.uint8).reshape(5, 4, 3)
    png = cv2.imencode(".png", img)[1].tobytes()
    dec = G.DecodePng(3).forward(np.array([png], dtype=object))
    assert tuple(dec.shape) == (5, 4, 3) and np.array_equal(dec.numpy().astype(np.uint8), img[..., ::-1])


def test_saver_registry_and_nodedef_builders(tmp_path):
    from bigdl_b200.interop import tensorflow as TF
    from bigdl_b200.nn import layer as L

    from bigdl_b200.nn.module import Module

    class Times(Module):
        def __init__(self, k):
            super().__init__()
            self.k = k

        def updateOutput(self, input):
            return input * self.k

    class TimesSaver(TF.BigDLToTensorflow):
        def toTFDef(self, module, inputs, byteOrder):
            assert len(inputs) == 1 and byteOrder == "little_endian"
            k = TF.Tensorflow.const(np.array(module.k, np.float32), module.getName() + "/k")
            return [TF.Tensorflow.multiply(inputs[0], k, module.getName()), k]

    m = L.Sequential().add(L.Linear(4, 3).setName("fc")).add(Times(3.0).setName("times")).add(L.Tanh().setName("act"))
    p = str(tmp_path / "m.pb")
    with pytest.raises(ValueError, match="unsupported module Times"):
        TF.TensorflowSaver.saveGraph(m, [("input", [2, 4])], p)
    TF.TensorflowSaver.register("my.pkg.Times", TimesSaver())
    try:
        outs = TF.TensorflowSaver.saveGraph(m, [("input", [2, 4])], p)
        g = TF.TensorflowLoader.load(p, ["input"], outs)
        x = torch.rand(2, 4)
        assert torch.allclose(g.forward(x), m.forward(x), atol=1e-6)
        assert {n.op for n in TF.TensorflowLoader.parse(p)} >= {"Placeholder", "MatMul", "Mul", "Tanh", "Const"}

        ph = TF.Tensorflow.placeholder(np.float32, [2, 4], "input")
        extra = TF.Tensorflow.identity(ph, "alias")
        p2 = str(tmp_path / "m2.pb")
        outs2 = TF.TensorflowSaver.saveGraphWithNodeDef(m, [ph], p2, extraNodes=[extra])
        g2 = TF.TensorflowLoader.load(p2, ["input"], outs2 + ["alias"])
        both = g2.forward(x)
        assert torch.allclose(both[1], m.forward(x), atol=1e-6) and torch.equal(both[2], x)
    finally:
        TF._CUSTOM_SAVERS.clear()

    T = TF.Tensorflow
    a, b2 = T.placeholder(np.float32, [2, 3], "a"), T.placeholder(np.float32, [3, 2], "b")
    mm = T.matmul(a, b2, "mm", transposeB=True)
    assert mm.op == "MatMul" and list(mm.input) == ["a", "b"] and mm.attr["transpose_b"].b and not mm.attr["transpose_a"].b
    cv = T.conv2D(a, b2, 2, 1, 3, 3, -1, -1, TF.TensorflowDataFormat.NHWC, "cv")
    assert list(cv.attr["strides"].list.i) == [1, 1, 2, 1] and cv.attr["padding"].s == b"SAME"
    mp = T.maxPool(a, 2, 3, 0, 0, 2, 2, "NCHW", "mp")
    assert list(mp.attr["ksize"].list.i) == [1, 1, 3, 2] and mp.attr["padding"].s == b"VALID"
    assert T.addN([a, b2], "s").attr["N"].i == 2 and T.squeeze(a, [1], "sq").attr["squeeze_dims"].list.i[0] == 1
    assert T.lrn(a, 2, 1.0, 1e-4, 0.75, "l").attr["depth_radius"].i == 2 and T.reduceMean(a, b2, True, "m").attr["keep_dims"].b
    for fn in (T.relu, T.tanh, T.sigmoid, T.softmax, T.logSoftmax, T.rsqrt, T.floor):
        assert list(fn(a, "n").input) == ["a"]


def test_tfutils(tmp_path):
    from bigdl_b200.interop import tensorflow as TF

    t = P.numpy_to_tensor(np.arange(6, dtype=np.float32).reshape(2, 3))
    assert TF.TFUtils.parseTensor(t).tolist() == [[0, 1, 2], [3, 4, 5]]
    f = tmp_path / "a.bin"
    f.write_bytes(b"xyz")
    out = TF.TFUtils.saveToHDFS([str(f)], str(tmp_path / "out"))
    assert open(out[0], "rb").read() == b"xyz"


def test_named_builtin_savers():
    from bigdl_b200.interop import tensorflow as TF
    from bigdl_b200.nn import layer as L

    ph = TF.Tensorflow.placeholder(np.float32, [2, 4], "x")
    nodes = TF.LinearToTF.toTFDef(L.Linear(4, 3).setName("fc"), [ph])
    assert (nodes[0].op, nodes[0].name) == ("BiasAdd", "fc") and {n.op for n in nodes} == {"BiasAdd", "MatMul", "Const"}
    a, b = TF.Tensorflow.placeholder(np.float32, [2, 4], "a"), TF.Tensorflow.placeholder(np.float32, [2, 4], "b")
    add = TF.CAddTableToTF.toTFDef(L.CAddTable().setName("sum"), [a, b])
    assert len(add) == 1 and list(add[0].input) == ["a", "b"]
    with pytest.raises(TypeError):
        TF.ReLUToTF.toTFDef(L.Tanh(), [ph])
    for name in ("InputToTF ReLUToTF TanhToTF SigmoidToTF SoftMaxToTF LogSoftMaxToTF SpatialConvolutionToTF SqueezeToTF "
                 "ReshapeToTF ViewToTF MaxpoolToTF AvgpoolToTF PaddingToTF DropoutToTF CMultTableToTF JoinTableToTF "
                 "MeanToTF ScaleToTF LRNToTF BatchNorm2DToTF TemporalConvolutionToTF").split():
        assert isinstance(getattr(TF, name), TF.BigDLToTensorflow)


def test_image_decoders_and_parse_single_example():
    import cv2

    from bigdl_b200.interop.tfrecord import encode_example
    from bigdl_b200.nn import tf_grad as G

    img = (np.arange(6 * 5 * 3) % 251).astype(np.uint8).reshape(6, 5, 3)
    bmp = cv2.imencode(".bmp", img)[1].tobytes()
    dec = G.DecodeBmp(3).forward(np.array([bmp], dtype=object))
    assert tuple(dec.shape) == (6, 5, 3) and np.array_equal(dec.numpy().astype(np.uint8), img[..., ::-1])
    jpg = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, 100])[1].tobytes()
    assert tuple(G.DecodeJpeg(3).forward(np.array([jpg], dtype=object)).shape) == (6, 5, 3)
    grey = G.DecodePng(1).forward(np.array([cv2.imencode(".png", img[..., 0])[1].tobytes()], dtype=object))
    assert grey.shape[-1] == 1 or grey.dim() == 2
    assert "DecodeGif" in _LOADERS and isinstance(G.DecodeGif(), G._TFOp)

    ex = encode_example({"x": np.array([1.5, 2.5], np.float32), "y": np.array([7], np.int64)})
    node = _node("ParseSingleExample")
    node.attr["dense_keys"].list.s.extend([b"x", b"missing", b"y"])
    node.attr["Tdense"].list.type.extend([1, 1, 9])
    for dims in ([2], [1], [1]):
        sh = node.attr["dense_shapes"].list.shape.add()
        for d in dims:
            sh.dim.add().size = d
    out = _LOADERS["ParseSingleExample"](node).forward(Table(np.array([ex], dtype=object), torch.zeros(0), torch.zeros(0)))
    vals = out.to_list()
    assert len(vals) == 2 and vals[0].tolist() == [1.5, 2.5] and vals[1].tolist() == [7] and vals[1].dtype == torch.int64
    op = G.ParseSingleExample(["float", torch.int64], [b"x", "y"], [[2], [1]])
    got = op.forward(np.array([ex], dtype=object)).to_list()
    assert got[0].tolist() == [1.5, 2.5] and got[1].tolist() == [7]
    assert G.ParseSingleExample(None, ["y"]).forward(ex)[1].tolist() == [7]          # types follow the record
