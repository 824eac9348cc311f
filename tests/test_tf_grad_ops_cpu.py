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
