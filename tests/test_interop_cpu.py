"""Model interop: Torch7 .t7 (T/utils/TorchFileSpec.scala style round trips)."""
import os
import struct

import numpy as np
import pytest
import torch

from bigdl_b200.interop import torchfile as T7
from bigdl_b200.nn import layer as L
from bigdl_b200.utils.table import Table


def test_t7_primitives_and_tensors(tmp_path):
    p = str(tmp_path / "a.t7")
    t = torch.arange(24, dtype=torch.float32).reshape(2, 3, 4)
    tab = Table()
    tab["x"], tab["flag"], tab["n"], tab["name"], tab[1] = t, True, 3.5, "hello", torch.tensor([1, 2, 3])
    T7.save(tab, p)
    back = T7.load(p)
    assert torch.equal(back["x"], t) and back["flag"] is True and back["n"] == 3.5 and back["name"] == "hello"
    assert back[1].dtype == torch.int64 and back[1].tolist() == [1, 2, 3]
    with pytest.raises(FileExistsError):
        T7.save(tab, p)
    # header layout: TYPE_TABLE, index 1, 5 entries
    assert struct.unpack("<iii", open(p, "rb").read(12)) == (3, 1, 5)
    # a non-contiguous tensor is stored by value
    T7.save(t.transpose(0, 2), p, overWrite=True)
    assert torch.equal(T7.load(p), t.transpose(0, 2))


def test_t7_module_roundtrip(tmp_path):
    torch.manual_seed(0)
    m = (L.Sequential()
         .add(L.SpatialConvolution(3, 8, 3, 3, 1, 1, 1, 1))
         .add(L.ReLU())
         .add(L.SpatialMaxPooling(2, 2, 2, 2))
         .add(L.SpatialCrossMapLRN(5, 1e-4, 0.75, 2.0))
         .add(L.Concat(2).add(L.Sequential().add(L.SpatialConvolution(8, 4, 1, 1)))
              .add(L.Sequential().add(L.SpatialConvolution(8, 4, 3, 3, 1, 1, 1, 1)).add(L.Threshold(0.1, 0.0))))
         .add(L.View(8 * 4 * 4))
         .add(L.Dropout(0.3))
         .add(L.Linear(128, 10))
         .add(L.LogSoftMax()))
    m.evaluate()
    x = torch.randn(2, 3, 8, 8)
    ref = m.forward(x).clone()
    p = str(tmp_path / "m.t7")
    m.saveTorch(p)
    m2 = L.Module.loadTorch(p).evaluate()
    assert [type(c).__name__ for c in m2.modules] == [type(c).__name__ for c in m.modules]
    assert torch.allclose(m2.forward(x), ref, atol=1e-6)
    with pytest.raises(ValueError):
        L.Sequential().add(L.SpatialConvolution(4, 4, 3, 3, n_group=2)).saveTorch(str(tmp_path / "g.t7"))


def test_t7_reads_torch_written_layout(tmp_path):
    """Hand-assemble the bytes Torch7 itself would write for nn.Linear (incl. a shared-storage back reference)."""
    out = bytearray()
    i32 = lambda v: out.extend(struct.pack("<i", v))
    i64 = lambda v: out.extend(struct.pack("<q", v))
    def s(v):
        i32(len(v)); out.extend(v.encode())
    def string(v):
        i32(2); s(v)
    i32(4); i32(1); s("V 1"); s("nn.Linear")
    i32(3); i32(2); i32(3)                       # table, index 2, 3 entries
    string("_type"); string("torch.FloatTensor")
    string("weight")
    i32(4); i32(3); s("V 1"); s("torch.FloatTensor")
    i32(2); i64(2); i64(3); i64(3); i64(1); i64(1)
    i32(4); i32(4); s("V 1"); s("torch.FloatStorage"); i64(8)
    out.extend(np.arange(8, dtype=np.float32).tobytes())
    string("bias")
    i32(4); i32(5); s("V 1"); s("torch.FloatTensor")
    i32(1); i64(2); i64(1); i64(7)               # bias = storage[6:8] of the SAME storage (index 4)
    i32(4); i32(4)
    p = tmp_path / "lin.t7"
    p.write_bytes(bytes(out))
    m = T7.load(str(p))
    assert isinstance(m, L.Linear) and m.weight.tolist() == [[0, 1, 2], [3, 4, 5]] and m.bias.tolist() == [6, 7]


def _branchy_model():
    torch.manual_seed(1)
    inp = L.Input("data")
    c1 = L.SpatialConvolution(3, 8, 3, 3, 1, 1, 1, 1).setName("conv1")(inp)
    bn = L.SpatialBatchNormalization(8).setName("bn1")(c1)
    r1 = L.ReLU().setName("relu1")(bn)
    p1 = L.SpatialMaxPooling(2, 2, 2, 2).setName("pool1")(r1)
    a = L.SpatialConvolution(8, 4, 1, 1).setName("branch_a")(p1)
    b = L.SpatialConvolution(8, 4, 3, 3, 1, 1, 1, 1).setName("branch_b")(p1)
    cat = L.JoinTable(2).setName("concat")([a, b])
    lrn = L.SpatialCrossMapLRN(3, 1e-3, 0.75, 1.0).setName("lrn")(cat)
    add = L.CAddTable().setName("sum")([lrn, p1])
    fl = L.View(8 * 4 * 4).setName("flat")(add)
    fc = L.Linear(128, 5).setName("fc")(fl)
    sm = L.SoftMax().setName("prob")(fc)
    g = L.Model([inp], [sm])
    g.node("bn1").element._buffers["runningMean"].uniform_(-0.5, 0.5)
    g.node("bn1").element._buffers["runningVar"].uniform_(0.5, 1.5)
    return g.evaluate()


def test_caffe_export_import_roundtrip(tmp_path):
    from bigdl_b200.interop.caffe import CaffeLoader

    g = _branchy_model()
    x = torch.randn(2, 3, 8, 8)
    ref = g.forward(x).clone()
    proto, model = str(tmp_path / "net.prototxt"), str(tmp_path / "net.caffemodel")
    g.saveCaffe(proto, model)
    txt = open(proto).read()
    assert 'type: "Convolution"' in txt and 'type: "BatchNorm"' in txt and 'type: "Scale"' in txt and "blobs" not in txt
    g2, crit = CaffeLoader.loadCaffe(proto, model)
    assert crit is None
    out = g2.evaluate().forward(x)
    assert torch.allclose(out, ref, atol=1e-5)
    # copy-by-name into a fresh model of the same structure
    fresh = _branchy_model()
    for m in fresh.all_modules():
        for p in m._parameters.values():
            p.data.zero_()
    L.Module.loadCaffe(fresh, proto, model, matchAll=False)
    assert torch.equal(fresh.node("fc").element.weight, g.node("fc").element.weight)
    extra = L.Sequential().add(L.Linear(3, 3).setName("not_in_caffe"))
    with pytest.raises(ValueError, match="cannot map"):
        L.Module.loadCaffe(extra, proto, model, matchAll=True)


def test_caffe_prototxt_semantics(tmp_path):
    """Hand-written prototxt: V2 syntax, in-place ReLU, ceil-mode pooling, loss layer → criterion."""
    from bigdl_b200.interop import caffe_pb as P
    from bigdl_b200.interop.caffe import CaffeLoader

    proto = tmp_path / "n.prototxt"
    proto.write_text('''
name: "tiny"
input: "data"
input_dim: 1 input_dim: 2 input_dim: 7 input_dim: 7
layer { name: "conv" type: "Convolution" bottom: "data" top: "conv"
        convolution_param { num_output: 3 kernel_size: 3 stride: 2 } }
layer { name: "relu" type: "ReLU" bottom: "conv" top: "conv" }
layer { name: "pool" type: "Pooling" bottom: "conv" top: "pool" pooling_param { pool: MAX kernel_size: 2 stride: 2 } }
layer { name: "ip" type: "InnerProduct" bottom: "pool" top: "ip" inner_product_param { num_output: 4 } }
layer { name: "acc" type: "Accuracy" bottom: "ip" bottom: "label" top: "acc" include { phase: TEST } }
layer { name: "loss" type: "SoftmaxWithLoss" bottom: "ip" bottom: "label" top: "loss" }
''')
    net = P.NetParameter()
    rs = np.random.RandomState(0)
    for name, shapes in (("conv", [(3, 2, 3, 3), (3,)]), ("ip", [(4, 12), (4,)])):
        l = net.layer.add()
        l.name = name
        for s in shapes:
            b = l.blobs.add()
            b.shape.dim.extend(s)
            b.data.extend(rs.randn(*s).astype(np.float32).reshape(-1).tolist())
    model = tmp_path / "n.caffemodel"
    model.write_bytes(net.SerializeToString())
    g, crit = CaffeLoader.loadCaffe(str(proto), str(model))
    assert type(crit).__name__ == "ClassNLLCriterion"
    x = torch.randn(2, 2, 7, 7)
    y = g.evaluate().forward(x)
    # conv 7->3 (stride 2), pool ceil((3-2)/2)+1 = 2 → 3*2*2 = 12 features
    assert tuple(y.shape) == (2, 4) and torch.allclose(y.sum(1), torch.ones(2), atol=1e-5)
    w = torch.tensor(np.asarray(net.layer[0].blobs[0].data, dtype=np.float32).reshape(3, 2, 3, 3))
    b = torch.tensor(np.asarray(net.layer[0].blobs[1].data, dtype=np.float32))
    h = torch.nn.functional.max_pool2d(torch.relu(torch.nn.functional.conv2d(x, w, b, 2)), 2, 2, ceil_mode=True)
    w2 = torch.tensor(np.asarray(net.layer[1].blobs[0].data, dtype=np.float32).reshape(4, 12))
    b2 = torch.tensor(np.asarray(net.layer[1].blobs[1].data, dtype=np.float32))
    assert torch.allclose(y, torch.softmax(h.reshape(2, -1) @ w2.t() + b2, 1), atol=1e-5)


def test_tensorflow_export_import_roundtrip(tmp_path):
    from bigdl_b200.interop.tensorflow import load_tf, save_tf, SUPPORTED_OPS

    assert len(SUPPORTED_OPS) > 120
    torch.manual_seed(2)
    m = (L.Sequential()
         .add(L.SpatialConvolution(3, 6, 3, 3, 1, 1, 1, 1).setName("conv1"))
         .add(L.SpatialBatchNormalization(6).setName("bn"))
         .add(L.ReLU().setName("relu"))
         .add(L.SpatialMaxPooling(2, 2, 2, 2).setName("pool"))
         .add(L.SpatialConvolution(6, 4, 3, 3, 2, 2, -1, -1).setName("conv_same"))
         .add(L.Tanh().setName("tanh"))
         .add(L.Reshape([4 * 2 * 2]).setName("flat"))
         .add(L.Linear(16, 5).setName("fc"))
         .add(L.SoftMax().setName("prob")))
    m.modules[1]._buffers["runningMean"].uniform_(-0.3, 0.3)
    m.modules[1]._buffers["runningVar"].uniform_(0.5, 1.5)
    m.evaluate()
    x = torch.randn(2, 3, 8, 8)
    ref = m.forward(x).clone()
    p = str(tmp_path / "model.pb")
    outs = m.saveTF([("input", [None, 3, 8, 8])], p)
    g = load_tf(p, ["input"], ["prob"])
    y = g.evaluate().forward(x.permute(0, 2, 3, 1).contiguous())         # the exported graph is NHWC
    assert torch.allclose(y, ref, atol=1e-5)
    # weight-slot constants came back as trainable variables → the imported graph can be fine-tuned
    w, gflat = g.getParameters()
    n_params = sum(p_.numel() for _, _, p_ in g.leaf_parameters())
    assert n_params == sum(p_.numel() for p_ in [m.modules[0].weight, m.modules[0].bias, m.modules[4].weight,
                                                 m.modules[4].bias, m.modules[7].weight, m.modules[7].bias]) + 12
    g.training()
    out = g.forward(x.permute(0, 2, 3, 1).contiguous())
    g.backward(x.permute(0, 2, 3, 1).contiguous(), torch.randn_like(out))
    assert float(gflat.abs().sum()) > 0


def test_tensorflow_graphdef_with_control_flow_and_ops():
    """Hand-built GraphDef: out = cond(pred, x*2, x+1) then reduce/concat/strided-slice ops."""
    from bigdl_b200.interop import tf_pb as P
    from bigdl_b200.interop.tensorflow import load_tf

    gd = P.GraphDef()

    def node(op, name, inputs=(), **attrs):
        n = gd.node.add()
        n.op, n.name = op, name
        n.input.extend(inputs)
        for k, v in attrs.items():
            if isinstance(v, np.ndarray):
                n.attr[k].tensor.CopyFrom(P.numpy_to_tensor(v))
            elif isinstance(v, bool):
                n.attr[k].b = v
            elif isinstance(v, int):
                n.attr[k].i = v
        return n

    node("Placeholder", "x")
    node("Placeholder", "pred")
    node("Switch", "sw", ["x", "pred"])
    node("Const", "two", value=np.array(2.0, np.float32))
    node("Const", "one", value=np.array(1.0, np.float32))
    node("Mul", "t", ["sw:1", "two"])
    node("Add", "f", ["sw", "one"])
    node("Merge", "m", ["f", "t"])
    node("Const", "axis", value=np.array([1], np.int32))
    node("Sum", "s", ["m", "axis"], keep_dims=True)
    node("Const", "cat_axis", value=np.array(1, np.int32))
    node("ConcatV2", "cat", ["m", "s", "cat_axis"])
    node("Const", "b", value=np.array([0, 1], np.int32))
    node("Const", "e", value=np.array([0, 3], np.int32))
    node("Const", "st", value=np.array([1, 1], np.int32))
    node("StridedSlice", "out", ["cat", "b", "e", "st"], begin_mask=1, end_mask=1)
    g = load_tf(gd.SerializeToString(), ["x", "pred"], ["out"])
    assert type(g).__name__ == "DynamicGraph"
    x = torch.tensor([[1.0, 2.0], [3.0, 4.0]])
    assert g.forward([x, torch.tensor(True)]).tolist() == [[4.0, 6.0], [8.0, 14.0]]
    assert g.forward([x, torch.tensor(False)]).tolist() == [[3.0, 5.0], [5.0, 9.0]]
    with pytest.raises(ValueError, match="unsupported TensorFlow op"):
        bad = P.GraphDef()
        bad.node.add(op="Placeholder", name="x")
        bad.node.add(op="SomeExoticOp", name="y", input=["x"])
        load_tf(bad.SerializeToString(), ["x"], ["y"])


def test_onnx_import_small_cnn():
    from bigdl_b200.interop import onnx_pb as OP
    from bigdl_b200.interop.onnx_loader import SUPPORTED_OPS, load_model_proto

    assert {"Conv", "Gemm", "BatchNormalization", "MaxPool", "Reshape", "Shape", "Unsqueeze", "Constant"} <= set(SUPPORTED_OPS)
    rs = np.random.RandomState(3)
    w = rs.randn(4, 3, 3, 3).astype(np.float32)
    b = rs.randn(4).astype(np.float32)
    sc, bb, mean, var = [rs.rand(4).astype(np.float32) + 0.5 for _ in range(4)]
    fw = rs.randn(5, 4 * 4 * 4).astype(np.float32)
    fb = rs.randn(5).astype(np.float32)
    m = OP.ModelProto()
    m.ir_version = 7
    m.opset_import.add().version = 11
    g = m.graph
    vi = g.input.add()
    vi.name = "x"
    for name, arr in (("w", w), ("b", b), ("sc", sc), ("bb", bb), ("mean", mean), ("var", var), ("fw", fw), ("fb", fb),
                      ("shape", np.array([0, -1], np.int64))):
        g.initializer.append(OP.from_numpy(arr, name))

    def node(op_type, ins, outs, **attrs):
        n = g.node.add()
        n.op_type = op_type
        n.input.extend(ins)
        n.output.extend(outs)
        for k, v in attrs.items():
            a = n.attribute.add()
            a.name = k
            if isinstance(v, list):
                a.ints.extend(v)
            elif isinstance(v, float):
                a.f = v
            else:
                a.i = v
    node("Conv", ["x", "w", "b"], ["c"], kernel_shape=[3, 3], pads=[1, 1, 1, 1], strides=[1, 1])
    node("BatchNormalization", ["c", "sc", "bb", "mean", "var"], ["bn"], epsilon=1e-5)
    node("Relu", ["bn"], ["r"])
    node("MaxPool", ["r"], ["p"], kernel_shape=[2, 2], strides=[2, 2])
    node("Reshape", ["p", "shape"], ["f"])
    node("Gemm", ["f", "fw", "fb"], ["y"], transB=1, alpha=1.0, beta=1.0)
    node("Softmax", ["y"], ["prob"], axis=1)
    g.output.add().name = "prob"
    model = load_model_proto(m.SerializeToString()).evaluate()
    x = torch.randn(2, 3, 8, 8)
    y = model.forward(x)
    h = F_ref = torch.nn.functional
    ref = h.conv2d(x, torch.tensor(w), torch.tensor(b), 1, 1)
    ref = h.batch_norm(ref, torch.tensor(mean), torch.tensor(var), torch.tensor(sc), torch.tensor(bb), False, 0.0, 1e-5)
    ref = h.max_pool2d(torch.relu(ref), 2, 2).reshape(2, -1) @ torch.tensor(fw).t() + torch.tensor(fb)
    assert torch.allclose(y, torch.softmax(ref, 1), atol=1e-5)
    n_params = sum(p.numel() for _, _, p in model.leaf_parameters())
    assert n_params == w.size + b.size + sc.size + bb.size + fw.size + fb.size      # mean/var stay constants


def test_keras_json_definition_and_weights(tmp_path):
    import json
    from bigdl_b200.interop.keras_converter import DefinitionLoader, WeightLoader, load_keras

    spec = {"class_name": "Sequential", "config": [
        {"class_name": "Convolution2D", "config": {"name": "conv", "nb_filter": 4, "nb_row": 3, "nb_col": 3,
                                                   "activation": "relu", "border_mode": "same", "dim_ordering": "th",
                                                   "batch_input_shape": [None, 1, 8, 8], "subsample": [1, 1]}},
        {"class_name": "MaxPooling2D", "config": {"name": "pool", "pool_size": [2, 2], "dim_ordering": "th"}},
        {"class_name": "Flatten", "config": {"name": "flat"}},
        {"class_name": "Dense", "config": {"name": "fc", "output_dim": 3, "activation": "softmax"}}]}
    p = tmp_path / "m.json"
    p.write_text(json.dumps(spec))
    rs = np.random.RandomState(0)
    wc, bc = rs.randn(4, 1, 3, 3).astype(np.float32), rs.randn(4).astype(np.float32)
    wd, bd = rs.randn(64, 3).astype(np.float32), rs.randn(3).astype(np.float32)
    np.savez(tmp_path / "w.npz", **{"conv/0": wc, "conv/1": bc, "fc/0": wd, "fc/1": bd})
    model = load_keras(str(p), str(tmp_path / "w.npz"))
    assert model.getOutputShape() == (None, 3)
    x = torch.randn(2, 1, 8, 8)
    y = model.forward(x)
    h = torch.nn.functional
    ref = h.max_pool2d(torch.relu(h.conv2d(x, torch.tensor(wc), torch.tensor(bc), 1, 1)), 2).reshape(2, -1)
    ref = torch.softmax(ref @ torch.tensor(wd) + torch.tensor(bd), 1)
    assert torch.allclose(y, ref, atol=1e-5)
    with pytest.raises(ValueError, match="Unsupported keras layer"):
        DefinitionLoader.from_config({"class_name": "Sequential", "config": [{"class_name": "Lambda", "config": {}}]})
    with pytest.raises(ImportError):
        WeightLoader.load_weights_from_hdf5(model, "x.h5")
    func = {"class_name": "Model", "config": {
        "layers": [{"class_name": "InputLayer", "name": "in", "config": {"batch_input_shape": [None, 6], "name": "in"}, "inbound_nodes": []},
                   {"class_name": "Dense", "name": "a", "config": {"name": "a", "output_dim": 4}, "inbound_nodes": [[["in", 0, 0]]]},
                   {"class_name": "Dense", "name": "b", "config": {"name": "b", "output_dim": 4}, "inbound_nodes": [[["in", 0, 0]]]},
                   {"class_name": "Merge", "name": "m", "config": {"name": "m", "mode": "sum"}, "inbound_nodes": [[["a", 0, 0], ["b", 0, 0]]]}],
        "input_layers": [["in", 0, 0]], "output_layers": [["m", 0, 0]]}}
    fm = DefinitionLoader.from_config(func)
    assert tuple(fm.forward(torch.randn(5, 6)).shape) == (5, 4)


def test_tfrecord_and_session(tmp_path):
    from bigdl_b200.dataset.base import DataSet, Sample
    from bigdl_b200.interop import tf_pb as P
    from bigdl_b200.interop.tf_session import Session
    from bigdl_b200.interop.tfrecord import FixedLengthRecordReader, TFRecordIterator, TFRecordWriter, encode_example
    from bigdl_b200.nn.criterion import MSECriterion
    from bigdl_b200.nn.tf import parse_example
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.optim.trigger import Trigger

    path = str(tmp_path / "data.tfrecord")
    with TFRecordWriter(path) as w:
        for i in range(3):
            w.write(encode_example({"x": [float(i), 2.0], "id": [i], "name": f"rec{i}"}))
    recs = [parse_example(r) for r in TFRecordIterator(path)]
    assert recs[2]["x"] == ("float", [2.0, 2.0]) and recs[1]["id"] == ("int64", [1]) and recs[0]["name"] == ("bytes", [b"rec0"])
    raw = tmp_path / "fixed.bin"
    raw.write_bytes(b"HH" + b"".join(bytes([i]) * 3 + b"-" for i in range(4)) + b"F")
    assert [r for r in FixedLengthRecordReader(str(raw), 1, 2, 1, 3)] == [bytes([i]) * 3 for i in range(4)]

    # y = x @ W + b as a TF graph with trainable constants; fit it to a known linear map through Session.train
    gd = P.GraphDef()
    def node(op, name, inputs=(), **attrs):
        n = gd.node.add()
        n.op, n.name = op, name
        n.input.extend(inputs)
        for k, v in attrs.items():
            n.attr[k].tensor.CopyFrom(P.numpy_to_tensor(v))
    node("Placeholder", "x")
    node("Const", "W", value=np.zeros((3, 2), np.float32))
    node("Const", "b", value=np.zeros(2, np.float32))
    node("MatMul", "mm", ["x", "W"])
    node("BiasAdd", "y", ["mm", "b"])
    rs = np.random.RandomState(0)
    X = rs.randn(64, 3).astype(np.float32)
    Wt = np.array([[1.0, -1.0], [0.5, 2.0], [-2.0, 0.0]], np.float32)
    Y = X @ Wt + np.array([0.5, -0.5], np.float32)
    ds = DataSet.array([Sample(X[i], Y[i]) for i in range(64)])
    sess = Session(gd)
    assert sess.placeholders() == ["x"]
    model = sess.train(["y"], ds, SGD(0.1), MSECriterion(), Trigger.maxEpoch(40), batch_size=16)
    pred = np.stack([np.asarray(p) for p in sess.predict(["y"], ds, batch_size=16)])
    assert np.abs(pred - Y).max() < 0.05
    out = sess.run(["y"], {"x": X[:2]})
    assert np.allclose(out.detach().numpy(), pred[:2], atol=1e-5)
    saved = sess.saveParameters(str(tmp_path / "vars.npz"))
    assert np.abs(np.load(saved)["W"] - Wt).max() < 0.05


def test_caffe_export_covers_the_activation_and_affine_layers(tmp_path):
    """saveCaffe → loadCaffeModel for the rest of the reference's toCaffe table (Converter.scala:297-329): ELU, PReLU,
    Deconvolution, Scale, Bias, AbsVal, Power, within-channel LRN, Threshold, Exp; weights above 256 elements take the
    raw packed-record path."""
    import torch
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.module import Module

    torch.manual_seed(0)
    m = (L.Sequential().add(L.SpatialConvolution(3, 4, 3, 3, 1, 1, 1, 1)).add(L.ELU(0.7))
         .add(L.SpatialFullConvolution(4, 6, 3, 3, 2, 2, 1, 1)).add(L.PReLU(6)).add(L.Scale([1, 6, 1, 1])).add(L.Abs())
         .add(L.Power(2.0, 0.5, 0.1)).add(L.SpatialWithinChannelLRN(3, 0.5, 0.75)).add(L.CAdd([1, 6, 1, 1]))
         .add(L.BinaryThreshold(0.2)).add(L.Exp()))
    for p in m.parameters()[0]:
        if p.abs().sum() == 0:
            p.data.normal_()
    m.evaluate()
    proto, weights = str(tmp_path / "n.prototxt"), str(tmp_path / "n.caffemodel")
    m.saveCaffe(proto, weights)
    back = Module.loadCaffeModel(proto, weights).evaluate()
    x = torch.randn(2, 3, 6, 6)
    assert torch.allclose(m.forward(x), back.forward(x), atol=1e-6)


def test_tf_export_keeps_default_module_names(tmp_path):
    """A model whose layers were never named exports under getName(), so the caller can ask loadTF for
    ``model.modules[-1].getName()`` as the reference's users do."""
    import torch
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.module import Module

    torch.manual_seed(0)
    m = L.Sequential().add(L.Linear(8, 6)).add(L.Tanh()).add(L.Linear(6, 3))
    p = str(tmp_path / "m.pb")
    m.saveTF([("input", [2, 8])], p)
    back = Module.loadTF(p, ["input"], [m.modules[-1].getName()])
    x = torch.randn(2, 8)
    assert torch.allclose(m.forward(x), back.forward(x), atol=1e-5)


def test_tf_export_sequence_and_affine_layers(tmp_path):
    """TemporalConvolution / Padding / Mean / Squeeze / Scale of the reference's saver table (TensorflowSaver.scala:
    171-194) survive saveTF → loadTF."""
    import torch
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.module import Module

    torch.manual_seed(0)
    m = (L.Sequential().add(L.TemporalConvolution(4, 5, 3, 1)).add(L.Tanh()).add(L.Padding(1, 2, 2, 0.5))
         .add(L.Padding(2, -1, 2)).add(L.Mean(2, squeeze=False)).add(L.Squeeze(2)))
    x = torch.randn(3, 9, 4)
    p = str(tmp_path / "seq.pb")
    m.saveTF([("input", [3, 9, 4])], p)
    back = Module.loadTF(p, ["input"], [m.modules[-1].getName()])
    assert torch.allclose(m.forward(x), back.forward(x), atol=1e-5)

    m2 = L.Sequential().add(L.SpatialConvolution(3, 4, 3, 3)).add(L.Scale([1, 4, 1, 1])).add(L.ReLU())
    for q in m2.modules[1].parameters()[0]:
        q.data.normal_()
    x = torch.randn(2, 3, 6, 6)
    p = str(tmp_path / "img.pb")
    m2.saveTF([("input", [2, 3, 6, 6])], p)
    back = Module.loadTF(p, ["input"], [m2.modules[-1].getName()])
    out = back.forward(x.permute(0, 2, 3, 1).contiguous()).permute(0, 3, 1, 2)
    assert torch.allclose(m2.forward(x), out, atol=1e-5)
