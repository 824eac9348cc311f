"""mkldnn-named API and the IR conversion utilities (T/nn/mkldnn/*Spec.scala, T/utils/intermediate/*Spec.scala)."""
import torch

from bigdl_b200.nn import layer as L
from bigdl_b200.nn import mkldnn as D
from bigdl_b200.utils.intermediate import BlasToIR, ConversionUtils, IRToBlas, IRToDnn


def test_dnn_sequential_compile_inference_folds_bn():
    torch.manual_seed(0)
    m = (D.Sequential().add(D.Input([2, 3, 8, 8])).add(D.SpatialConvolution(3, 8, 3, 3, 1, 1, 1, 1))
         .add(D.SpatialBatchNormalization(8)).add(D.ReLU()).add(D.MaxPooling(2, 2, 2, 2)).add(D.Output()))
    m.modules[2]._buffers["runningMean"].uniform_(-0.2, 0.2)
    m.modules[2]._buffers["runningVar"].uniform_(0.5, 1.5)
    x = torch.randn(2, 3, 8, 8)
    ref = m.evaluate().forward(x).clone()
    m.compile(D.Phase.InferencePhase, device="cpu")
    assert getattr(m.modules[2], "_folded", False)
    assert torch.allclose(m.forward(x), ref, atol=1e-5)


def test_reorder_memory_layout_and_dtype():
    r = D.ReorderMemory(D.HeapData([2, 4, 3, 3], D.Memory.Format.nchw), D.NativeData([2, 4, 3, 3], D.Memory.Format.nhwc,
                                                                                     D.Memory.DataType.BF16))
    y = r.forward(torch.randn(2, 4, 3, 3))
    assert y.dtype == torch.bfloat16 and y.is_contiguous(memory_format=torch.channels_last)


def test_dnn_rnn_shapes():
    assert tuple(D.RNN("vanilla_lstm", 6, 5, layers=2).forward(torch.randn(3, 4, 6)).shape) == (3, 4, 5)
    assert tuple(D.RNN("vanilla_gru", 6, 5, direction="bidirectional_concat").forward(torch.randn(3, 4, 6)).shape) == (3, 4, 10)


def test_ir_roundtrip_sequential_and_graph():
    torch.manual_seed(1)
    seq = (L.Sequential().add(L.Linear(6, 8)).add(L.ReLU())
           .add(L.Sequential().add(L.Linear(8, 8)).add(L.Tanh())).add(L.Linear(8, 3)))
    ir = BlasToIR.convert(seq)
    assert [n.element.op for n in ir.nodes] == ["Linear", "ReLU", "Linear", "Tanh", "Linear"]      # nested chain flattened
    assert "Linear" in ir.summary()
    x = torch.randn(4, 6)
    ref = seq.forward(x).clone()
    g = IRToBlas.convert(ir)
    assert torch.allclose(g.forward(x), ref)
    dnn = IRToDnn.convert(ir)
    assert isinstance(dnn, D.DnnGraph) and torch.allclose(dnn.compile(D.Phase.InferencePhase, device="cpu").forward(x), ref)
    # weights are shared with the source model, not copied
    assert g.exec_order[1].element is seq.modules[0]
    # graph with a branch
    inp = L.Input()
    a = L.Linear(4, 4)(inp)
    b = L.Linear(4, 4)(inp)
    out = L.CAddTable()([a, b])
    model = L.Model([inp], [out])
    ir2 = BlasToIR.convert(model)
    assert len(ir2) == 4 and ir2.nodes[3].prev == [1, 2]
    xx = torch.randn(2, 4)
    assert torch.allclose(IRToBlas.convert(ir2).forward(xx), model.forward(xx))


def test_conversion_utils_cpu_is_identity():
    m = L.Sequential().add(L.Linear(3, 2))
    assert ConversionUtils.convert(m, device="cpu") is m
    assert ConversionUtils.convertBack(m) is m


def test_to_graph_shares_layers_and_matches():
    torch.manual_seed(2)
    m = (L.Sequential().add(L.Linear(4, 6)).add(L.ReLU())
         .add(L.ConcatTable(L.Linear(6, 3), L.Sequential().add(L.Linear(6, 3)).add(L.Tanh()))).add(L.CAddTable())
         .add(L.Concat(2).add(L.Linear(3, 2)).add(L.Linear(3, 5))))
    x = torch.randn(3, 4)
    ref = m.forward(x).clone()
    g = m.toGraph()
    assert type(g).__name__ == "Graph" and torch.allclose(g.forward(x), ref)
    assert any(n.element is m.modules[0] for n in g.exec_order)
    assert g.toGraph() is g


def test_reorder_manager_tensor_mmap_and_quantization_rule():
    import torch
    from bigdl_b200.nn import mkldnn as D
    from bigdl_b200.quantized import Desc, QuantSerializer, Quantization
    from bigdl_b200.nn.layer import SpatialConvolution

    rm = D.ReorderManager()
    a = D.HeapData([2, 4, 3, 3], D.Memory.Format.nchw)
    b = D.NativeData([2, 4, 3, 3], D.Memory.Format.nhwc)
    assert rm.needReorder(a, b) and not rm.needReorder(a, a.cloneFormat())
    rm.register(a, b)
    x = torch.randn(2, 4, 3, 3)
    y = rm.infer(a, b, x)
    assert torch.equal(x, y) and y.is_contiguous(memory_format=torch.channels_last)
    assert rm.infer(b, a, x) is x           # not registered: passes through

    mm = D.TensorMMap([3, 2], device="cpu")
    mm.dense.copy_(torch.arange(6.).reshape(3, 2))
    assert torch.equal(mm.sync().native(), mm.dense)
    mm.native().mul_(2)
    assert torch.equal(mm.syncToHeap().dense, 2 * torch.arange(6.).reshape(3, 2))

    assert Quantization.quantize(0.5, 1.0, -2.0) == 32 and abs(Quantization.dequantize(32, 1.0, -2.0) - 0.5039) < 1e-3
    w = torch.randn(6, 4, 3, 3)
    q, mx, mn = Quantization.quantize_tensor(w)
    assert q.dtype == torch.int8 and int(q.abs().max()) == 127
    assert (Quantization.dequantize_tensor(q, mx, mn) - w).abs().max() <= float(w.abs().max()) / 127 * 0.51
    assert Quantization.loss(w) < 0.02
    d = Desc.of(SpatialConvolution(4, 6, 3, 3))
    assert (d.nOutputPlane, d.nInputPlane, d.kernelH) == (6, 4, 3)


def test_typed_ir_operators_and_converter():
    import torch

    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.mkldnn import DnnGraph
    from bigdl_b200.utils import config
    from bigdl_b200.utils import intermediate as IR

    m = L.Sequential().add(L.SpatialConvolution(3, 4, 3, 3, 1, 1, 1, 1)).add(L.ReLU()).add(L.SpatialMaxPooling(2, 2, 2, 2)) \
        .add(L.View(4 * 4 * 4)).add(L.Linear(64, 5)).add(L.SoftMax())
    ir = IR.BlasToIR.convert(m)
    ops = [n.element.operator() for n in ir.nodes]
    assert isinstance(ops[0], IR.IRSpatialConvolution) and ops[0].n_output_plane == 4 and ops[0].kernel_w == 3
    assert isinstance(ops[1], IR.IRReLU) and isinstance(ops[2], IR.IRSpatialMaxPooling)
    assert isinstance(ops[3], IR.IRGeneralModule) and isinstance(ops[3].model, L.View)       # no dedicated operator
    assert isinstance(ops[4], IR.IRLinear) and isinstance(ops[5], IR.IRSoftMax) and ops[5].name() == "IRSoftMax"
    assert ops[1] == IR.IRReLU(**ops[1].attrs()) and ops[1] != ops[5] and "IRLinear(" in repr(ops[4])
    assert IR.ConvertBase.convertLayerCheck(m.modules[0]) and not IR.ConvertBase.convertLayerCheck(m.modules[3])
    el = IR.ConvertBase.convertLayer(m.modules[4])
    assert el.op == "Linear" and IR.ConvertBase.convertLayer(el) is m.modules[4]

    x = torch.rand(2, 3, 8, 8)
    want = m.forward(x)
    old = config.get("bigdl.engineType")
    try:
        config.set("bigdl.engineType", "mkldnn")
        g = IR.IRConverter(ir).toGraph()
        assert isinstance(g, DnnGraph) and torch.allclose(g.forward(x), want, atol=1e-6)
        config.set("bigdl.engineType", "mklblas")
        g2 = IR.IRConverter(ir).toGraph()
        assert not isinstance(g2, DnnGraph) and torch.allclose(g2.forward(x), want, atol=1e-6)
    finally:
        config.set("bigdl.engineType", old)
