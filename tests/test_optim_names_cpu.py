"""LineSearch / WolfeLineSearch, PRAUCResult, EvaluateMethods, MAPUtil and the ground-truth region classes
(S/optim/{LineSearch,PrecisionRecallAUC,EvaluateMethods,ValidationMethod,AbstractOptimizer}.scala)."""
import numpy as np
import pytest
import torch

from bigdl_b200.optim import optimizer as O
from bigdl_b200.optim import validation as Vd
from bigdl_b200.optim.optim_method import LBFGS, LineSearch, WolfeLineSearch


def _rosenbrock(x):
    x = x.detach().clone().requires_grad_(True)
    f = ((1 - x[:-1]) ** 2).sum() + 100 * ((x[1:] - x[:-1] ** 2) ** 2).sum()
    f.backward()
    return float(f.detach()), x.grad


def test_wolfe_line_search_conditions_and_lbfgs():
    x = torch.tensor([-1.2, 1.0, 0.5], dtype=torch.float64)
    f, g = _rosenbrock(x)
    d = -g
    gtd = float(g.dot(d))
    ls = WolfeLineSearch()
    assert isinstance(ls, LineSearch)
    f1, g1, x1, t, evals = ls(_rosenbrock, x, 1.0, d, f, g, gtd)
    assert evals >= 1 and t > 0 and torch.allclose(x1, x + t * d)
    assert f1 <= f + 1e-4 * t * gtd                     # sufficient decrease
    assert abs(float(g1.dot(d))) <= 0.9 * abs(gtd)      # curvature
    assert f1 == pytest.approx(_rosenbrock(x1)[0])

    runs = {}
    for name, search in (("fixed", None), ("wolfe", WolfeLineSearch())):
        x = torch.zeros(4, dtype=torch.float64)
        opt = LBFGS(max_iter=200, max_eval=4000, linesearch=search, tolfun=1e-10, tolx=1e-12)
        x, hist = opt.optimize(_rosenbrock, x)
        assert torch.allclose(x, torch.ones(4, dtype=torch.float64), atol=1e-4)
        runs[name] = len(hist)
    assert runs["wolfe"] < runs["fixed"]
    with pytest.raises(NotImplementedError):
        LineSearch()(None, None, 1, None, 0, None, 0)


def test_prauc_result_matches_reference_loop():
    scores = [0.9, 0.8, 0.7, 0.6, 0.55, 0.4, 0.3]
    labels = [1, 0, 1, 1, 0, 0, 1]
    res = Vd.PRAUCResult(list(zip(scores, labels)))
    # the reference's running-sum form
    tp = fp = 0.0
    area, prev_p, prev_r = 0.0, 1.0, 0.0
    total = sum(labels)
    for s, l in sorted(zip(scores, labels), reverse=True):
        if tp == total:
            break
        tp, fp = tp + (l == 1), fp + (l != 1)
        p, r = tp / (tp + fp), tp / total
        area += (r - prev_r) * (p + prev_p)
        prev_p, prev_r = p, r
    assert res.result() == (pytest.approx(area / 2), 7)
    both = res + Vd.PRAUCResult([0.95], [1.0])
    assert both.result()[1] == 8 and "Precision Recall AUC" in repr(both)
    m = Vd.PrecisionRecallAUC()(torch.tensor(scores), torch.tensor(labels))
    assert isinstance(m, Vd.PRAUCResult) and m.result()[0] == pytest.approx(area / 2)


def test_evaluate_methods():
    out = torch.tensor([[0.1, 0.7, 0.2], [0.5, 0.2, 0.3], [0.2, 0.3, 0.5]])
    assert Vd.EvaluateMethods.calcAccuracy(out, torch.tensor([2.0, 1.0, 1.0])) == (2, 3)
    assert Vd.EvaluateMethods.calcAccuracy(out[0], torch.tensor([2.0])) == (1, 1)
    big = torch.arange(24, dtype=torch.float32).reshape(3, 8)
    assert Vd.EvaluateMethods.calcTop5Accuracy(big, torch.tensor([8.0, 3.0, 4.0])) == (2, 3)
    assert Vd.EvaluateMethods.calcTop5Accuracy(big[0], torch.tensor([1.0])) == (0, 1)
    with pytest.raises(ValueError):
        Vd.EvaluateMethods.calcAccuracy(torch.zeros(2, 2, 2), torch.zeros(2))


def test_map_util_and_regions():
    assert Vd.MAPUtil.findTopK(2, [[0, 0.3], [0, 0.9], [0, 0.5]], 1) == [(1, pytest.approx(0.9)), (2, 0.5)]
    gt = [{"bboxes": torch.tensor([[0.0, 0, 9, 9], [20.0, 20, 29, 29]]), "classes": torch.tensor([[1.0, 1.0], [0, 1]]),
           "is_crowd": torch.zeros(2)}]
    images, cnt = Vd.MAPUtil.gtTablesToGroundTruthRegions(gt, 3, 2, False, False)
    assert cnt == [0, 1, 0] and isinstance(images[0][0], Vd.GroundTruthBBox) and images[0][1].diff == 1.0
    assert (images[0][0].xmin, images[0][0].ymax) == (0.0, 9.0)
    predict = [[[] for _ in range(3)] for _ in range(2)]
    Vd.MAPUtil.parseDetection(images[0], 1, 0.9, 0, 0, 9, 9, None, 3, [0.5, 0.75], predict)
    Vd.MAPUtil.parseDetection(images[0], 1, 0.8, 0, 0, 9, 9, None, 3, [0.5, 0.75], predict)   # GT taken → miss
    Vd.MAPUtil.parseDetection(images[0], 1, 0.7, 20, 20, 29, 29, None, 3, [0.5, 0.75], predict)   # difficult → ignored
    assert predict[0][1] == [(0.9, True), (0.8, False)] and predict[1][1] == [(0.9, True), (0.8, False)]
    g = Vd.GroundTruthBBox(False, 1, 0, 0.0, 0, 0, 9, 9)
    assert g.getIOURate(0, 0, 9, 4) == pytest.approx(0.5) and g.canOccupy(0)
    g.occupy(0)
    assert not g.canOccupy(0)
    crowd = Vd.GroundTruthBBox(True, 1, 0, 1.0, 0, 0, 9, 9)
    crowd.occupy(0)
    assert crowd.canOccupy(0) and crowd.getIOURate(0, 0, 4, 4) == pytest.approx(1.0)   # intersection over detection
    seen = []
    out = torch.tensor([[2, 1, 0.9, 1, 2, 3, 4, 2, 0.8, 5, 6, 7, 8], [0] + [0] * 12], dtype=torch.float32)
    Vd.MAPUtil.parseSegmentationTensorResult(out, lambda *a: seen.append(a))
    assert seen == [(0, 1, pytest.approx(0.9), 1, 2, 3, 4), (0, 2, pytest.approx(0.8), 5, 6, 7, 8)]
    assert isinstance(Vd.GroundTruthRLE(1, 0, 0.0, object()), Vd.GroundTruthRegion)
    assert issubclass(O.LocalOptimizer, O.AbstractOptimizer) and issubclass(O.DistriOptimizer, O.AbstractOptimizer)


def test_future_result_and_sumsquare():
    from concurrent.futures import ThreadPoolExecutor

    from bigdl_b200.parallel.all_reduce_parameter import FutureResult, Util

    with ThreadPoolExecutor(2) as ex:
        got = FutureResult([ex.submit(lambda i=i: i * i) for i in range(4)] + [7]).waitResult()
    assert got == [0, 1, 4, 9, 7]
    x = torch.randn(1001)
    assert Util.getSumsquareInParallel(x, 4) == pytest.approx(float((x.double() ** 2).sum()))


def test_model_broadcast_factory_and_cache():
    import pickle

    from bigdl_b200.nn.layer import Linear
    from bigdl_b200.parallel import model_broadcast as MB
    from bigdl_b200.utils import config

    b = MB.ModelBroadcast()
    assert isinstance(b, MB.ModelBroadcastImp) and not b.applyProtoBuffer and len(b.uuid()) == 36
    config.set("bigdl.ModelBroadcastFactory", "protobuf")
    try:
        assert MB.ModelBroadcast().applyProtoBuffer
        config.set("bigdl.ModelBroadcastFactory", "bigdl_b200.parallel.model_broadcast.DefaultModelBroadcastFactory")
        assert not MB.ModelBroadcast().applyProtoBuffer
    finally:
        config.set("bigdl.ModelBroadcastFactory", "default")
    m = Linear(3, 2)
    m.forward(torch.rand(4, 3))
    b.broadcast(m)
    assert b.value() is m and b.value(shareWeight=False) is not m
    assert MB.CachedModels.get(b.uuid()) == [m]
    info = pickle.loads(pickle.dumps(MB.ModelInfo(b.uuid(), m)))
    assert info.model is m
    MB.CachedModels.add("old", Linear(2, 2))
    MB.CachedModels.deleteAll(b.uuid())
    assert MB.CachedModels.get("old") == [] and MB.CachedModels.get(b.uuid()) == [m]
    MB.CachedModels.deleteKey(b.uuid())
    assert MB.CachedModels.get(b.uuid()) == [] and m.output is None      # released


def test_abstractnn_small_types():
    from bigdl_b200.nn import abstractnn as A
    from bigdl_b200.nn import layer as L
    from bigdl_b200.utils.table import T

    assert A.DataFormat("nhwc") is A.DataFormat.NHWC and A.DataFormat.NCHW == "NCHW" and str(A.DataFormat.NHWC) == "NHWC"
    assert A.DataFormat.NCHW.getHWCDims(4) == (3, 4, 2) and A.DataFormat.NHWC.getHWCDims(3) == (1, 2, 3)
    with pytest.raises(ValueError):
        A.DataFormat("chwn")
    t = torch.zeros(2)
    assert A.Activity.isTensor(t) and not A.Activity.isTable(t) and A.Activity.isTable(T(t)) and A.Activity.toTensor(t) is t
    assert A.Activity.isTable(A.Activity.allocate("table")) and A.Activity.allocate("tensor", torch.int32).dtype == torch.int32
    with pytest.raises(TypeError):
        A.Activity.toTable(t)
    with pytest.raises(RuntimeError, match="gradInput of conv1 is empty"):
        A.EmptyGradInput("conv1").toTensor()

    class Same(A.IdentityOutputShape):
        def isKerasStyle(self):
            return True

    s = Same()
    with pytest.raises(RuntimeError):
        s.getOutputShape()
    assert s.build((None, 4)) == (None, 4) and s.getOutputShape() == (None, 4) and s.getInputShape() == (None, 4)
    with pytest.raises(A.InvalidLayer):
        s.validateInput([L.Linear(2, 2)])
    with pytest.raises(RuntimeError):
        A.InferShape().getInputShape()

    class Mine(A.Initializable):
        resets = 0

        def reset(self):
            self.resets += 1

    m = Mine().setInitMethod("w", "b")
    assert (m.weightInitMethod, m.biasInitMethod, m.resets) == ("w", "b", 1)
    assert getattr(A.SizeAverageStatus, "True") == "True" and A.AbstractModule is L.Module if hasattr(L, "Module") else True


def test_transformer_building_blocks_and_tensor_tree():
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.criterion import NormMode
    from bigdl_b200.nn.layers_attention import position_signal

    x = torch.randn(2, 5, 8)
    pe = L.PositionEncode().forward(x)
    assert tuple(pe.shape) == (5, 8) and torch.allclose(pe, position_signal(5, 8, x.device))
    sh = L.PositionEncodeWithShift().forward(x)
    assert torch.allclose(sh[:, 0], pe[0].expand(2, 8)) and torch.allclose(sh[:, 1:], x[:, :-1] + pe[1:])
    ids = torch.tensor([[3.0, 4.0, 0.0], [5.0, 0.0, 0.0]])
    pm = L.PaddingMask().forward(ids)
    assert tuple(pm.shape) == (2, 1, 1, 3) and pm[0, 0, 0].tolist() == [0.0, 0.0, -1e9]
    sm = L.SelfAttentionMask().forward(x)
    assert tuple(sm.shape) == (1, 1, 5, 5) and float(sm[0, 0, 0, 1]) == -1e9 and float(sm[0, 0, 3, 2]) == 0.0
    heads = L.SplitHeads(8, 2, True).forward(x)
    assert tuple(heads.shape) == (2, 2, 5, 4) and torch.allclose(heads[0, 1, 3], x[0, 3, 4:] * 0.5)
    assert torch.allclose(L.CombineHeads().forward(L.SplitHeads(8, 2).forward(x)), x)
    parts = L.SplitTensor(3, 2).forward(x)
    assert len(parts) == 2 and torch.equal(parts[2], x[..., 4:])
    vp = L.VectorProduct(8)
    assert torch.allclose(vp.forward(x), x)
    vp.weight.data.fill_(2.0)
    vp.bias.data.fill_(1.0)
    y = vp.forward(x)
    assert torch.allclose(y, 2 * x + 1)
    vp.backward(x, torch.ones_like(x))
    assert torch.allclose(vp.parameters()[1][0], x.sum((0, 1)), atol=1e-5)
    assert L.LanguageModel == L.TransformerType.LanguageModel and NormMode.BATCH_SIZE == "BATCH_SIZE"

    t = L.TensorTree(torch.zeros(3, 3))
    t.addChild(1, 2)
    t.addChild(1, 3)
    t.markAsRoot(1)
    t.markAsLeaf(2, 1)
    t.markAsLeaf(3, 2)
    assert t.children(1) == [2, 3, -1] and t.getRoot() == 1 and t.hasChild(1) and t.noChild(2) and t.leafIndex(3) == 2
    assert t.exists(3) and not t.exists(4) and not t.isPadding(1) and t.nodeNumber == 3 and t.size == [3, 3]
    with pytest.raises(ValueError):
        L.TensorTree(torch.zeros(3))


def test_keras_wrappers_and_utils():
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.keras import layers as K
    from bigdl_b200.nn.keras.topology import Sequential as KSequential

    m = KSequential()
    m.add(K.Dense(6, input_shape=(4,)))
    m.add(K.KerasLayerWrapper(L.Sequential().add(L.Linear(6, 3)).add(L.ReLU())))
    m.add(K.KerasIdentityWrapper(L.Tanh()))
    assert m.getOutputShape() == (None, 3) and tuple(m.forward(torch.rand(5, 4)).shape) == (5, 3)
    with pytest.raises(ValueError):
        K.KerasLayerWrapper(K.Dense(2))
    assert K.KerasUtils.getPadsFromBorderMode("same") == (-1, -1) and K.KerasUtils.toBigDLFormat("tf") == "NHWC"
    assert K.KerasUtils.toBigDLFormat5D("th") == "CHANNEL_FIRST" and K.KerasUtils.addBatch((3, 4)) == (None, 3, 4)
    assert K.KerasUtils.removeBatch((None, 3)) == (3,) and isinstance(K.KerasUtils.getActivation("relu"), L.ReLU)
    assert K.KerasUtils.getInitMethod("glorot_uniform") is not None
    with pytest.raises(ValueError):
        K.KerasUtils.toBigDLFormat("xx")
    assert issubclass(K.MaxPooling2D, K.Pooling2D) and issubclass(K.GlobalMaxPooling1D, K.GlobalPooling1D)


def test_recurrent_batch_norm_params(tmp_path):
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.layers_recurrent import BatchNormParams
    from bigdl_b200.nn.module import Module

    torch.manual_seed(0)
    rec = L.Recurrent(BatchNormParams(eps=1e-3, momentum=0.2)).add(L.LSTM(5, 4))
    assert len(rec.modules) == 2 and rec.bnorm.modules[0].n_output == 16
    x = torch.randn(3, 6, 5)
    y = rec.forward(x)
    assert tuple(y.shape) == (3, 6, 4)
    # the normalised projection: statistics are taken over batch AND time
    cell, bn = rec.cell, rec.bnorm.modules[0]
    proj = cell.i2g.forward(x.reshape(18, 5))
    mu, var = proj.mean(0), proj.var(0, unbiased=False)
    want = (proj - mu) / torch.sqrt(var + 1e-3) * bn.weight.detach() + bn.bias.detach()
    h = c = torch.zeros(3, 4)
    outs = []
    for t in range(6):
        g = want.reshape(3, 6, 16)[:, t] + h @ cell.h2g.weight.detach().t()
        i, cand, f, o = g.chunk(4, 1)
        c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(cand)
        h = torch.sigmoid(o) * torch.tanh(c)
        outs.append(h)
    assert torch.allclose(y, torch.stack(outs, 1), atol=1e-5)
    g = rec.backward(x, torch.ones_like(y))
    assert tuple(g.shape) == (3, 6, 5) and float(bn.parameters()[1][0].abs().sum()) > 0
    # 2 LSTM weights + bias, BN gain + shift
    assert len(rec.parameters()[0]) == 5

    p = str(tmp_path / "rec.bigdl")
    rec.evaluate()
    y_eval = rec.forward(x)
    rec.saveModule(p, overWrite=True)
    back = Module.loadModule(p)
    assert back.bnorm is not None and back.batch_norm_params.eps == pytest.approx(1e-3)
    assert torch.allclose(back.evaluate().forward(x), y_eval, atol=1e-6)
    from bigdl_b200.nn.layers_recurrent import Cell

    class Bare(Cell):                       # no input projection → nothing to normalise
        input_size = 5

    with pytest.raises(ValueError, match="does not support BatchNormalization"):
        L.Recurrent(True).add(Bare())
    assert L.BiRecurrent(batch_norm_params=True).add(L.GRU(5, 4)).forward(x).shape[-1] == 4


def test_detection_output_param(tmp_path):
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.module import Module

    prm = L.DetectionOutputParam(nClasses=5, keepTopK=7, confThresh=0.2)
    a, b = L.DetectionOutputSSD(prm), L.DetectionOutputSSD.apply(prm, False)
    assert (a.n_classes, a.keep_top_k, a.conf_thresh, a.conf_post_process, b.conf_post_process) == (5, 7, 0.2, True, False)
    f = str(tmp_path / "d.bigdl")
    a.saveModule(f, overWrite=True)
    back = Module.loadModule(f)
    assert (back.n_classes, back.keep_top_k) == (5, 7)


def test_per_kind_quantizers():
    from bigdl_b200.nn import layer as L
    from bigdl_b200.quantized import layers as QL
    from bigdl_b200.quantized.quantizer import (CellQuantizer, ContainerQuantizer, GraphQuantizer, ModuleQuantizer,
                                                Quantizable, Quantizer)

    relu = L.ReLU()
    assert ModuleQuantizer.quantize(relu) is relu and issubclass(GraphQuantizer, Quantizable)
    seq = L.Sequential().add(L.Linear(4, 3)).add(L.ReLU())
    out = ContainerQuantizer.quantize(seq)
    assert out is seq and isinstance(seq.modules[0], QL.QuantizedModule) and isinstance(seq.modules[1], L.ReLU)
    inp = L.Input()
    fc = L.Linear(4, 2).inputs(inp)
    g = GraphQuantizer.quantize(L.Graph(inp, L.ReLU().inputs(fc)))
    assert any(isinstance(m, QL.QuantizedModule) for m in g.all_modules())
    x = torch.rand(3, 4)
    assert tuple(g.forward(x).shape) == (3, 2)

    class Holder(L.Identity):
        def __init__(self, cell):
            super().__init__()
            self.cell = cell

    h = CellQuantizer.quantize(Holder(L.Linear(4, 4)))
    assert isinstance(h.cell, QL.QuantizedModule)
    assert isinstance(Quantizer.quantize(L.Sequential().add(L.Linear(4, 3))).modules[0], QL.QuantizedModule)


def test_serializer_entry_points(tmp_path):
    from bigdl_b200.nn import layer as L
    from bigdl_b200.serialization.serializer import ModuleLoader, ModulePersister, ModuleSerializer

    torch.manual_seed(0)
    src = L.Sequential().add(L.Linear(4, 3).setName("fc1")).add(L.ReLU()).add(L.Linear(3, 2).setName("fc2"))
    f = str(tmp_path / "m.bigdl")
    ModulePersister.saveToFile(f, None, src, True)
    back = ModuleLoader.loadFromFile(f)
    x = torch.rand(5, 4)
    assert torch.allclose(back.forward(x), src.forward(x))

    dst = L.Sequential().add(L.Linear(4, 3).setName("fc1")).add(L.ReLU()).add(L.Linear(3, 2).setName("fc2"))
    ModuleLoader.loadFromDefinition(dst, f, layers={"fc1"})
    assert torch.equal(dst.modules[0].weight.data, src.modules[0].weight.data)
    assert not torch.equal(dst.modules[2].weight.data, src.modules[2].weight.data)
    ModuleLoader.loadFromDefinition(dst, f)
    assert torch.allclose(dst.forward(x), src.forward(x))
    other = L.Sequential().add(L.Linear(4, 3).setName("nope"))
    with pytest.raises(KeyError):
        ModuleLoader.loadFromDefinition(other, f)

    d = str(tmp_path / "def.prototxt")
    ModulePersister.saveModelDefinitionToFile(d, src, True)
    text = open(d).read()
    assert "fc1" in text and "Linear" in text and "float_data" not in text and len(text) < 4000
    with pytest.raises(FileExistsError):
        ModulePersister.saveModelDefinitionToFile(d, src)
    assert torch.allclose(ModuleSerializer.load(ModuleSerializer.serialize(src)).forward(x), src.forward(x))

    class Twice(L.Identity):
        def updateOutput(self, input):
            return input * 2

    ModuleSerializer.registerModule("test.Twice", Twice)
    with pytest.raises(ValueError):
        ModuleSerializer.registerModule("test.Other")


def test_utils_object_names():
    from bigdl_b200.cli.convert_model import ConvertModel
    from bigdl_b200.interop.torchfile import TorchFile
    from bigdl_b200.parallel.distri_parameter_synchronizer import (BlockManagerParameterSynchronizer, SyncMeta,
                                                                   SyncState)
    from bigdl_b200.parallel.parameter_synchronizer import Event, ParameterSynchronizer
    from bigdl_b200.utils.engine import EngineType, MklBlas
    from bigdl_b200.utils.logger_filter import LoggerFilter

    s = BlockManagerParameterSynchronizer(0, 1)
    w, g = torch.zeros(4), torch.ones(4)
    s.init("fc", 4, 1, w, g)
    b = s.buckets["fc"]
    assert isinstance(b, SyncMeta) and SyncState.of(b) == SyncState.INIT and s.partitionID == 0 and s.totalPartition == 1
    s.put("fc")
    s.get("fc")
    assert SyncState.of(b) == SyncState.PUT_AGGREGATED
    ParameterSynchronizer.register("k", 1)
    assert isinstance(ParameterSynchronizer._events["k"], Event)
    ParameterSynchronizer.reset("k")
    assert MklBlas == EngineType.MklBlas == "mklblas" and callable(LoggerFilter.redirectSparkInfoLogs)
    assert callable(TorchFile.load) and callable(ConvertModel.main)


def test_module_api_parity_methods(tmp_path):
    from bigdl_b200.nn import layer as L
    from bigdl_b200.tensor.types import FloatType
    from bigdl_b200.transform.vision import image as V
    from bigdl_b200.dataset.base import Sample
    from bigdl_b200.optim.validation import Top1Accuracy

    torch.manual_seed(0)
    m = L.Sequential().add(L.SpatialConvolution(1, 2, 3, 3).setName("conv")).add(L.SpatialBatchNormalization(2).setName("bn")) \
        .add(L.ReLU()).add(L.View(2 * 2 * 2)).add(L.Linear(8, 3).setName("fc")).add(L.LogSoftMax())
    assert m.apply("fc") is m.modules[4] and m.apply("nope") is None
    assert [x.getName() for x in m.findModules("Linear")] == ["fc"] and m.findModules("Sequential") == [m]
    wb = m.getWeightsBias()
    assert len(wb) == 6 and m.getNumericType() == FloatType
    m.setWeightsBias([torch.full_like(w, 0.5) for w in wb])
    assert float(m.modules[4].weight.mean()) == 0.5
    with pytest.raises(ValueError):
        m.setWeightsBias(wb[:2])
    extra = m.getExtraParameter()
    assert len(extra) == 2
    m.setExtraParameter([torch.full((2,), 3.0), torch.full((2,), 4.0)])
    assert m.modules[1]._buffers["runningMean"].tolist() == [3.0, 3.0]
    with pytest.raises(ValueError):
        m.setExtraParameter([torch.zeros(2)])
    assert L.ReLU().getExtraParameter() is None and L.ReLU().setExtraParameter(None) is not None

    other = L.Sequential().add(L.SpatialConvolution(1, 2, 3, 3).setName("conv")).add(L.Linear(8, 3).setName("fc"))
    other.loadModelWeights(m, matchAll=False)
    assert float(other.modules[1].weight.mean()) == 0.5
    with pytest.raises(KeyError):
        L.Sequential().add(L.Linear(8, 3).setName("zzz")).loadModelWeights(m)
    wpath = str(tmp_path / "w.bin")
    m.saveWeights(wpath, True)
    fresh = L.Sequential().add(L.Linear(8, 3).setName("fc"))
    fresh.loadWeights(wpath)
    assert float(fresh.modules[0].weight.mean()) == 0.5
    dpath = str(tmp_path / "def.txt")
    m.saveDefinition(dpath, True)
    assert "conv" in open(dpath).read()
    assert m.setLine(" | ").line == " | " and isinstance(m.getHashCode(), int) and m.checkEngineType() is m

    feats = []
    for i in range(4):
        f = V.ImageFeature(np.zeros((4, 4, 1), np.uint8))
        f["sample"] = Sample(torch.rand(1, 4, 4), torch.tensor([float(i % 3 + 1)]))
        feats.append(f)
    bad = V.ImageFeature(np.zeros((4, 4, 1), np.uint8))
    bad.isValid = False
    res = m.evaluate().evaluateImage(V.LocalImageFrame(feats + [bad]), [Top1Accuracy()], 2)
    assert res[0][0].result()[1] == 4

    inp = L.Input()
    g = L.Graph(inp, m.getEndNodes([inp]))
    assert len(g.getForwardExecutions()) == 7 and g.getSortedForwardExecutions()[0] is inp
    assert g.getStopGradientLayers() == set() and g.resetModules() is g and g.buildBackwardGraph() is g
    assert len(g.toIRgraph()) == 7 and g.setInputFormats(["NCHW"]).inputFormats == ["NCHW"]
    x = torch.rand(2, 1, 4, 4)
    assert g.findInput(inp, x) is x and torch.allclose(g.evaluate().forward(x), m.forward(x))

    rec = L.Recurrent().add(L.LSTM(5, 4, 0.2))
    assert rec.getCell() is rec.cell and rec.cell.hiddenSizeOfPreTopo == 16 and len(rec.findDropouts()) == 1
    hid = rec.cell.hidResize(None, 3)
    assert len(hid) == 2 and tuple(hid[1].shape) == (3, 4)
    assert tuple(L.GRU(5, 4).hidResize(None, 2).shape) == (2, 4)
    from bigdl_b200.optim.regularizer import L2Regularizer
    cell = L.RnnCell(5, 4, wRegularizer=L2Regularizer(0.1))
    cell.regluarized(False)
    assert not cell.i2h.wRegularizer.isRegualrized if hasattr(cell.i2h.wRegularizer, "isRegualrized") else True
    cell.regluarized(True)


def test_cell_dropout_follows_evaluate():
    from bigdl_b200.nn import layer as L

    for cell in (L.LSTM(5, 4, 0.5), L.GRU(5, 4, 0.5)):
        rec = L.Recurrent().add(cell)
        x = torch.rand(2, 3, 5)
        rec.evaluate()
        assert torch.equal(rec.forward(x), rec.forward(x).clone())           # no dropout noise at inference
        rec.training()
        assert not torch.equal(rec.forward(x).clone(), rec.forward(x).clone())


def test_nn_utils_basemodule_and_primitives():
    import torch.nn.functional as F

    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn import utils as U
    from bigdl_b200.nn.module import BaseModule
    from bigdl_b200.utils.table import T

    a = T(torch.ones(2, 2), T(torch.ones(3)))
    b = T(torch.full((2, 2), 2.0), T(torch.full((3,), 3.0)))
    U.recursiveAdd(a, 0.5, b)
    assert a[1].tolist() == [[2, 2], [2, 2]] and a[2][1].tolist() == [2.5, 2.5, 2.5]
    U.recursiveCopy(a, b)
    assert a[2][1].tolist() == [3, 3, 3]
    U.recursiveFill(a, 7.0)
    assert float(a[1].sum()) == 28.0
    z = U.recursiveResizeAs(None, b)
    assert tuple(z[1].shape) == (2, 2) and float(z[2][1].sum()) == 0.0 and U.recursiveResizeAs(z, b)[1] is z[1]
    with pytest.raises(ValueError):
        U.recursiveAdd(torch.ones(1), 1.0, b)
    t1 = T()
    U.zeroTableCopy(t1, T(torch.ones(2), "tag"))
    assert t1[1].tolist() == [0, 0] and t1[2] == "tag"
    assert U.Utils.shuffle(torch.arange(6).reshape(2, 3), [1, 0]).tolist() == [[0, 3], [1, 4], [2, 5]]
    with pytest.raises(ValueError):
        U.shuffle(torch.zeros(2, 2), [0, 0])
    src, dst = L.Linear(3, 2).setName("a"), L.Linear(3, 2)
    U.copyModule(src, dst)
    assert torch.equal(dst.weight.data, src.weight.data) and dst.getName() == "a"
    m = L.Sequential().add(src).add(L.ReLU().setName("r"))
    assert set(U.getNamedModules(m)) >= {"a", "r"}
    m.forward(torch.rand(2, 3))
    f, bwd = U.calculateFwdBwdTime(m.getTimes())
    assert f > 0 and bwd >= 0

    x = torch.randn(2, 3, 6, 5)
    cols = L.NNPrimitive.im2col(x, 3, 2, 2, 1, 1, 0)
    w = torch.randn(4, 3, 3, 2)
    want = F.conv2d(x, w, stride=(2, 1), padding=(1, 0))
    assert torch.allclose((w.reshape(4, -1) @ cols).reshape(want.shape), want, atol=1e-5)
    back = L.NNPrimitive.col2im(torch.ones_like(cols), (6, 5), 3, 2, 2, 1, 1, 0)
    assert tuple(back.shape) == (2, 3, 6, 5) and float(back.max()) >= 2.0        # overlapping patches are summed

    class TwoLayer(BaseModule):
        def buildModel(self):
            return L.Sequential().add(L.Linear(3, 4)).add(L.Tanh())

    bm = TwoLayer()
    y = bm.forward(torch.rand(5, 3))
    assert tuple(y.shape) == (5, 4) and len(bm.parameters()[0]) == 2
    bm.evaluate()
    assert not bm.model.isTraining()
    g = bm.backward(torch.rand(5, 3), torch.ones(5, 4))
    assert tuple(g.shape) == (5, 3)
    assert L.ErrorInfo.constrainInputAsVectorOrBatch and L.Scheduler and L.FrameManager


def test_prauc_reference_spec_value():
    """T/optim/ValidationSpec.scala "precision recall auc": 0.7916667 on 4 samples; empty input is an error."""
    r = Vd.PrecisionRecallAUC()(torch.tensor([0.1, 0.4, 0.35, 0.8]), torch.tensor([0.0, 0.0, 1.0, 1.0]))
    assert r.result() == (pytest.approx(0.7916667, abs=1e-6), 4)
    with pytest.raises(ValueError):
        Vd.PrecisionRecallAUC()(torch.zeros(0), torch.zeros(0))
    with pytest.raises(ValueError):
        Vd.PrecisionRecallAUC()(torch.zeros(3), torch.zeros(2))


def test_hit_ratio_ndcg_reference_spec_values():
    """T/optim/ValidationSpec.scala "HR@10", "ndcg", "CongituousResult"."""
    o = torch.arange(1, 1001).float() / 1000
    t = torch.zeros(1000)
    t[999] = 1
    kw = {"neg_num": 999} if "neg_num" in Vd.HitRatio.__init__.__code__.co_varnames else {}
    hr, nd = (Vd.HitRatio(**kw), Vd.NDCG(**kw)) if kw else (Vd.HitRatio(10, 999), Vd.NDCG(10, 999))
    assert hr(o, t).result()[0] == 1.0 and nd(o, t).result()[0] == pytest.approx(1.0)
    o[999] = 0.9988
    assert hr(o, t).result()[0] == 1.0 and nd(o, t).result()[0] == pytest.approx(0.63092977, abs=1e-6)
    o[999] = 0.9888
    assert hr(o, t).result()[0] == 0.0 and nd(o, t).result()[0] == 0.0
    r = (Vd.ContiguousResult(0.2, 2, "HR@10") + Vd.ContiguousResult(0.1, 1, "HR@10")).result()
    assert r[0] == pytest.approx(0.1) and r[1] == 3
