"""Quantiser on CPU: module swapping, emulated fp8 arithmetic accuracy, calibration bookkeeping, backward is an error.
Mirrors S/test/.../nn/quantized/{QuantizableSpec,LinearSpec,SpatialConvolutionSpec}.scala."""
import numpy as np
import pytest
import torch

from bigdl_b200.nn import layer as L
from bigdl_b200 import quantized as QZ
from bigdl_b200.quantized import Quantizer


def _lenet_like():
    return (L.Sequential()
            .add(L.SpatialConvolution(3, 16, 3, 3, 1, 1, 1, 1))
            .add(L.SpatialBatchNormalization(16))
            .add(L.ReLU())
            .add(L.SpatialConvolution(16, 32, 3, 3, 2, 2, 1, 1))
            .add(L.ReLU())
            .add(L.SpatialMaxPooling(2, 2, 2, 2))
            .add(L.Reshape([32 * 4 * 4]))
            .add(L.Linear(32 * 4 * 4, 64))
            .add(L.ReLU())
            .add(L.Linear(64, 10)))


def test_quantize_swaps_layers_and_keeps_original():
    torch.manual_seed(0)
    m = _lenet_like()
    x = torch.randn(4, 3, 16, 16)
    m.evaluate()
    ref = m.forward(x).clone()
    q = m.quantize()
    kinds = [type(c).__name__ for c in q.modules]
    assert isinstance(q.modules[0], QZ.SpatialConvolution) and isinstance(q.modules[7], QZ.Linear), kinds
    assert isinstance(m.modules[0], L.SpatialConvolution) and not isinstance(m.modules[0], QZ.QuantizedModule)
    out = q.forward(x)
    rel = float((out - ref).norm() / ref.norm())
    assert rel < 0.08, rel
    # the float model is untouched
    assert torch.allclose(m.forward(x), ref)


def test_backward_raises():
    q = L.Sequential().add(L.Linear(8, 4)).quantize()
    x = torch.randn(2, 8)
    q.forward(x)
    with pytest.raises(ValueError, match="quantized model"):
        q.modules[0].backward(x, torch.randn(2, 4))


def test_graph_quantize():
    torch.manual_seed(1)
    inp = L.Input()
    c1 = L.SpatialConvolution(16, 16, 1, 1)(inp)
    r1 = L.ReLU()(c1)
    c2 = L.SpatialConvolution(16, 16, 3, 3, 1, 1, 1, 1)(r1)
    add = L.CAddTable()([c2, inp])
    g = L.Model([inp], [add]).evaluate()
    x = torch.randn(2, 16, 8, 8)
    ref = g.forward(x).clone()
    q = g.quantize()
    names = [type(n.element).__module__ for n in q.exec_order]
    assert sum(n.endswith("quantized.layers") for n in names) == 2
    out = q.forward(x)
    assert float((out - ref).norm() / ref.norm()) < 0.08


def test_calibrate_sets_static_scales_and_links():
    torch.manual_seed(2)
    m = _lenet_like().evaluate()
    q = m.quantize()
    xs = [torch.randn(4, 3, 16, 16) for _ in range(3)]
    Quantizer.calibrate(q, xs)
    qs = Quantizer.quantized_layers(q)
    assert all(l.static_scale for l in qs)
    want = max(float(x.abs().max()) for x in xs)
    assert abs(qs[0].getInputScales()[0][0] - want) < 1e-5
    # conv0 (+folded BN + fused ReLU) feeds conv1 directly: linked; linear0 → linear1 linked
    assert qs[0]._emit_q is qs[1] and qs[2]._emit_q is qs[3]
    assert qs[1]._emit_q is None            # max-pool in between
    ref = m.forward(xs[0]).clone()
    out = q.forward(xs[0])
    assert float((out - ref).norm() / ref.norm()) < 0.1


def test_weight_quantisation_error_bound():
    torch.manual_seed(3)
    lin = L.Linear(64, 32)
    q = QZ.Linear.from_float(lin)
    w, wd = lin.weight.detach(), q.dequantized_weight()
    # e4m3 has 3 mantissa bits: relative error <= 2^-4 per element for normal values, absolute error bounded by the
    # sub-normal step (2^-9 * absmax/448 * 448) for tiny ones
    err = (w - wd).abs()
    bound = w.abs() * 2 ** -4 + q._buffers["weight_absmax"].reshape(-1, 1) * 2 ** -9
    assert bool((err <= bound + 1e-9).all())


def test_quantized_save_load_roundtrip(tmp_path):
    torch.manual_seed(4)
    m = L.Sequential().add(L.Linear(16, 8)).add(L.ReLU()).add(L.Linear(8, 4)).evaluate()
    q = m.quantize()
    x = torch.randn(3, 16)
    ref = q.forward(x).clone()
    p = str(tmp_path / "q.bigdl")
    q.saveModule(p, overWrite=True)
    q2 = L.Module.loadModule(p).evaluate()
    assert isinstance(q2.modules[0], QZ.Linear)
    assert torch.allclose(q2.forward(x), ref, atol=1e-6)


def test_calc_scales_feed_static_quantisation(tmp_path):
    """MklInt8Convertible flow: calcScales on the float model → scales saved in the .bigdl file → quantize() picks them
    up as static activation scales (S/example/mkldnn/int8/GenerateInt8Scales.scala)."""
    torch.manual_seed(5)
    m = _lenet_like().evaluate()
    m.setWeightDimMask(1, overrideSubmodules=True)
    x = torch.randn(4, 3, 16, 16)
    m.calcScales(x)
    conv0, fc = m.modules[0], m.modules[9]
    assert abs(conv0.getInputScales()[0][0] - float(x.abs().max())) < 1e-6
    assert len(conv0.getWeightScales()[0]) == 16                      # mask 1 → one scale per output channel
    assert abs(fc.getOutputScales()[0][0] - float(m.forward(x).abs().max())) < 1e-5
    assert m.getInputScales()[0][0] == conv0.getInputScales()[0][0]   # containers record their own activity too
    p = str(tmp_path / "scaled.bigdl")
    m.saveModule(p, overWrite=True)
    back = L.Module.loadModule(p)
    assert back.modules[0].getInputScales() == conv0.getInputScales()
    q = back.evaluate().quantize()
    qs = Quantizer.quantized_layers(q)
    assert all(l.static_scale for l in qs) and qs[0]._emit_q is qs[1]
    ref = m.forward(x)
    out = q.forward(x)
    assert float((out - ref).norm() / ref.norm()) < 0.1


def test_quantize_model_that_was_already_fused():
    """Fusion flags set before quantize() (a model that lived on the GPU) must not leave a ReLU on a folded BN."""
    from bigdl_b200.nn.fusion import fuse

    torch.manual_seed(6)
    m = _lenet_like().evaluate()
    x = torch.randn(4, 3, 16, 16)
    ref = m.forward(x).clone()
    fuse(m)
    assert m.modules[1]._fuse_relu                      # BN + ReLU fused
    q = m.quantize()
    Quantizer.calibrate(q, [x])
    assert not q.modules[1]._fuse_relu and q.modules[0]._fuse_relu      # the ReLU moved into the conv epilogue
    out = q.forward(x)
    assert float((out - ref).norm() / ref.norm()) < 0.1


def test_quant_serializer_roundtrip():
    import torch
    from bigdl_b200.nn.layer import Linear
    from bigdl_b200.quantized import QuantSerializer
    from bigdl_b200.quantized.layers import Linear as QLinear

    torch.manual_seed(0)
    a, b = QLinear.from_float(Linear(16, 8)), QLinear.from_float(Linear(16, 8))
    x = torch.randn(4, 16)
    assert not torch.allclose(a.forward(x), b.forward(x))
    payload = QuantSerializer.serialize(a)
    assert payload["weight_q"].dtype.name == "uint8" and payload["desc"] == {"outputSize": 8, "inputSize": 16}
    QuantSerializer.deserialize(b, payload)
    assert torch.equal(a.forward(x), b.forward(x))


def _trained_mlp(seed=3, steps=150):
    """A small classifier trained on synthetic clustered data, so that top-1 agreement is a meaningful statistic (the
    reference's integration check trains/loads real models: T/integration/Quantization.scala:88-148)."""
    import torch
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.nn.layer import Linear, LogSoftMax, ReLU, Sequential
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.utils.random_generator import RNG

    RNG.setSeed(seed)
    g = torch.Generator().manual_seed(seed)
    centers = torch.randn(10, 96, generator=g) * 2.0
    # feature scales spread over three orders of magnitude: the case per-tensor scales handle worst
    fscale = torch.exp2(torch.linspace(-6, 4, 96))

    def batch(n):
        y = torch.randint(0, 10, (n,), generator=g)
        x = (centers[y] + torch.randn(n, 96, generator=g)) * fscale
        return x, (y + 1).float()

    m = Sequential().add(Linear(96, 128)).add(ReLU()).add(Linear(128, 64)).add(ReLU()).add(Linear(64, 10)).add(LogSoftMax())
    with torch.no_grad():
        m.modules[0].weight.div_(fscale)           # first layer absorbs the feature scales (a trained net would learn this)
    crit, om = ClassNLLCriterion(), SGD(0.05, momentum=0.9)
    w, gr = m.getParameters()
    m.training()
    for _ in range(steps):
        x, y = batch(64)
        m.zeroGradParameters()
        out = m.forward(x)
        crit.forward(out, y)
        m.backward(x, crit.backward(out, y))
        om.update(w, gr)
        m._mark_weights_dirty()
    m.evaluate()
    return m, batch


def test_top1_agreement_of_quantised_models():
    """Quantised vs float predictions on held-out data: >= 99.5 % top-1 agreement for both scaling modes (the reference
    asks for 'same top-1 accuracy'), and the block-scaled (MX) Linear is at least as close to the float logits as the
    per-tensor one on inputs whose features span many octaves."""
    import torch
    from bigdl_b200.quantized.quantizer import Quantizer

    m, batch = _trained_mlp()
    x, y = batch(2000)
    ref = m.forward(x).float()
    acc = float((ref.argmax(1) + 1 == y).float().mean())
    assert acc > 0.9, acc                                  # the float model actually learned the task
    errs = {}
    for mode in (False, True):
        q = Quantizer.quantize(m, block_scale=mode)
        out = q.forward(x).float()
        agree = float((out.argmax(1) == ref.argmax(1)).float().mean())
        assert agree >= 0.995, (mode, agree)
        errs[mode] = float((out - ref).norm() / ref.norm())
    assert errs[True] <= errs[False] * 1.05, errs
    from bigdl_b200.quantized.layers import Linear as QLinear
    assert all(l.block_scale for l in Quantizer.quantized_layers(Quantizer.quantize(m, block_scale=True)) if isinstance(l, QLinear))
