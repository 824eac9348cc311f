"""Breadth tests in the style of the reference's per-layer specs: every listed layer (a) survives a ``.bigdl`` round trip
with identical outputs (ModuleSerializationTest) and (b) passes a finite-difference check of its input — and, where it
has any, parameter — gradients (GradientChecker)."""
import pytest
import torch

from bigdl_b200.nn import layer as L
from bigdl_b200.nn.module import Module
from bigdl_b200.utils.gradient_checker import GradientChecker
from bigdl_b200.utils.random_generator import RNG
from bigdl_b200.utils.table import T

IMG = (2, 4, 6, 6)
VEC = (3, 8)
SEQ = (2, 5, 4)


def _img():
    return torch.randn(*IMG)


def _vec():
    return torch.randn(*VEC)


def _pos(shape):
    return torch.rand(*shape) + 0.5


# name -> (constructor, input maker, smooth enough for a finite-difference check)
CASES = {
    "Linear": (lambda: L.Linear(8, 5), _vec, True),
    "Bilinear": (lambda: L.Bilinear(8, 8, 3), lambda: T(_vec(), _vec()), True),
    "Add": (lambda: L.Add(8), _vec, True),
    "CAdd": (lambda: L.CAdd([1, 8]), _vec, True),
    "Mul": (lambda: L.Mul(), _vec, True),
    "CMul": (lambda: L.CMul([1, 8]), _vec, True),
    "Cosine": (lambda: L.Cosine(8, 4), _vec, True),
    "Euclidean": (lambda: L.Euclidean(8, 4), _vec, True),
    "Maxout": (lambda: L.Maxout(8, 4, 3), _vec, False),
    "Highway": (lambda: L.Highway(8), _vec, True),
    "SpatialConvolution": (lambda: L.SpatialConvolution(4, 6, 3, 3, 1, 1, 1, 1), _img, True),
    "SpatialConvolution_group": (lambda: L.SpatialConvolution(4, 6, 3, 3, 2, 2, 1, 1, 2), _img, True),
    "SpatialDilatedConvolution": (lambda: L.SpatialDilatedConvolution(4, 5, 3, 3, 1, 1, 2, 2, 2, 2), _img, True),
    "SpatialFullConvolution": (lambda: L.SpatialFullConvolution(4, 3, 3, 3, 2, 2, 1, 1, 1, 1), _img, True),
    "SpatialSeparableConvolution": (lambda: L.SpatialSeparableConvolution(4, 6, 2, 3, 3), _img, True),
    "TemporalConvolution": (lambda: L.TemporalConvolution(4, 6, 3), lambda: torch.randn(*SEQ), True),
    "VolumetricConvolution": (lambda: L.VolumetricConvolution(2, 3, 2, 2, 2), lambda: torch.randn(2, 2, 4, 5, 5), True),
    "LocallyConnected2D": (lambda: L.LocallyConnected2D(4, 6, 6, 3, 3, 3), _img, True),
    "SpatialBatchNormalization": (lambda: L.SpatialBatchNormalization(4), _img, True),
    "BatchNormalization": (lambda: L.BatchNormalization(8), _vec, True),
    "LayerNormalization": (lambda: L.LayerNormalization(8), _vec, True),
    "Normalize": (lambda: L.Normalize(2.0), _vec, True),
    "SpatialCrossMapLRN": (lambda: L.SpatialCrossMapLRN(3, 0.1, 0.75, 1.0), _img, True),
    "SpatialMaxPooling": (lambda: L.SpatialMaxPooling(2, 2, 2, 2), _img, False),
    "SpatialAveragePooling": (lambda: L.SpatialAveragePooling(3, 3, 2, 2, 1, 1), _img, True),
    "ResizeBilinear": (lambda: L.ResizeBilinear(9, 7), _img, True),
    "UpSampling2D": (lambda: L.UpSampling2D([2, 2]), _img, True),
    "SpatialZeroPadding": (lambda: L.SpatialZeroPadding(1, 2, 1, 0), _img, True),
    "Cropping2D": (lambda: L.Cropping2D([1, 1], [1, 2]), _img, True),
    "ReLU6": (lambda: L.ReLU6(), _vec, False),
    "LeakyReLU": (lambda: L.LeakyReLU(0.1), _vec, False),
    "PReLU": (lambda: L.PReLU(4), _img, False),
    "SReLU": (lambda: L.SReLU([4, 6, 6]), _img, False),
    "ELU": (lambda: L.ELU(1.0), _vec, True),
    "Sigmoid": (lambda: L.Sigmoid(), _vec, True),
    "Tanh": (lambda: L.Tanh(), _vec, True),
    "SoftMax": (lambda: L.SoftMax(), _vec, True),
    "LogSoftMax": (lambda: L.LogSoftMax(), _vec, True),
    "SoftPlus": (lambda: L.SoftPlus(), _vec, True),
    "SoftSign": (lambda: L.SoftSign(), _vec, True),
    "LogSigmoid": (lambda: L.LogSigmoid(), _vec, True),
    "TanhShrink": (lambda: L.TanhShrink(), _vec, True),
    "Power": (lambda: L.Power(2.0, 1.5, 0.5), lambda: _pos(VEC), True),
    "Sqrt": (lambda: L.Sqrt(), lambda: _pos(VEC), True),
    "Square": (lambda: L.Square(), _vec, True),
    "Log": (lambda: L.Log(), lambda: _pos(VEC), True),
    "Exp": (lambda: L.Exp(), _vec, True),
    "AddConstant": (lambda: L.AddConstant(1.5), _vec, True),
    "MulConstant": (lambda: L.MulConstant(-2.0), _vec, True),
    "Reshape": (lambda: L.Reshape([2, 4]), _vec, True),
    "View": (lambda: L.View(4, 2).setNumInputDims(1), _vec, True),
    "Transpose": (lambda: L.Transpose([(2, 3)]), _img, True),
    "Narrow": (lambda: L.Narrow(2, 2, 3), _vec, True),
    "Select": (lambda: L.Select(2, 3), _vec, True),
    "Replicate": (lambda: L.Replicate(3, 2), _vec, True),
    "Sum": (lambda: L.Sum(2), _vec, True),
    "Mean": (lambda: L.Mean(2), _vec, True),
    "Padding": (lambda: L.Padding(2, 2, 2, 0.0), _vec, True),
    "Squeeze": (lambda: L.Squeeze(2), lambda: torch.randn(3, 1, 5), True),
    "Unsqueeze": (lambda: L.Unsqueeze(2), _vec, True),
    "JoinTable": (lambda: L.JoinTable(2, 2), lambda: T(_vec(), _vec()), True),
    "CAddTable": (lambda: L.CAddTable(), lambda: T(_vec(), _vec()), True),
    "CMulTable": (lambda: L.CMulTable(), lambda: T(_vec(), _vec()), True),
    "CSubTable": (lambda: L.CSubTable(), lambda: T(_vec(), _vec()), True),
    "CDivTable": (lambda: L.CDivTable(), lambda: T(_vec(), _pos(VEC)), True),
    "DotProduct": (lambda: L.DotProduct(), lambda: T(_vec(), _vec()), True),
    "PairwiseDistance": (lambda: L.PairwiseDistance(2), lambda: T(_vec(), _vec()), True),
    "CosineDistance": (lambda: L.CosineDistance(), lambda: T(_vec(), _vec()), True),
    "MM": (lambda: L.MM(), lambda: T(torch.randn(2, 3, 4), torch.randn(2, 4, 5)), True),
    "MV": (lambda: L.MV(), lambda: T(torch.randn(2, 3, 4), torch.randn(2, 4)), True),
    "Recurrent_LSTM": (lambda: L.Recurrent().add(L.LSTM(4, 5)), lambda: torch.randn(*SEQ), True),
    "Recurrent_GRU": (lambda: L.Recurrent().add(L.GRU(4, 5)), lambda: torch.randn(*SEQ), True),
    "Recurrent_RnnCell": (lambda: L.Recurrent().add(L.RnnCell(4, 5, L.Tanh())), lambda: torch.randn(*SEQ), True),
    "TimeDistributed": (lambda: L.TimeDistributed(L.Linear(4, 3)), lambda: torch.randn(*SEQ), True),
    "Sequential_ConcatTable": (lambda: L.Sequential().add(L.ConcatTable().add(L.Linear(8, 4)).add(L.Linear(8, 4)))
                               .add(L.CAddTable()), _vec, True),
    "Concat": (lambda: L.Concat(2).add(L.Linear(8, 3)).add(L.Linear(8, 2)), _vec, True),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_layer_serialization_roundtrip(name, tmp_path):
    RNG.setSeed(5)
    torch.manual_seed(5)
    ctor, make_input, _ = CASES[name]
    m = ctor()
    x = make_input()
    m.evaluate()
    y = m.forward(x)
    path = str(tmp_path / "layer.bigdl")
    m.saveModule(path, None, True)
    m2 = Module.loadModule(path)
    m2.evaluate()
    y2 = m2.forward(x)
    ya = y.to_list() if hasattr(y, "to_list") else [y]
    yb = y2.to_list() if hasattr(y2, "to_list") else [y2]
    assert type(m2) is type(m) and len(ya) == len(yb)
    for a, b in zip(ya, yb):
        assert a.shape == b.shape and torch.allclose(a, b, atol=1e-6), name


@pytest.mark.parametrize("name", sorted(n for n, c in CASES.items() if c[2]))
def test_layer_gradients_match_finite_differences(name):
    RNG.setSeed(6)
    torch.manual_seed(6)
    ctor, make_input, _ = CASES[name]
    m = ctor()
    x = make_input()
    checker = GradientChecker(1e-2, 2e-2)
    assert checker.checkLayer(m, x), f"{name}: input gradient"
    if m.parameters()[0]:
        assert checker.checkWeight(m, x), f"{name}: parameter gradient"


# ------------------------------------------------------------------------------------------------- criterions
from bigdl_b200.nn import criterion as C  # noqa: E402


def _probs(shape):
    return torch.softmax(torch.randn(*shape), dim=-1)


def _labels(n, k):
    return torch.randint(1, k + 1, (n,)).float()


def _pm1(shape):
    return (torch.randint(0, 2, shape).float() * 2 - 1)


CRITERIONS = {
    "ClassNLL": (lambda: C.ClassNLLCriterion(), lambda: torch.log_softmax(torch.randn(4, 5), 1), lambda: _labels(4, 5)),
    "ClassNLL_weights": (lambda: C.ClassNLLCriterion(torch.tensor([0.5, 1.0, 2.0, 1.5, 0.1])),
                         lambda: torch.log_softmax(torch.randn(4, 5), 1), lambda: _labels(4, 5)),
    "CrossEntropy": (lambda: C.CrossEntropyCriterion(), lambda: torch.randn(4, 5), lambda: _labels(4, 5)),
    "MSE": (lambda: C.MSECriterion(), lambda: torch.randn(4, 5), lambda: torch.randn(4, 5)),
    "Abs": (lambda: C.AbsCriterion(), lambda: torch.randn(4, 5), lambda: torch.randn(4, 5) + 3),
    "BCE": (lambda: C.BCECriterion(), lambda: torch.rand(4, 5) * 0.8 + 0.1, lambda: torch.randint(0, 2, (4, 5)).float()),
    "SmoothL1": (lambda: C.SmoothL1Criterion(), lambda: torch.randn(4, 5), lambda: torch.randn(4, 5) + 3),
    "Margin": (lambda: C.MarginCriterion(), lambda: torch.randn(4, 5) * 3, lambda: _pm1((4, 5))),
    "MarginRanking": (lambda: C.MarginRankingCriterion(), lambda: T(torch.randn(6) * 3, torch.randn(6) * 3),
                      lambda: _pm1((6,))),
    "MultiMargin": (lambda: C.MultiMarginCriterion(), lambda: torch.randn(4, 5) * 3, lambda: _labels(4, 5)),
    "MultiLabelSoftMargin": (lambda: C.MultiLabelSoftMarginCriterion(), lambda: torch.randn(4, 5),
                             lambda: torch.randint(0, 2, (4, 5)).float()),
    "SoftMargin": (lambda: C.SoftMarginCriterion(), lambda: torch.randn(4, 5), lambda: _pm1((4, 5))),
    "HingeEmbedding": (lambda: C.HingeEmbeddingCriterion(2.0), lambda: torch.rand(4, 5) * 4, lambda: _pm1((4, 5))),
    "CosineEmbedding": (lambda: C.CosineEmbeddingCriterion(0.1), lambda: T(torch.randn(4, 5), torch.randn(4, 5)),
                        lambda: _pm1((4,))),
    "CosineDistance": (lambda: C.CosineDistanceCriterion(), lambda: torch.randn(4, 5), lambda: torch.randn(4, 5)),
    "CosineProximity": (lambda: C.CosineProximityCriterion(), lambda: torch.randn(4, 5), lambda: torch.randn(4, 5)),
    "DistKLDiv": (lambda: C.DistKLDivCriterion(), lambda: torch.log(_probs((4, 5))), lambda: _probs((4, 5))),
    "KullbackLeiblerDivergence": (lambda: C.KullbackLeiblerDivergenceCriterion(), lambda: _probs((4, 5)),
                                  lambda: _probs((4, 5))),
    "Poisson": (lambda: C.PoissonCriterion(), lambda: torch.rand(4, 5) + 0.5, lambda: torch.rand(4, 5) * 3),
    "DiceCoefficient": (lambda: C.DiceCoefficientCriterion(), lambda: torch.rand(4, 5), lambda: torch.randint(0, 2, (4, 5)).float()),
    "L1Cost": (lambda: C.L1Cost(), lambda: torch.randn(4, 5) + 2.5, lambda: torch.zeros(4, 5)),
    "MeanAbsolutePercentage": (lambda: C.MeanAbsolutePercentageCriterion(), lambda: torch.rand(4, 5) + 2,
                               lambda: torch.rand(4, 5) + 0.5),
    "MeanSquaredLogarithmic": (lambda: C.MeanSquaredLogarithmicCriterion(), lambda: torch.rand(4, 5) + 0.5,
                               lambda: torch.rand(4, 5) + 0.5),
    "CategoricalCrossEntropy": (lambda: C.CategoricalCrossEntropy(), lambda: _probs((4, 5)),
                                lambda: torch.nn.functional.one_hot(torch.randint(0, 5, (4,)), 5).float()),
    "DotProduct": (lambda: C.DotProductCriterion(), lambda: torch.randn(4, 5), lambda: torch.randn(4, 5)),
    "ClassSimplex": (lambda: C.ClassSimplexCriterion(5), lambda: torch.randn(4, 5), lambda: _labels(4, 5)),
    "SoftmaxWith": (lambda: C.SoftmaxWithCriterion(), lambda: torch.randn(2, 5, 3, 3), lambda: torch.randint(1, 6, (2, 1, 3, 3)).float()),
    "TimeDistributed": (lambda: C.TimeDistributedCriterion(C.MSECriterion(), True), lambda: torch.randn(3, 4, 5),
                        lambda: torch.randn(3, 4, 5)),
    "Parallel": (lambda: C.ParallelCriterion().add(C.MSECriterion(), 0.5).add(C.AbsCriterion(), 2.0),
                 lambda: T(torch.randn(4, 5), torch.randn(4, 3)), lambda: T(torch.randn(4, 5), torch.randn(4, 3) + 3)),
    "Multi": (lambda: C.MultiCriterion().add(C.MSECriterion(), 0.5).add(C.SmoothL1Criterion(), 2.0),
              lambda: torch.randn(4, 5), lambda: torch.randn(4, 5) + 3),
}


@pytest.mark.parametrize("name", sorted(CRITERIONS))
def test_criterion_gradient_matches_finite_differences(name):
    RNG.setSeed(7)
    torch.manual_seed(7)
    make, make_in, make_tgt = CRITERIONS[name]
    crit, x, t = make(), make_in(), make_tgt()
    loss = float(crit.forward(x, t))
    assert loss == loss          # not NaN
    g = crit.backward(x, t)
    xs = x.to_list() if hasattr(x, "to_list") else [x]
    gs = g.to_list() if hasattr(g, "to_list") else [g]
    eps = 1e-3
    for xi, gi in zip(xs, gs):
        flat = xi.reshape(-1)
        assert flat.data_ptr() == xi.data_ptr()
        for i in torch.randperm(flat.numel())[:24]:
            old = float(flat[i])
            flat[i] = old + eps
            up = float(crit.forward(x, t))
            flat[i] = old - eps
            dn = float(crit.forward(x, t))
            flat[i] = old
            num, ana = (up - dn) / (2 * eps), float(gi.reshape(-1)[i])
            assert abs(num - ana) <= 2e-2 * max(1.0, abs(ana), abs(num)), (name, int(i), num, ana)


# ------------------------------------------------------------------------------------------------- rarely used layers
def _vol():
    return torch.randn(2, 3, 4, 6, 6)


SMOKE = {
    "BinaryThreshold": (lambda: L.BinaryThreshold(0.1), _vec, (3, 8)),
    "RReLU": (lambda: L.RReLU(), _vec, (3, 8)),
    "Clamp": (lambda: L.Clamp(-1, 1), _vec, (3, 8)),
    "SpatialDropout1D": (lambda: L.SpatialDropout1D(0.3), lambda: torch.randn(2, 5, 4), (2, 5, 4)),
    "SpatialDropout2D": (lambda: L.SpatialDropout2D(0.3), _img, IMG),
    "SpatialDropout3D": (lambda: L.SpatialDropout3D(0.3), _vol, (2, 3, 4, 6, 6)),
    "GaussianDropout": (lambda: L.GaussianDropout(0.3), _vec, (3, 8)),
    "GaussianNoise": (lambda: L.GaussianNoise(0.3), _vec, (3, 8)),
    "GaussianSampler": (lambda: L.GaussianSampler(), lambda: T(_vec(), _vec()), (3, 8)),
    "NegativeEntropyPenalty": (lambda: L.NegativeEntropyPenalty(0.1), lambda: torch.softmax(_vec(), 1), (3, 8)),
    "ActivityRegularization": (lambda: L.ActivityRegularization(0.1, 0.1), _vec, (3, 8)),
    "MapTable": (lambda: L.MapTable(L.Linear(8, 3)), lambda: T(_vec(), _vec()), [(3, 3), (3, 3)]),
    "Bottle": (lambda: L.Bottle(L.Linear(8, 3), 2, 2), lambda: torch.randn(2, 3, 8), (2, 3, 3)),
    "BifurcateSplitTable": (lambda: L.BifurcateSplitTable(2), _vec, [(3, 4), (3, 4)]),
    "FlattenTable": (lambda: L.FlattenTable(), lambda: T(_vec(), T(_vec(), _vec())), [(3, 8)] * 3),
    "NarrowTable": (lambda: L.NarrowTable(2, 2), lambda: T(_vec(), _vec(), _vec()), [(3, 8)] * 2),
    "Pack": (lambda: L.Pack(1), lambda: T(_vec(), _vec()), (2, 3, 8)),
    "CMinTable": (lambda: L.CMinTable(), lambda: T(_vec(), _vec()), (3, 8)),
    "CAveTable": (lambda: L.CAveTable(), lambda: T(_vec(), _vec()), (3, 8)),
    "MixtureTable": (lambda: L.MixtureTable(), lambda: T(torch.softmax(torch.randn(3, 2), 1), T(_vec(), _vec())), (3, 8)),
    "InferReshape": (lambda: L.InferReshape([-1, 4]), _vec, (6, 4)),
    "Contiguous": (lambda: L.Contiguous(), lambda: _vec().t(), (8, 3)),
    "Index": (lambda: L.Index(1), lambda: T(_vec(), torch.tensor([1., 3.])), (2, 8)),
    "Masking": (lambda: L.Masking(0.0), lambda: torch.randn(2, 5, 4), (2, 5, 4)),
    "ExpandSize": (lambda: L.ExpandSize([3, 8]), lambda: torch.randn(1, 8), (3, 8)),
    "Reverse": (lambda: L.Reverse(2), _vec, (3, 8)),
    "Cropping3D": (lambda: L.Cropping3D([1, 1], [1, 1], [1, 1]), _vol, (2, 3, 2, 4, 4)),
    "Min": (lambda: L.Min(2), _vec, (3,)),
    "CrossProduct": (lambda: L.CrossProduct(), lambda: T(_vec(), _vec(), _vec()), (3, 3)),
    "NormalizeScale": (lambda: L.NormalizeScale(2.0, scale=20.0, size=[1, 4, 1, 1]), _img, IMG),
    "SpatialContrastiveNormalization": (lambda: L.SpatialContrastiveNormalization(4), _img, IMG),
    "SpatialShareConvolution": (lambda: L.SpatialShareConvolution(4, 5, 3, 3), _img, (2, 5, 4, 4)),
    "SpatialConvolutionMap": (lambda: L.SpatialConvolutionMap(L.SpatialConvolutionMap.full(4, 3), 3, 3), _img, (2, 3, 4, 4)),
    "VolumetricFullConvolution": (lambda: L.VolumetricFullConvolution(3, 2, 2, 2, 2), _vol, (2, 2, 5, 7, 7)),
    "VolumetricAveragePooling": (lambda: L.VolumetricAveragePooling(2, 2, 2), _vol, (2, 3, 2, 3, 3)),
    "UpSampling3D": (lambda: L.UpSampling3D([2, 2, 2]), _vol, (2, 3, 8, 12, 12)),
    "FeedForwardNetwork": (lambda: L.FeedForwardNetwork(8, 16, 0.0), lambda: torch.randn(2, 3, 8), (2, 3, 8)),
}


@pytest.mark.parametrize("name", sorted(SMOKE))
def test_rarely_used_layers_forward_backward(name):
    RNG.setSeed(9)
    torch.manual_seed(9)
    ctor, make_input, shape = SMOKE[name]
    m = ctor()
    x = make_input()
    y = m.forward(x)
    ys = y.to_list() if hasattr(y, "to_list") else [y]
    assert [tuple(t.shape) for t in ys] == (shape if isinstance(shape, list) else [tuple(shape)]), name
    assert all(torch.isfinite(t.float()).all() for t in ys)
    go = T(*[torch.ones_like(t) for t in ys]) if hasattr(y, "to_list") else torch.ones_like(y)
    gi = m.backward(x, go)
    gis = gi.to_list() if hasattr(gi, "to_list") else [gi]
    xs = [t for t in (x.to_list() if hasattr(x, "to_list") else [x])]
    flat_x = []
    for t in xs:
        flat_x.extend(t.to_list() if hasattr(t, "to_list") else [t])
    flat_g = []
    for t in gis:
        flat_g.extend(t.to_list() if hasattr(t, "to_list") else [t])
    for a, b in zip(flat_x, flat_g):
        if b is not None and a.is_floating_point():
            assert tuple(a.shape) == tuple(b.shape), name
