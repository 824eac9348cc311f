"""Vision transformers and criterions that no other test touches (S/transform/vision/image/augmentation/*,
S/nn/{MultiLabelMarginCriterion,L1HingeEmbeddingCriterion,TransformerCriterion,TimeDistributedMaskCriterion,
SmoothL1CriterionWithWeights}.scala), against simple oracles."""
import numpy as np
import torch
import torch.nn.functional as F

from bigdl_b200.nn import criterion as C
from bigdl_b200.nn import layer as L
from bigdl_b200.transform.vision import image as V
from bigdl_b200.utils.random_generator import RNG
from bigdl_b200.utils.table import T


def _feat():
    return V.ImageFeature(np.random.RandomState(0).randint(0, 255, (40, 60, 3)).astype(np.uint8))


def _apply(tr, f=None):
    f = f or _feat()
    out = tr.transform(f)               # the public entry point (handles validity flags, wraps transformMat)
    return out if out is not None else f


def test_vision_transformers_shapes_and_values():
    RNG.setSeed(1)
    assert np.asarray(_apply(V.MatToFloats(40, 60, 3))["floats"]).shape == (40, 60, 3)
    assert min(_apply(V.RandomAspectScale([20, 30], 1, 100))["mat"].shape[:2]) in (20, 30)
    assert _apply(V.FixedCrop(0.25, 0.25, 0.75, 0.75))["mat"].shape == (20, 30, 3)
    src = _feat()
    crop = _apply(V.FixedCrop(10, 5, 40, 25, normalized=False))
    assert crop["mat"].shape == (20, 30, 3) and np.array_equal(crop["mat"], src["mat"][5:25, 10:40])
    assert _apply(V.RandomCropper(32, 24, True))["mat"].shape == (24, 32, 3)
    flipped = _apply(V.RandomTransformer(V.HFlip(), 1.0))
    assert np.array_equal(flipped["mat"], src["mat"][:, ::-1])
    same = _apply(V.RandomTransformer(V.HFlip(), 0.0))
    assert np.array_equal(same["mat"], src["mat"])
    scaled = _apply(V.ChannelScaledNormalizer(123, 117, 104, 0.5))["mat"]
    assert scaled.dtype == np.float32
    means = np.full(40 * 60 * 3, 10.0, np.float32)
    pn = _apply(V.PixelNormalizer(means))["mat"]
    assert np.allclose(pn.reshape(-1), src["mat"].astype(np.float32).reshape(-1) - 10.0)
    f = _feat()
    f["roi"] = np.array([0, 0, 0.2, 0.2, 0.8, 0.8], dtype=np.float32)
    V.DetectionCrop("roi").transformMat(f)
    assert f["mat"].shape == (24, 36, 3)
    f = _feat()
    f["label"] = V.RoiLabel([1., 2.], [[5., 5., 30., 20.], [20., 10., 55., 35.]])
    V.RoiNormalize().transformMat(f)
    assert float(f["label"].bboxes.max()) <= 1.0
    V.RandomSampler().transformMat(f)
    V.RoiProject().transformMat(f)
    assert f["label"].bboxes.shape[1] == 4 and f["label"].bboxes.shape[0] <= 2


def test_extra_criterions_against_formulas():
    torch.manual_seed(2)
    x, t = torch.randn(4, 5), torch.randn(4, 5)
    tgt = torch.tensor([[1., 3, 0, 0, 0], [2, 0, 0, 0, 0], [5, 4, 0, 0, 0], [1, 0, 0, 0, 0]])
    ml = C.MultiLabelMarginCriterion()
    ref = F.multilabel_margin_loss(x, tgt.long() - 1)                      # torch: 0-based, -1 terminated
    assert abs(float(ml.forward(x, tgt)) - float(ref)) < 1e-5
    xr = x.clone().requires_grad_()
    F.multilabel_margin_loss(xr, tgt.long() - 1).backward()
    assert torch.allclose(ml.backward(x, tgt), xr.grad, atol=1e-5)
    # one pair per call (target is a single +1 / -1), like the reference
    l1 = float((x - t).abs().sum())
    assert abs(float(C.L1HingeEmbeddingCriterion(1.0).forward(T(x, t), torch.tensor([1.]))) - l1) < 1e-4
    assert abs(float(C.L1HingeEmbeddingCriterion(100.0).forward(T(x, t), torch.tensor([-1.]))) - (100.0 - l1)) < 1e-3
    assert float(C.L1HingeEmbeddingCriterion(1.0).forward(T(x, t), torch.tensor([-1.]))) == 0.0
    tc = C.TransformerCriterion(C.MSECriterion(), L.Square(), L.Square())
    assert abs(float(tc.forward(x, t)) - float(F.mse_loss(x ** 2, t ** 2))) < 1e-5
    assert tuple(tc.backward(x, t).shape) == (4, 5)
    logp = torch.log_softmax(torch.randn(2, 3, 4), 2)
    labels = torch.tensor([[1., 2, 0], [3, 0, 0]])                          # 0 = padding
    got = float(C.TimeDistributedMaskCriterion(C.ClassNLLCriterion(paddingValue=0), 0).forward(logp, labels))
    picked = -(logp[0, 0, 0] + logp[0, 1, 1] + logp[1, 0, 2])
    assert abs(got - float(picked) / 3) < 1e-5
    sl = C.SmoothL1CriterionWithWeights(1.0, 4)
    iw, ow = torch.rand(4, 5), torch.rand(4, 5)
    d = (x - t) * iw
    ref = (torch.where(d.abs() < 1, 0.5 * d * d, d.abs() - 0.5) * ow).sum() / 4
    assert abs(float(sl.forward(x, T(t, iw, ow))) - float(ref)) < 1e-4
