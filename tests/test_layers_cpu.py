"""Per-layer checks on the CPU path against hand-computed values or plain torch references (the role of T/nn/*Spec and
T/torch/*Spec in the reference), plus finite-difference gradient checks (T/nn/GradientChecker.scala)."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from bigdl_b200.nn import layer as L
from bigdl_b200.nn import criterion as C
from bigdl_b200.utils.random_generator import RNG
from bigdl_b200.utils.table import T, Table


def grad_check(module, x, eps=1e-3, tol=2e-2):
    """Central finite differences of sum(forward(x)) w.r.t. x vs backward(x, ones)."""
    x = x.double().float()
    y = module.forward(x)
    g = module.backward(x, torch.ones_like(y))
    num = torch.zeros_like(x)
    flat = x.reshape(-1)
    for i in range(flat.numel()):
        old = float(flat[i])
        flat[i] = old + eps
        a = float(module.forward(x).double().sum())
        flat[i] = old - eps
        b = float(module.forward(x).double().sum())
        flat[i] = old
        num.reshape(-1)[i] = (a - b) / (2 * eps)
    assert torch.allclose(g, num, rtol=tol, atol=tol), (g - num).abs().max()


def test_linear_matches_formula_and_grad():
    RNG.setSeed(1)
    m = L.Linear(5, 3)
    x = torch.randn(4, 5)
    y = m.forward(x)
    assert torch.allclose(y, x @ m.weight.detach().t() + m.bias.detach(), atol=1e-6)
    grad_check(m, torch.randn(2, 5))
    assert abs(float(m.weight.detach().abs().max())) <= 1 / math.sqrt(5) + 1e-6


def test_conv_same_padding_and_groups():
    RNG.setSeed(2)
    m = L.SpatialConvolution(4, 6, 3, 3, 2, 2, -1, -1, 2)
    x = torch.randn(2, 4, 9, 9)
    y = m.forward(x)
    assert y.shape == (2, 6, 5, 5)           # ceil(9/2)
    grad_check(L.SpatialConvolution(2, 2, 3, 3, 1, 1, 1, 1), torch.randn(1, 2, 4, 4))


def test_pooling_ceil_mode_sizes():
    x = torch.randn(1, 3, 112, 112)
    assert L.SpatialMaxPooling(3, 3, 2, 2).ceil().forward(x).shape == (1, 3, 56, 56)
    assert L.SpatialMaxPooling(3, 3, 2, 2).forward(x).shape == (1, 3, 55, 55)
    assert L.SpatialMaxPooling(3, 3, 1, 1, 1, 1).ceil().forward(x).shape == (1, 3, 112, 112)
    assert L.SpatialAveragePooling(7, 7, 1, 1).forward(torch.randn(2, 8, 7, 7)).shape == (2, 8, 1, 1)
    a = L.SpatialAveragePooling(2, 2, 2, 2).forward(torch.arange(16.).reshape(1, 1, 4, 4))
    assert torch.allclose(a.reshape(-1), torch.tensor([2.5, 4.5, 10.5, 12.5]))


def test_lrn_matches_torch():
    x = torch.randn(2, 8, 5, 5) * 3
    y = L.SpatialCrossMapLRN(5, 1e-4, 0.75, 1.0).forward(x)
    assert torch.allclose(y, F.local_response_norm(x, 5, 1e-4, 0.75, 1.0), atol=1e-6)


def test_batchnorm_running_stats_use_unbiased_variance():
    RNG.setSeed(3)
    bn = L.BatchNormalization(4)
    x = torch.randn(16, 4) * 2 + 1
    bn.forward(x)
    assert torch.allclose(bn.runningMean, 0.1 * x.mean(0), atol=1e-5)
    assert torch.allclose(bn.runningVar, 0.9 + 0.1 * x.var(0, unbiased=True), atol=1e-5)
    bn.evaluate()
    y = bn.forward(x)
    ref = (x - bn.runningMean) / torch.sqrt(bn.runningVar + 1e-5) * bn.weight.detach() + bn.bias.detach()
    assert torch.allclose(y, ref, atol=1e-5)


def test_dropout_scaling_and_eval():
    RNG.setSeed(4)
    d = L.Dropout(0.3)
    x = torch.ones(20000)
    y = d.forward(x)
    assert abs(float((y == 0).float().mean()) - 0.3) < 0.02
    assert abs(float(y[y != 0][0]) - 1 / 0.7) < 1e-5
    d.evaluate()
    assert torch.equal(d.forward(x), x)


def test_table_ops_and_containers():
    a, b = torch.randn(3, 4), torch.randn(3, 4)
    assert torch.allclose(L.CAddTable().forward(T(a, b)), a + b)
    assert torch.allclose(L.CSubTable().forward(T(a, b)), a - b)
    assert torch.allclose(L.CMulTable().forward(T(a, b)), a * b)
    assert torch.allclose(L.CMaxTable().forward(T(a, b)), torch.maximum(a, b))
    assert torch.allclose(L.JoinTable(2).forward(T(a, b)), torch.cat([a, b], 1))
    parts = L.SplitTable(1).forward(a)
    assert len(parts.to_list()) == 3 and torch.allclose(parts[2], a[1])
    assert torch.allclose(L.SelectTable(2).forward(T(a, b)), b)
    assert torch.allclose(L.Narrow(2, 2, 2).forward(a), a[:, 1:3])
    assert torch.allclose(L.Select(1, 2).forward(a), a[1])
    assert L.Reshape([2, 2]).forward(a).shape == (3, 2, 2)
    assert L.View(12).forward(a).shape == (12,)
    assert L.Transpose([(1, 2)]).forward(a).shape == (4, 3)
    assert L.Squeeze(1).forward(torch.zeros(1, 3)).shape == (3,)
    assert L.Unsqueeze(1).forward(a).shape == (1, 3, 4)
    assert L.Replicate(5, 1).forward(a).shape == (5, 3, 4)
    assert torch.allclose(L.Sum(2).forward(a), a.sum(1))
    assert torch.allclose(L.Mean(1).forward(a), a.mean(0))
    assert torch.allclose(L.Max(2).forward(a), a.max(1).values)
    assert torch.allclose(L.MM().forward(T(a, b.t())), a @ b.t())
    assert torch.allclose(L.DotProduct().forward(T(a, b)), (a * b).sum(1))
    ct = L.ConcatTable().add(L.Identity()).add(L.MulConstant(2.0))
    out = ct.forward(a)
    assert torch.allclose(out[2], 2 * a)
    pt = L.ParallelTable().add(L.AddConstant(1.0)).add(L.Negative())
    out = pt.forward(T(a, b))
    assert torch.allclose(out[1], a + 1) and torch.allclose(out[2], -b)
    cc = L.Concat(2).add(L.Identity()).add(L.Identity())
    assert cc.forward(a).shape == (3, 8)


def test_activations_match_torch():
    x = torch.randn(4, 6)
    pairs = [(L.ReLU(), F.relu(x)), (L.ReLU6(), F.relu6(x)), (L.Tanh(), torch.tanh(x)), (L.Sigmoid(), torch.sigmoid(x)),
             (L.ELU(0.5), F.elu(x, 0.5)), (L.LeakyReLU(0.1), F.leaky_relu(x, 0.1)), (L.SoftPlus(), F.softplus(x)),
             (L.SoftSign(), F.softsign(x)), (L.HardTanh(), F.hardtanh(x)), (L.LogSigmoid(), F.logsigmoid(x)),
             (L.SoftShrink(0.5), F.softshrink(x, 0.5)), (L.HardShrink(0.5), F.hardshrink(x, 0.5)),
             (L.Abs(), x.abs()), (L.Square(), x * x), (L.Exp(), x.exp()), (L.TanhShrink(), x - torch.tanh(x)),
             (L.HardSigmoid(), torch.clamp(0.2 * x + 0.5, 0, 1)), (L.Power(2, 2.0, 1.0), (1 + 2 * x) ** 2),
             (L.Threshold(0.1, -1.0), torch.where(x > 0.1, x, torch.full_like(x, -1.0))),
             (L.SoftMax(), F.softmax(x, 1)), (L.LogSoftMax(), F.log_softmax(x, 1)), (L.SoftMin(), F.softmin(x, 1))]
    for m, ref in pairs:
        assert torch.allclose(m.forward(x), ref, atol=1e-6), type(m).__name__
    xp = torch.randn(2, 6)
    xp[xp.abs() < 0.05] = 0.1              # keep the finite-difference step away from the kink at 0
    grad_check(L.PReLU(6), xp)
    g = L.GradientReversal(2.0)
    g.forward(x)
    assert torch.allclose(g.backward(x, torch.ones_like(x)), -2 * torch.ones_like(x))


def test_lookup_table_is_one_based():
    RNG.setSeed(5)
    lt = L.LookupTable(10, 4)
    idx = torch.tensor([[1, 10], [3, 3]]).float()
    y = lt.forward(idx)
    assert torch.allclose(y[0, 0], lt.weight.detach()[0]) and torch.allclose(y[0, 1], lt.weight.detach()[9])


def test_recurrent_lstm_gru_shapes_and_state():
    RNG.setSeed(6)
    x = torch.randn(3, 5, 4)
    for cell, h in ((L.LSTM(4, 6), 6), (L.GRU(4, 6), 6), (L.RnnCell(4, 6, L.Tanh()), 6), (L.LSTMPeephole(4, 6), 6)):
        r = L.Recurrent().add(cell)
        y = r.forward(x)
        assert y.shape == (3, 5, h)
        gx = r.backward(x, torch.ones_like(y))
        assert gx.shape == x.shape and float(gx.abs().sum()) > 0
    # LSTM against a hand-rolled reference with the same gate order [in, cand, forget, out]
    cell = L.LSTM(4, 6)
    r = L.Recurrent().add(cell)
    y = r.forward(x)
    Wi, bi, Wh = cell.i2g.weight.detach(), cell.i2g.bias.detach(), cell.h2g.weight.detach()
    h = torch.zeros(3, 6)
    c = torch.zeros(3, 6)
    for t in range(5):
        g = x[:, t] @ Wi.t() + bi + h @ Wh.t()
        i, z, f, o = torch.sigmoid(g[:, :6]), torch.tanh(g[:, 6:12]), torch.sigmoid(g[:, 12:18]), torch.sigmoid(g[:, 18:])
        c = i * z + f * c
        h = torch.tanh(c) * o
        assert torch.allclose(y[:, t], h, atol=1e-5)
    bi_r = L.BiRecurrent(L.JoinTable(3)).add(L.GRU(4, 5))
    assert bi_r.forward(x).shape == (3, 5, 10)
    td = L.TimeDistributed(L.Linear(4, 2))
    assert td.forward(x).shape == (3, 5, 2)
    conv = L.Recurrent().add(L.ConvLSTMPeephole(2, 3, 3, 3))
    assert conv.forward(torch.randn(2, 4, 2, 6, 6)).shape == (2, 4, 3, 6, 6)


def test_attention_and_transformer():
    RNG.setSeed(7)
    torch.manual_seed(7)
    att = L.Attention(8, 2, 0.0)
    x = torch.randn(2, 5, 8)
    y = att.forward(T(x, x, L.causal_bias(5, x.device)))
    assert y.shape == (2, 5, 8)
    tr = L.Transformer(20, 8, 2, 16, 2, 0.0, 0.0, 0.0, with_share_weights_linear=True)
    ids = torch.randint(1, 21, (2, 6)).float()
    out = tr.forward(ids)
    assert out.shape == (2, 6, 20)
    # causality: changing a later token does not change earlier outputs
    ids2 = ids.clone()
    ids2[:, 5] = (ids2[:, 5] % 20) + 1
    out2 = tr.forward(ids2)
    assert torch.allclose(out[:, :5], out2[:, :5], atol=1e-5)
    bs = L.SequenceBeamSearch(20, 3, 0.6, 5, 2, 0, 2, 8)
    tt = L.Transformer(20, 8, 2, 16, 1, 0.0, 0.0, 0.0, with_share_weights_linear=True,
                       transformer_type=L.TransformerType.Translation, beam_search=bs)
    tt.evaluate()
    res = tt.forward(T(ids, None))
    seqs, scores = res[1], res[2]
    assert seqs.shape[0] == 2 and seqs.shape[1] == 3 and scores.shape == (2, 3)


def test_detection_helpers():
    boxes = torch.tensor([[0., 0, 10, 10], [1, 1, 11, 11], [50, 50, 60, 60]])
    keep = L.Nms(0.5).forward(T(boxes, torch.tensor([0.9, 0.8, 0.7])))
    assert keep.tolist() == [1, 3]
    anc = L.Anchor([0.5, 1.0, 2.0], [8.0, 16.0, 32.0])
    a = anc.generateAnchors(4, 3, 16)
    assert a.shape == (4 * 3 * 9, 4)
    assert torch.allclose(anc.basic[0], torch.tensor([-84., -40., 99., 55.]))    # classic py-faster-rcnn anchor
    pb = L.PriorBox([30.0], [60.0], [2.0], img_size=300, step=8.0)
    out = pb.forward(torch.zeros(1, 8, 38, 38))
    assert out.shape == (1, 2, 38 * 38 * 4 * 4)
    fpn = L.FPN([8, 16, 32], 8)
    feats = T(torch.randn(1, 8, 32, 32), torch.randn(1, 16, 16, 16), torch.randn(1, 32, 8, 8))
    ps = fpn.forward(feats)
    assert [tuple(p.shape[-2:]) for p in ps.to_list()] == [(32, 32), (16, 16), (8, 8)]
    ra = L.RoiAlign(1.0, 2, 2, 2)
    f = torch.arange(16.).reshape(1, 1, 4, 4)
    r = ra.forward(T(f, torch.tensor([[0., 0., 4., 4.]])))
    assert r.shape == (1, 1, 2, 2) and abs(float(r.mean()) - 7.5) < 0.6
    rp = L.RoiPooling(2, 2, 1.0)
    r = rp.forward(T(f, torch.tensor([[0., 0., 0., 3., 3.]])))
    assert r.reshape(-1).tolist() == [5., 7., 13., 15.]


def test_misc_layers_smoke():
    RNG.setSeed(8)
    x = torch.randn(4, 6)
    assert L.Bilinear(6, 5, 3).forward(T(x, torch.randn(4, 5))).shape == (4, 3)
    assert L.Cosine(6, 3).forward(x).shape == (4, 3)
    assert L.Euclidean(6, 3).forward(x).shape == (4, 3)
    assert L.Maxout(6, 3, 4).forward(x).shape == (4, 3)
    assert L.Highway(6).forward(x).shape == (4, 6)
    assert L.CMul([1, 6]).forward(x).shape == (4, 6) and L.CAdd([1, 6]).forward(x).shape == (4, 6)
    assert L.Scale([1, 6]).forward(x).shape == (4, 6)
    assert torch.allclose(L.Normalize(2).forward(x).norm(dim=1), torch.ones(4), atol=1e-4)
    ln = L.LayerNormalization(6).forward(x)
    assert torch.allclose(ln.mean(1), torch.zeros(4), atol=1e-5)
    img = torch.randn(2, 3, 8, 8)
    assert L.SpatialFullConvolution(3, 5, 3, 3, 2, 2, 1, 1, 1, 1).forward(img).shape == (2, 5, 16, 16)
    assert L.SpatialSeparableConvolution(3, 6, 2, 3, 3, 1, 1, 1, 1).forward(img).shape == (2, 6, 8, 8)
    assert L.SpatialDilatedConvolution(3, 4, 3, 3, 1, 1, 2, 2, 2, 2).forward(img).shape == (2, 4, 8, 8)
    assert L.TemporalConvolution(6, 4, 3).forward(torch.randn(2, 10, 6)).shape == (2, 8, 4)
    assert L.VolumetricConvolution(2, 3, 3, 3, 3).forward(torch.randn(1, 2, 5, 6, 7)).shape == (1, 3, 3, 4, 5)
    assert L.LocallyConnected2D(3, 8, 8, 4, 3, 3).forward(img).shape == (2, 4, 6, 6)
    assert L.LocallyConnected1D(10, 6, 4, 3).forward(torch.randn(2, 10, 6)).shape == (2, 8, 4)
    assert L.UpSampling2D([2, 2]).forward(img).shape == (2, 3, 16, 16)
    assert L.ResizeBilinear(4, 4).forward(img).shape == (2, 3, 4, 4)
    assert L.SpatialZeroPadding(1, 1, 2, 2).forward(img).shape == (2, 3, 12, 10)
    assert L.Cropping2D([1, 1], [2, 2]).forward(img).shape == (2, 3, 6, 4)
    assert L.TemporalMaxPooling(2).forward(torch.randn(2, 10, 6)).shape == (2, 5, 6)
    assert L.VolumetricMaxPooling(2, 2, 2).forward(torch.randn(1, 2, 4, 4, 4)).shape == (1, 2, 2, 2, 2)
    assert L.SpatialSubtractiveNormalization(3).forward(img).shape == img.shape
    assert L.SpatialDivisiveNormalization(3).forward(img).shape == img.shape
    assert L.SpatialWithinChannelLRN(3).forward(img).shape == img.shape
    sp = L.SparseLinear(6, 3)
    assert sp.forward(x.to_sparse()).shape == (4, 3)
    lts = L.LookupTableSparse(10, 4, "mean")
    ids = torch.tensor([[1., 3., 0.], [2., 0., 0.]]).to_sparse()
    assert lts.forward(ids).shape == (2, 4)
    m = L.SReLU([6])
    assert m.forward(x).shape == x.shape
    l1 = L.L1Penalty(1)
    l1.forward(x)
    assert torch.allclose(l1.backward(x, torch.zeros_like(x)), torch.sign(x))
    tree = L.BinaryTreeLSTM(5, 4)
    emb = torch.randn(1, 3, 5)
    trees = torch.tensor([[[2, 3, -1], [0, 0, 1], [4, 5, 0], [0, 0, 2], [0, 0, 3]]]).float()
    assert tree.forward(T(emb, trees)).shape == (1, 5, 4)


def test_inception_v2_and_maskrcnn_models():
    """Model-zoo additions: BN-Inception (S/models/inception/Inception_v2.scala) and Mask R-CNN inference
    (S/models/maskrcnn/MaskRCNN.scala)."""
    import torch

    from bigdl_b200.models.inception_v2 import Inception_v2, Inception_v2_NoAuxClassifier
    from bigdl_b200.models.maskrcnn import MaskRCNN, MaskRCNNParams
    from bigdl_b200.utils.table import Table

    m = Inception_v2_NoAuxClassifier(20).evaluate()
    assert tuple(m.forward(torch.randn(1, 3, 224, 224)).shape) == (1, 20)
    assert tuple(Inception_v2(7).evaluate().forward(torch.randn(1, 3, 224, 224)).shape) == (1, 21)
    torch.manual_seed(0)
    p = MaskRCNNParams(preNmsTopNTest=50, postNmsTopNTest=20, maxPerImage=5, boxScoreThresh=0.0)
    det = MaskRCNN(64, 32, 5, p).evaluate()
    out = det.forward(Table(torch.randn(1, 3, 128, 160), torch.tensor([[128.0, 160.0, 256.0, 320.0]])))
    r = out[1]
    n = r["bboxes"].shape[0]
    assert 0 < n <= 5 and r["classes"].shape[0] == n and len(r["masks"]) == n
    assert r["masks"][0].height == 256 and r["masks"][0].width == 320      # decoded at the ORIGINAL image size
