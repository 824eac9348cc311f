"""Criterions vs torch references; OptimMethods on Rosenbrock-style problems and schedule value sequences
(T/nn/*CriterionSpec, T/optim/{SGD,Adam,Adagrad,LBFGS,...}Spec.scala)."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from bigdl_b200.nn import criterion as C
from bigdl_b200.nn import layer as L
from bigdl_b200.optim import optim_method as O
from bigdl_b200.optim.trigger import (EveryEpoch, MaxEpoch, MaxIteration, MinLoss, SeveralIteration, Trigger)
from bigdl_b200.utils.table import T


def test_classnll_one_based_weights_padding():
    logp = F.log_softmax(torch.randn(5, 4), 1)
    tgt = torch.tensor([1., 4., 2., 2., 3.])
    assert abs(C.ClassNLLCriterion().forward(logp, tgt) - float(F.nll_loss(logp, tgt.long() - 1))) < 1e-6
    w = torch.tensor([1., 2., 3., 4.])
    ref = F.nll_loss(logp, tgt.long() - 1, weight=w)
    assert abs(C.ClassNLLCriterion(w).forward(logp, tgt) - float(ref)) < 1e-6
    pad = C.ClassNLLCriterion(paddingValue=0)
    t2 = torch.tensor([1., 0., 2., 0., 3.])
    keep = torch.tensor([0, 2, 4])
    assert abs(pad.forward(logp, t2) - float(F.nll_loss(logp[keep], t2[keep].long() - 1))) < 1e-6
    with pytest.raises(ValueError):
        C.ClassNLLCriterion().forward(logp, torch.tensor([1., 5., 1., 1., 1.]))
    g = C.ClassNLLCriterion().backward(logp, tgt)
    assert abs(float(g[0, 0]) + 0.2) < 1e-6 and float(g[0, 1]) == 0
    probs = logp.exp()
    assert abs(C.ClassNLLCriterion(logProbAsInput=False).forward(probs, tgt) -
               float(F.nll_loss(logp, tgt.long() - 1))) < 1e-5


def test_criterions_against_torch():
    x, t = torch.rand(6, 3) * 0.8 + 0.1, torch.rand(6, 3)
    assert abs(C.MSECriterion().forward(x, t) - float(F.mse_loss(x, t))) < 1e-6
    assert abs(C.AbsCriterion().forward(x, t) - float(F.l1_loss(x, t))) < 1e-6
    assert abs(C.BCECriterion().forward(x, (t > 0.5).float()) - float(F.binary_cross_entropy(x, (t > 0.5).float()))) < 1e-5
    assert abs(C.SmoothL1Criterion().forward(x, t) - float(F.smooth_l1_loss(x, t))) < 1e-6
    y = torch.sign(torch.randn(6, 3))
    assert abs(C.SoftMarginCriterion().forward(x, y) - float(F.soft_margin_loss(x, y))) < 1e-6
    assert abs(C.MarginCriterion().forward(x, y) - float(torch.clamp(1 - x * y, min=0).mean())) < 1e-6
    assert abs(C.HingeEmbeddingCriterion(1.0).forward(x, y) - float(F.hinge_embedding_loss(x, y))) < 1e-6
    logits = torch.randn(6, 5)
    cls = torch.randint(1, 6, (6,)).float()
    assert abs(C.CrossEntropyCriterion().forward(logits, cls) - float(F.cross_entropy(logits, cls.long() - 1))) < 1e-6
    assert abs(C.MultiMarginCriterion().forward(logits, cls) - float(F.multi_margin_loss(logits, cls.long() - 1))) < 1e-6
    lp = F.log_softmax(logits, 1)
    q = F.softmax(torch.randn(6, 5), 1)
    assert abs(C.DistKLDivCriterion().forward(lp, q) - float(F.kl_div(lp, q, reduction="mean"))) < 1e-6
    a, b = torch.randn(6, 4), torch.randn(6, 4)
    yy = torch.sign(torch.randn(6))
    assert abs(C.CosineEmbeddingCriterion(0.1).forward(T(a, b), yy) -
               float(F.cosine_embedding_loss(a, b, yy, margin=0.1))) < 1e-5
    assert abs(C.MarginRankingCriterion(0.5).forward(T(a[:, 0], b[:, 0]), yy) -
               float(F.margin_ranking_loss(a[:, 0], b[:, 0], yy, margin=0.5))) < 1e-6
    pc = C.ParallelCriterion().add(C.MSECriterion(), 0.5).add(C.AbsCriterion(), 2.0)
    assert abs(pc.forward(T(x, x), T(t, t)) - (0.5 * float(F.mse_loss(x, t)) + 2 * float(F.l1_loss(x, t)))) < 1e-5
    mc = C.MultiCriterion().add(C.MSECriterion()).add(C.AbsCriterion())
    assert abs(mc.forward(x, t) - float(F.mse_loss(x, t) + F.l1_loss(x, t))) < 1e-5
    seq = F.log_softmax(torch.randn(2, 4, 5), 2)
    st = torch.randint(1, 6, (2, 4)).float()
    td = C.TimeDistributedCriterion(C.ClassNLLCriterion(), True)
    ref = np.mean([float(F.nll_loss(seq[:, i], st[:, i].long() - 1)) for i in range(4)])
    assert abs(td.forward(seq, st) - ref) < 1e-5
    g = C.MSECriterion().backward(x, t)
    assert torch.allclose(g, 2 * (x - t) / x.numel(), atol=1e-6)
    for crit, inp, tg in [(C.L1Cost(), x, None), (C.KLDCriterion(), T(a, b), None), (C.PoissonCriterion(), x, t),
                          (C.MeanAbsolutePercentageCriterion(), x, t + 0.1), (C.MeanSquaredLogarithmicCriterion(), x, t),
                          (C.CosineProximityCriterion(), a, b), (C.DiceCoefficientCriterion(), x, (t > 0.5).float()),
                          (C.CategoricalCrossEntropy(), F.softmax(logits, 1), F.one_hot(cls.long() - 1, 5).float()),
                          (C.ClassSimplexCriterion(5), torch.randn(6, 5), cls), (C.GaussianCriterion(), T(a, b), a),
                          (C.DotProductCriterion(), x, t), (C.PGCriterion(), x, t),
                          (C.KullbackLeiblerDivergenceCriterion(), q, F.softmax(logits, 1)),
                          (C.MultiLabelSoftMarginCriterion(), logits, (torch.rand(6, 5) > 0.5).float()),
                          (C.CosineDistanceCriterion(), a, b), (C.SoftmaxWithCriterion(), logits, cls)]:
        v = crit.forward(inp, tg)
        assert math.isfinite(v), type(crit).__name__


def _rosenbrock(x):
    a, b = x[:-1], x[1:]
    f = float((100 * (b - a * a) ** 2 + (1 - a) ** 2).sum())
    g = torch.zeros_like(x)
    g[:-1] += -400 * a * (b - a * a) - 2 * (1 - a)
    g[1:] += 200 * (b - a * a)
    return f, g


@pytest.mark.parametrize("make,iters,tol", [
    (lambda: O.Adam(0.05), 2500, 1e-2), (lambda: O.Adagrad(0.5), 4000, 0.2), (lambda: O.RMSprop(0.01), 4000, 0.2),
    (lambda: O.Adamax(0.05), 4000, 0.1), (lambda: O.SGD(1e-3, momentum=0.9, dampening=0.0), 6000, 0.05),
    (lambda: O.Adadelta(0.9, 1e-6), 6000, 1.0),
])
def test_optim_methods_minimise_rosenbrock(make, iters, tol):
    x = torch.tensor([-1.2, 1.0])
    om = make()
    f0 = _rosenbrock(x)[0]
    for _ in range(iters):
        om.optimize(_rosenbrock, x)
    assert _rosenbrock(x)[0] < min(tol, f0)


def test_lbfgs_rosenbrock():
    x = torch.tensor([-1.2, 1.0, -1.2, 1.0])
    om = O.LBFGS(max_iter=200, max_eval=400, tolfun=1e-10, tolx=1e-12, learningrate=0.2)
    for _ in range(10):
        _, hist = om.optimize(_rosenbrock, x)
    assert _rosenbrock(x)[0] < 1e-3


def test_sgd_semantics_and_negative_rate():
    x = torch.tensor([1.0, 2.0])
    om = O.SGD(0.1, learningrate_decay=0.5, weightdecay=0.1, momentum=0.9)
    om.update(x, torch.tensor([0.5, -0.5]))
    # first step: v = g + wd*x ; x -= clr*v with clr = 0.1/(1+0*0.5)
    assert torch.allclose(x, torch.tensor([1.0 - 0.1 * 0.6, 2.0 - 0.1 * (-0.3)]))
    assert om.getLearningRate() == -0.1
    om.update(x, torch.zeros(2))
    assert abs(om.getLearningRate() + 0.1 / 1.5) < 1e-9            # evalCounter advanced
    assert om.dampening == 0.9                                     # dampening defaults to momentum
    with pytest.raises(ValueError):
        O.SGD(0.1, momentum=0.9, nesterov=True)


def test_schedules_value_sequences():
    def rates(schedule, n, lr=1.0, epochs=None):
        om = O.SGD(lr, learningrate_schedule=schedule)
        out = []
        for i in range(n):
            if epochs:
                om.state["epoch"] = epochs[i]
            schedule.updateHyperParameter(om)
            out.append(-schedule.currentRate)
        return out
    assert rates(O.Step(2, 0.1), 5) == pytest.approx([1, 1, 0.1, 0.1, 0.01])
    assert rates(O.MultiStep([1, 3], 0.5), 5) == pytest.approx([1, 0.5, 0.5, 0.25, 0.25])
    assert rates(O.Poly(2.0, 4), 6) == pytest.approx([1, 0.5625, 0.25, 0.0625, 0, 0])
    assert rates(O.Exponential(2, 0.5), 3) == pytest.approx([1, 0.5 ** 0.5, 0.5])
    assert rates(O.Exponential(2, 0.5, True), 4) == pytest.approx([1, 1, 0.5, 0.5])
    assert rates(O.NaturalExp(1, 1.0), 3) == pytest.approx([1, math.exp(-1), math.exp(-2)])
    assert rates(O.Warmup(0.1), 3, lr=0.5) == pytest.approx([0.5, 0.6, 0.7])
    assert rates(O.EpochStep(2, 0.1), 4, epochs=[1, 2, 3, 4]) == pytest.approx([1, 0.1, 0.1, 0.01])
    assert rates(O.EpochDecay(lambda e: e - 1), 3, epochs=[1, 2, 3]) == pytest.approx([1, 0.1, 0.01])
    seq = O.SequentialSchedule(10).add(O.Warmup(0.1), 3).add(O.Poly(1.0, 10), 6)
    r = rates(seq, 7, lr=0.1)
    assert r[:4] == pytest.approx([0.1, 0.2, 0.3, 0.4])
    # Poly counts raw iterations (no exclusion, S/optim/SGD.scala:307) from the rate reached by the warm-up
    assert r[4:] == pytest.approx([0.4 * 0.6, 0.4 * 0.5, 0.4 * 0.4])
    ew = O.EpochDecayWithWarmUp(2, 0.1, lambda e: 0 if e < 3 else 1)
    assert rates(ew, 4, lr=0.1, epochs=[1, 1, 2, 3]) == pytest.approx([0.1, 0.2, 0.3, 0.03])
    pl = O.Plateau("Loss", factor=0.5, patience=1, epsilon=0.0)
    om = O.SGD(1.0, learningrate_schedule=pl)
    seen = []
    for epoch, loss in enumerate([1.0, 1.0, 1.0, 1.0, 1.0], start=1):
        om.state["epoch"], om.state["Loss"] = epoch, loss
        pl.updateHyperParameter(om)
        seen.append(-pl.currentRate)
    assert seen[0] == 1.0 and seen[-1] < 1.0
    es = O.EpochSchedule([O.Regime(1, 2, {"learningRate": 0.5}), O.Regime(3, 4, {"learningRate": 0.05, "weightDecay": 1e-3})])
    om = O.SGD(1.0, learningrate_schedule=es)
    om.state["epoch"] = 3
    es.updateHyperParameter(om)
    assert om.learningRate == 0.05 and om.weightDecay == 1e-3 and es.currentRate == -0.05


def test_triggers():
    st = T()
    st["epoch"], st["neval"], st["Loss"] = 1, 1, 10.0
    assert not MaxEpoch(1)(st) and not MaxIteration(1)(st)
    st["neval"] = 2
    assert MaxIteration(1)(st)
    st["epoch"] = 2
    assert MaxEpoch(1)(st)
    ee = EveryEpoch()
    st["epoch"] = 1
    assert not ee(st)
    st["epoch"] = 2
    assert ee(st) and not ee(st)
    assert SeveralIteration(2)(st) and not SeveralIteration(3)(st)
    st["Loss"] = 0.01
    assert MinLoss(0.1)(st)
    assert Trigger.and_(MinLoss(0.1), MaxEpoch(1))(st) and Trigger.or_(MinLoss(0.001), MaxEpoch(1))(st)


def test_regularizers_and_freeze_scale():
    from bigdl_b200.optim.regularizer import L1L2Regularizer, L2Regularizer
    from bigdl_b200.utils.random_generator import RNG

    RNG.setSeed(9)
    m = L.Linear(3, 2, wRegularizer=L2Regularizer(0.5))
    x = torch.randn(4, 3)
    m.forward(x)
    m.backward(x, torch.zeros(4, 2))
    assert torch.allclose(m.gradWeight, 0.5 * m.weight.detach(), atol=1e-6)
    m2 = L.Linear(3, 2)
    m2.setScaleW(2.0).setScaleB(0.0)
    m2.forward(x)
    m2.backward(x, torch.ones(4, 2))
    assert torch.allclose(m2.gradWeight, 2 * torch.ones(2, 4) @ x, atol=1e-5)
    assert float(m2.gradBias.abs().sum()) == 0
    seq = L.Sequential().add(L.Linear(3, 3).setName("a")).add(L.Linear(3, 2).setName("b"))
    seq.freeze("a")
    seq.forward(x)
    seq.backward(x, torch.ones(4, 2))
    assert float(seq.find("a").gradWeight.abs().sum()) == 0 and float(seq.find("b").gradWeight.abs().sum()) > 0
    seq.unFreeze()
    w, g = seq.getParameters()
    assert w.numel() % 8 == 0 and g.shape == w.shape
    with pytest.raises(ValueError):
        shared = L.Linear(2, 2)
        L.Sequential().add(shared).add(shared).checkDuplicate()
