"""Validation methods, triggers and initialisation methods that no other test exercises (S/optim/ValidationMethod.scala
MAE :1103, HitRatio :882, NDCG :949, PrecisionRecallAUC; S/optim/Trigger.scala; S/nn/InitializationMethod.scala)."""
import math

import torch

from bigdl_b200.nn.initialization_method import BilinearFiller, MsraFiller, Ones, RandomNormal, Xavier
from bigdl_b200.nn.layer import Linear, SpatialConvolution, SpatialFullConvolution
from bigdl_b200.optim import validation as V
from bigdl_b200.optim.optim_method import Adam, ParallelAdam
from bigdl_b200.optim.trigger import MaxScore, MinLoss, Trigger, TriggerAnd, TriggerOr
from bigdl_b200.utils.random_generator import RNG
from bigdl_b200.utils.table import T


def test_mae_hitratio_ndcg_and_pr_auc():
    mae = V.MAE()(torch.tensor([[1.0], [2.0], [4.0]]), torch.tensor([[1.5], [2.0], [2.0]]))
    assert abs(mae.result()[0] - (0.5 + 0.0 + 2.0) / 3) < 1e-6
    # two users, 1 positive + 3 negatives each; the positive is ranked 1st for user A and 3rd for user B
    scores = torch.tensor([0.9, 0.1, 0.2, 0.3, 0.4, 0.8, 0.7, 0.1])
    labels = torch.tensor([1.0, 0, 0, 0, 1.0, 0, 0, 0])
    hr2 = V.HitRatio(2, 3)(scores, labels)
    assert abs(hr2.result()[0] - 0.5) < 1e-6 and hr2.result()[1] == 2
    assert abs(V.HitRatio(3, 3)(scores, labels).result()[0] - 1.0) < 1e-6
    ndcg = V.NDCG(3, 3)(scores, labels).result()[0]
    assert abs(ndcg - (1.0 + math.log(2) / math.log(4)) / 2) < 1e-5       # ranks 1 and 3 -> 1/log2(2), 1/log2(4)
    merged = V.HitRatio(2, 3)(scores[:4], labels[:4]) + V.HitRatio(2, 3)(scores[4:], labels[4:])
    assert abs(merged.result()[0] - 0.5) < 1e-6
    perfect = V.PrecisionRecallAUC()(torch.tensor([0.9, 0.8, 0.2, 0.1]), torch.tensor([1.0, 1.0, 0.0, 0.0]))
    assert abs(perfect.result()[0] - 1.0) < 1e-6
    worst = V.PrecisionRecallAUC()(torch.tensor([0.9, 0.8, 0.2, 0.1]), torch.tensor([0.0, 0.0, 1.0, 1.0]))
    assert worst.result()[0] < 0.6
    both = V.PrecisionRecallAUC()(torch.tensor([0.9, 0.8]), torch.tensor([1.0, 1.0])) + \
        V.PrecisionRecallAUC()(torch.tensor([0.2, 0.1]), torch.tensor([0.0, 0.0]))
    assert abs(both.result()[0] - 1.0) < 1e-6


def test_triggers_compose():
    s = T(epoch=3, neval=120, Loss=0.4, score=0.91)
    assert MaxScore(0.9)(s) and not MaxScore(0.95)(s)
    assert MinLoss(0.5)(s) and not MinLoss(0.3)(s)
    assert TriggerAnd(MaxScore(0.9), MinLoss(0.5))(s) and not TriggerAnd(MaxScore(0.9), MinLoss(0.3))(s)
    assert TriggerOr(MaxScore(0.95), MinLoss(0.5))(s) and not TriggerOr(MaxScore(0.95), MinLoss(0.3))(s)
    assert TriggerAnd(Trigger.maxIteration(100), Trigger.maxEpoch(2), MinLoss(0.5))(s)   # maxEpoch fires once epoch > max


def test_initialisation_methods_statistics():
    RNG.setSeed(11)
    lin = Linear(400, 300)
    lin.setInitMethod(RandomNormal(1.0, 0.5), Ones())
    assert abs(float(lin.weight.mean()) - 1.0) < 0.02 and abs(float(lin.weight.std()) - 0.5) < 0.02
    assert torch.all(lin.bias == 1)
    lin.setInitMethod(Xavier())
    bound = math.sqrt(6.0 / (400 + 300))
    assert float(lin.weight.abs().max()) <= bound + 1e-6 and float(lin.weight.abs().max()) > 0.9 * bound
    conv = SpatialConvolution(16, 32, 3, 3)
    conv.setInitMethod(MsraFiller(False))
    assert abs(float(conv.weight.std()) - math.sqrt(2.0 / (16 * 9))) < 0.01        # fan-in variant
    up = SpatialFullConvolution(3, 3, 4, 4, 2, 2, 1, 1)
    up.setInitMethod(BilinearFiller())
    k = up.weight[0, 0]
    assert torch.allclose(k, k.flip(0)) and torch.allclose(k, k.flip(1)) and float(k.max()) <= 1.0    # symmetric tent
    assert abs(float(k[1, 1]) - 0.5625) < 1e-6


def test_parallel_adam_matches_adam():
    torch.manual_seed(0)
    w1, w2 = torch.randn(1000), None
    w2 = w1.clone()
    a, p = Adam(1e-2), ParallelAdam(1e-2, parallel_num=4)
    for _ in range(3):
        g = torch.randn(1000)
        a.update(w1, g.clone())
        p.update(w2, g.clone())
    assert torch.allclose(w1, w2, atol=1e-6)
