"""Config #1 of BASELINE.json: LeNet-5 on synthetic MNIST-shape data, Optimizer in local mode on CPU."""
import numpy as np
import torch

from bigdl_b200.dataset.base import DataSet, Sample
from bigdl_b200.models.lenet import LeNet5, LeNet5_graph
from bigdl_b200.nn.criterion import ClassNLLCriterion
from bigdl_b200.optim.optim_method import SGD
from bigdl_b200.optim.optimizer import LocalOptimizer, Optimizer
from bigdl_b200.optim.trigger import MaxEpoch, MaxIteration
from bigdl_b200.optim.validation import Top1Accuracy
from bigdl_b200.utils.random_generator import RNG


def _data(n=256, seed=0):
    rs = np.random.RandomState(seed)
    y = rs.randint(1, 11, size=n)
    x = rs.randn(n, 28, 28).astype(np.float32) * 0.1
    for i in range(n):       # class-dependent blob so the task is learnable
        c = y[i] - 1
        x[i, 2 * c:2 * c + 6, 4:24] += 1.0
    return [Sample(x[i], np.array([y[i]], dtype=np.float32)) for i in range(n)]


def test_lenet_param_count():
    m = LeNet5(10)
    w, g = m.getParameters()
    assert sum(p.numel() for p in m.torch_parameters()) == 22278
    assert w.numel() >= 22278 and w.numel() % 8 == 0


def test_lenet_trains_and_validates():
    RNG.setSeed(1)
    torch.manual_seed(1)
    model = LeNet5(10)
    train = DataSet.array(_data(256, 0))
    opt = Optimizer(model, train, ClassNLLCriterion(), MaxEpoch(6), 32, SGD(learningrate=0.05, momentum=0.9,
                                                                             dampening=0.0))
    assert isinstance(opt, LocalOptimizer)
    opt.setValidation(MaxIteration(1), DataSet.array(_data(64, 1)), [Top1Accuracy()], 32)
    opt.optimize()
    res = model.evaluate(DataSet.array(_data(128, 2)), [Top1Accuracy()], 32)
    acc = res[0][0].result()[0]
    assert acc > 0.9, acc
    assert opt.state["epoch"] == 7


def test_graph_equals_sequential():
    RNG.setSeed(3)
    a = LeNet5(10)
    b = LeNet5_graph(10)
    b.setWeights(a.getWeights())
    x = torch.randn(4, 28, 28)
    ya, yb = a.forward(x), b.forward(x)
    assert torch.allclose(ya, yb, atol=1e-6)
    go = torch.randn(4, 10)
    ga, gb = a.backward(x, go), b.backward(x, go)
    assert torch.allclose(ga, gb, atol=1e-6)
    for pa, pb in zip(a.parameters()[1], b.parameters()[1]):
        assert torch.allclose(pa, pb, atol=1e-6)
