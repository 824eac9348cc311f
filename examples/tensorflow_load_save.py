"""TensorFlow interop examples (reference: S/example/tensorflow/loadandsave/{Load,Save}.scala and
S/example/tensorflow/transferlearning/TransferLearning.scala).

* save: export a model as a frozen GraphDef (`saveTF`), the file `Load.scala` / a TensorFlow session would consume
* load: read a GraphDef back with named input / output endpoints (`loadTF`) and compare outputs
* transfer learning: cut the loaded graph at a feature endpoint, freeze it, attach a new classifier and train only that

    python examples/tensorflow_load_save.py
"""
from __future__ import annotations

import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def lenet_like():
    from bigdl_b200.nn import layer as L

    return (L.Sequential()
            .add(L.SpatialConvolution(1, 6, 5, 5).setName("conv1")).add(L.Tanh().setName("tanh1"))
            .add(L.SpatialMaxPooling(2, 2, 2, 2).setName("pool1"))
            .add(L.SpatialConvolution(6, 12, 5, 5).setName("conv2")).add(L.Tanh().setName("tanh2"))
            .add(L.SpatialMaxPooling(2, 2, 2, 2).setName("pool2"))
            .add(L.View(12 * 4 * 4).setName("flat"))
            .add(L.Linear(192, 32).setName("fc1")).add(L.Tanh().setName("feature"))
            .add(L.Linear(32, 10).setName("fc2")).add(L.LogSoftMax().setName("output")))


def main(argv=None):
    from bigdl_b200.dataset.base import DataSet, Sample
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.nn.module import Module
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import Trigger
    from bigdl_b200.utils.engine import Engine

    Engine.init(device="cpu")
    torch.manual_seed(0)
    from bigdl_b200.utils.random_generator import RNG

    RNG.setSeed(7)                       # layer initialisation and the epoch shuffle draw from the framework's generator
    d = tempfile.mkdtemp()
    pb = os.path.join(d, "lenet.pb")
    # ---- Save.scala: model.saveTF(Seq(("input", Seq(4, 1, 28, 28))), path)
    model = lenet_like().evaluate()
    model.saveTF([("input", [None, 1, 28, 28])], pb)
    print(f"saved {os.path.getsize(pb)} bytes of GraphDef to {pb}")
    # ---- Load.scala: Module.loadTF(path, Seq("input"), Seq("output"))
    loaded = Module.loadTF(pb, ["input"], ["output"]).evaluate()
    x = torch.randn(4, 1, 28, 28)
    diff = float((loaded.forward(x.permute(0, 2, 3, 1).contiguous()) - model.forward(x)).abs().max())
    print(f"max |loaded - original| = {diff:.2e}")
    # ---- TransferLearning.scala: features from an inner endpoint, new head trained on top
    features = Module.loadTF(pb, ["input"], ["feature"]).evaluate()
    features.freeze()
    n = 200
    xs = torch.randn(n, 28, 28, 1)
    with torch.no_grad():
        f = features.forward(xs)
    w_true = torch.randn(f.shape[1], 3)
    ys = (f @ w_true).argmax(1) + 1                              # a task that IS linearly solvable from the features
    head = L.Sequential().add(L.Linear(f.shape[1], 3)).add(L.LogSoftMax())
    net = L.Sequential().add(features).add(head)
    before = [p.detach().clone() for p in features.parameters()[0]]
    samples = [Sample(xs[i], torch.tensor([float(ys[i])])) for i in range(n)]
    Optimizer(model=net, training_set=DataSet.array(samples), criterion=ClassNLLCriterion(),
              optim_method=SGD(learningrate=0.5), end_trigger=Trigger.maxEpoch(40), batch_size=50).optimize()
    net.evaluate()
    acc = float((net.forward(xs).argmax(1) + 1 == ys).float().mean())
    frozen_ok = all(torch.equal(b, p.detach()) for b, p in zip(before, features.parameters()[0]))
    print(f"transfer-learning accuracy {acc:.3f}; loaded graph weights unchanged: {frozen_ok}")
    return diff, acc, frozen_ok


if __name__ == "__main__":
    main()
