"""LeNet-5 trained in a single process without any cluster (reference: S/example/lenetLocal/{Train,Test,Predict}.scala,
P/models/local_lenet/local_lenet.py). Reads the MNIST idx files from --folder when present, otherwise trains on a
synthetic digit-like dataset so the example runs offline.

    python examples/lenet_local.py [-f /tmp/mnist] [-b 128] [-e 2] [--checkpoint /tmp/lenet]
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def synthetic_mnist(n: int, seed: int = 0):
    """Ten fixed random 28x28 prototypes plus noise: learnable in a few hundred iterations."""
    protos = np.random.RandomState(1234).rand(10, 28, 28).astype(np.float32)      # shared by train and test
    rs = np.random.RandomState(seed)
    labels = rs.randint(0, 10, size=n)
    images = 2.0 * (protos[labels] - 0.5) + 0.3 * rs.randn(n, 28, 28).astype(np.float32)
    return images.reshape(n, 1, 28, 28), labels.astype(np.float32) + 1          # labels are 1-based


def load(folder: str, kind: str, n_synth: int):
    from bigdl_b200.dataset import mnist

    try:
        x, y = mnist.read_data_sets(folder, kind)
        mean, std = (mnist.TRAIN_MEAN, mnist.TRAIN_STD) if kind == "train" else (mnist.TEST_MEAN, mnist.TEST_STD)
        return ((x.astype(np.float32).reshape(-1, 1, 28, 28) - mean) / std), y.astype(np.float32) + 1
    except FileNotFoundError:
        return synthetic_mnist(n_synth, seed=0 if kind == "train" else 1)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("-f", "--folder", default="/tmp/mnist")
    ap.add_argument("-b", "--batchSize", type=int, default=128)
    ap.add_argument("-e", "--maxEpoch", type=int, default=4)
    ap.add_argument("-r", "--learningRate", type=float, default=0.05)
    ap.add_argument("--checkpoint", default=None)
    ap.add_argument("--device", default="cpu")
    a = ap.parse_args(argv)
    from bigdl_b200.dataset.base import DataSet
    from bigdl_b200.models.lenet import LeNet5
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import Trigger
    from bigdl_b200.optim.validation import Loss, Top1Accuracy
    from bigdl_b200.utils.engine import Engine

    Engine.init(device=a.device)
    xtr, ytr = load(a.folder, "train", 4096)
    xte, yte = load(a.folder, "test", 1024)
    model = LeNet5(10).to(a.device)
    opt = Optimizer(model=model, training_set=DataSet.from_ndarrays(xtr, ytr), criterion=ClassNLLCriterion(),
                    optim_method=SGD(learningrate=a.learningRate, learningrate_decay=0.0002),
                    end_trigger=Trigger.maxEpoch(a.maxEpoch), batch_size=a.batchSize)
    opt.setValidation(Trigger.everyEpoch(), DataSet.from_ndarrays(xte, yte), [Top1Accuracy(), Loss(ClassNLLCriterion())],
                      a.batchSize)
    if a.checkpoint:
        opt.setCheckpoint(a.checkpoint, Trigger.everyEpoch())
    opt.optimize()
    res = model.evaluate(DataSet.from_ndarrays(xte, yte), [Top1Accuracy()], a.batchSize)
    acc = res[0][0].result()[0]
    print(f"Top1Accuracy on the test split: {acc:.4f}")
    pred = model.predictClass(DataSet.from_ndarrays(xte[:8]), 8)
    print("predicted classes of the first 8 test images:", [int(p) for p in pred], "labels:", yte[:8].astype(int).tolist())
    return acc


if __name__ == "__main__":
    main()
