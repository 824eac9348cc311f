"""Single-process throughput harness (S/models/utils/LocalOptimizerPerf.scala): same as DistriOptimizerPerf with the
LocalOptimizer; runs on the CPU when no GPU is present (the LeNet-5 "local mode" configuration)."""
from __future__ import annotations

import argparse
import json
import time

import torch

from ._common import build_model, synthetic_batch


def main(argv=None):
    ap = argparse.ArgumentParser(description="Performance Test of Local Optimizer")
    ap.add_argument("-b", "--batchSize", type=int, default=128)
    ap.add_argument("-i", "--iteration", type=int, default=80)
    ap.add_argument("-m", "--model", default="inception_v1",
                    choices=["inception_v1", "inception_v2", "vgg16", "vgg19", "resnet50", "lenet5"])
    ap.add_argument("-d", "--inputdata", default="random", choices=["constant", "random"])
    ap.add_argument("--device", default=None)
    a = ap.parse_args(argv)
    from ..dataset.base import FixedMiniBatchDataSet, MiniBatch
    from ..nn.criterion import ClassNLLCriterion
    from ..optim.optim_method import SGD
    from ..optim.optimizer import LocalOptimizer
    from ..optim.trigger import Trigger
    from ..utils.engine import Engine

    dev = torch.device(a.device or ("cuda" if torch.cuda.is_available() else "cpu"))
    Engine.init(device=dev.type)
    model, shape = build_model(a.model)
    model = model.to(dev)
    x, y = synthetic_batch(a.batchSize, shape, a.inputdata)
    ds = FixedMiniBatchDataSet(MiniBatch(x, y), epoch_size=10 ** 9)
    opt = LocalOptimizer(model, ds, ClassNLLCriterion(), Trigger.maxIteration(a.iteration), a.batchSize, SGD(0.01))
    marks = []
    opt.setIterationCallback(lambda state, loss: marks.append(time.perf_counter()))
    opt.optimize()
    if dev.type == "cuda":
        torch.cuda.synchronize()
    steady = marks[len(marks) // 2:]
    per = (steady[-1] - steady[0]) / max(len(steady) - 1, 1)
    print(json.dumps({"tool": "LocalOptimizerPerf", "model": a.model, "device": str(dev), "batch": a.batchSize,
                      "iterations": len(marks), "ms_per_iteration": round(per * 1e3, 3),
                      "records_per_second": round(a.batchSize / per, 1)}))


if __name__ == "__main__":
    main()
