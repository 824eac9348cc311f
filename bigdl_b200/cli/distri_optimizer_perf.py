"""DistriOptimizer throughput harness (S/models/utils/DistriOptimizerPerf.scala:30-160): a fixed synthetic mini-batch
per rank, ClassNLLCriterion, default SGD, ``maxEpoch`` passes over a 10000-record virtual epoch; throughput is what the
optimizer logs per iteration. Launch one process per GPU:

    torchrun --nproc-per-node 8 -m bigdl_b200.cli.distri_optimizer_perf -m inception_v1 -b 128 -i 50
"""
from __future__ import annotations

import argparse
import json
import time

import torch

from ._common import build_model, init_distributed, synthetic_batch


def main(argv=None):
    ap = argparse.ArgumentParser(description="Performance Test of Distribute Optimizer")
    ap.add_argument("-b", "--batchSize", type=int, default=128, help="batch size PER GPU")
    ap.add_argument("-e", "--maxEpoch", type=int, default=None)
    ap.add_argument("-i", "--maxIteration", type=int, default=50)
    ap.add_argument("-t", "--type", default="float", choices=["float", "double"])
    ap.add_argument("-m", "--model", default="inception_v1",
                    choices=["inception_v1", "inception_v2", "vgg16", "vgg19", "resnet50"])
    ap.add_argument("-d", "--inputdata", default="random", choices=["constant", "random"])
    a = ap.parse_args(argv)
    from ..dataset.base import FixedMiniBatchDataSet, MiniBatch
    from ..nn.criterion import ClassNLLCriterion
    from ..optim.optim_method import SGD
    from ..optim.optimizer import DistriOptimizer, LocalOptimizer
    from ..optim.trigger import Trigger
    from ..utils.engine import Engine

    rank, world, dev = init_distributed()
    Engine.init()
    model, shape = build_model(a.model)
    model = model.to(dev)
    x, y = synthetic_batch(a.batchSize, shape, a.inputdata)
    ds = FixedMiniBatchDataSet(MiniBatch(x, y), epoch_size=10000)
    end = Trigger.maxEpoch(a.maxEpoch) if a.maxEpoch else Trigger.maxIteration(a.maxIteration)
    cls = DistriOptimizer if world > 1 else LocalOptimizer
    opt = cls(model, ds, ClassNLLCriterion(), end, a.batchSize * world, SGD(learningrate=0.01))
    marks = []
    opt.setIterationCallback(lambda state, loss: marks.append(time.perf_counter()))
    opt.optimize()
    if dev.type == "cuda":
        torch.cuda.synchronize()
    if rank == 0 and len(marks) > 10:
        steady = marks[len(marks) // 2:]
        per = (steady[-1] - steady[0]) / (len(steady) - 1)
        print(json.dumps({"tool": "DistriOptimizerPerf", "model": a.model, "gpus": world, "batch_per_gpu": a.batchSize,
                          "iterations": len(marks), "ms_per_iteration": round(per * 1e3, 3),
                          "records_per_second": round(a.batchSize * world / per, 1)}))


if __name__ == "__main__":
    main()
