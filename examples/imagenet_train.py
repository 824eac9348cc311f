"""ImageNet-style training of Inception-v1 / ResNet-50 / VGG-16 with the DistriOptimizer (reference:
S/models/inception/Train.scala:40-120 — SGD with Warmup + Poly(0.5); S/models/resnet/TrainImageNet.scala:60-160 — SGD
momentum 0.9, warm-up then EpochDecay, optional LARS; S/models/vgg/Train.scala).

Data: Hadoop sequence files written by ``python -m bigdl_b200.cli.imagenet_seqfile_generator`` when --folder is given,
otherwise a synthetic 224x224 dataset (there is no network to fetch ImageNet here).

    scripts/bigdl-b200-run.sh -n 8 examples/imagenet_train.py --model resnet -b 64 --maxIteration 200
"""
from __future__ import annotations

import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def build(name, classes):
    from bigdl_b200.cli._common import build_model

    key = {"inception": "inception_v1", "resnet": "resnet50", "vgg": "vgg16"}.get(name, name)
    return build_model(key, classes)[0]


def optim_method(a, iters_per_epoch, world):
    from bigdl_b200.optim.optim_method import SGD, EpochBoundaries, EpochDecayWithWarmUp, Poly, SequentialSchedule, Warmup

    base = a.learningRate
    if a.model == "resnet":
        # linear warm-up to maxLr over warmupEpoch epochs, then x0.1 at epochs 30/60/80 (TrainImageNet.scala:117-135)
        warm_iters = a.warmupEpoch * iters_per_epoch
        delta = (a.maxLr - base) / max(warm_iters, 1)
        sched = EpochDecayWithWarmUp(warm_iters, delta, EpochBoundaries(30, 60, 80))
        return SGD(base, 0.0, a.weightDecay, 0.9, 0.0, False, learningrate_schedule=sched)
    max_iter = a.maxIteration or a.maxEpoch * iters_per_epoch
    warm_iters = a.warmupEpoch * iters_per_epoch
    if warm_iters > 0:
        sched = SequentialSchedule(iters_per_epoch)
        sched.add(Warmup((a.maxLr - base) / warm_iters), warm_iters).add(Poly(0.5, max_iter), max_iter - warm_iters)
    else:
        sched = Poly(0.5, max_iter)
    return SGD(base, 0.0, a.weightDecay, 0.9, 0.0, False, learningrate_schedule=sched)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="inception", choices=["inception", "resnet", "vgg"])
    ap.add_argument("-f", "--folder", default=None, help="folder with train/*.seq and val/*.seq")
    ap.add_argument("-b", "--batchSize", type=int, default=64, help="per GPU")
    ap.add_argument("--classNum", type=int, default=1000)
    ap.add_argument("--maxEpoch", type=int, default=1)
    ap.add_argument("--maxIteration", type=int, default=40)
    ap.add_argument("--learningRate", type=float, default=0.01)
    ap.add_argument("--maxLr", type=float, default=0.05)
    ap.add_argument("--warmupEpoch", type=int, default=0)
    ap.add_argument("--weightDecay", type=float, default=1e-4)
    ap.add_argument("--checkpoint", default=None)
    ap.add_argument("--checkpointIteration", type=int, default=0,
                    help="snapshot every N iterations instead of every epoch (models/inception/Options.scala)")
    ap.add_argument("--resume", default=None, help="checkpoint folder: continue from its newest model.N / state.N "
                    "(the reference's --model / --state snapshot pair)")
    ap.add_argument("--summary", default=None, help="TensorBoard log dir")
    ap.add_argument("--syncBN", action="store_true")
    a = ap.parse_args(argv)
    from bigdl_b200.cli._common import init_distributed
    from bigdl_b200.dataset.base import DataSet, FixedMiniBatchDataSet, MiniBatch
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import Trigger
    from bigdl_b200.optim.validation import Top1Accuracy, Top5Accuracy
    from bigdl_b200.utils.engine import Engine

    rank, world, dev = init_distributed()
    Engine.init()
    model, resumed_om = build(a.model, a.classNum), None
    if a.resume:
        from bigdl_b200.nn.module import Module
        from bigdl_b200.optim.optim_method import OptimMethod
        from bigdl_b200.serialization.snapshot import latest_file

        mf, sf = latest_file(a.resume, "model."), latest_file(a.resume, "state.")
        if mf is None or sf is None:
            raise SystemExit(f"--resume {a.resume}: no model.N / state.N pair found")
        model, resumed_om = Module.load(mf), OptimMethod.load(sf)
        if rank == 0:
            print(f"resuming from {mf} (iteration {resumed_om.state['neval'] - 1}, epoch {resumed_om.state['epoch']})")
    model = model.to(dev)
    if a.syncBN:
        from bigdl_b200.parallel import enable_sync_batchnorm

        enable_sync_batchnorm(model)
    val = None
    if a.folder:
        from bigdl_b200.dataset.image import MTLabeledBGRImgToBatch
        from bigdl_b200.dataset.seqfile import seq_file_dataset

        batcher = MTLabeledBGRImgToBatch(224, 224, a.batchSize, (123.0, 117.0, 104.0), (1.0, 1.0, 1.0), to_rgb=True, pin=True)
        train = seq_file_dataset(os.path.join(a.folder, "train"), a.classNum).transform(batcher)
        if os.path.isdir(os.path.join(a.folder, "val")):
            vb = MTLabeledBGRImgToBatch(224, 224, a.batchSize, (123.0, 117.0, 104.0), (1.0, 1.0, 1.0), random_crop=False,
                                        random_flip=False, to_rgb=True)
            val = seq_file_dataset(os.path.join(a.folder, "val"), a.classNum).transform(vb)
        iters_per_epoch = max(train.size() // (a.batchSize * world), 1)
    else:
        x = torch.rand(a.batchSize, 3, 224, 224)
        y = torch.randint(1, a.classNum + 1, (a.batchSize,)).float()
        train = FixedMiniBatchDataSet(MiniBatch(x, y), epoch_size=a.batchSize * world * 100, distributed=world > 1)
        iters_per_epoch = 100
    end = Trigger.maxIteration(a.maxIteration) if a.maxIteration else Trigger.maxEpoch(a.maxEpoch)
    opt = Optimizer(model=model, training_set=train, criterion=ClassNLLCriterion(),
                    optim_method=resumed_om or optim_method(a, iters_per_epoch, world), end_trigger=end, batch_size=a.batchSize * world)
    if val is not None:
        opt.setValidation(Trigger.everyEpoch(), val, [Top1Accuracy(), Top5Accuracy()], a.batchSize)
    if a.checkpoint:
        opt.setCheckpoint(a.checkpoint, Trigger.severalIteration(a.checkpointIteration) if a.checkpointIteration
                          else Trigger.everyEpoch())
    if a.summary and rank == 0:
        from bigdl_b200.visualization import TrainSummary, ValidationSummary

        opt.setTrainSummary(TrainSummary(a.summary, a.model))
        opt.setValidationSummary(ValidationSummary(a.summary, a.model))
    opt.optimize()
    if rank == 0:
        print(f"finished {opt.state['neval'] - 1} iterations, last loss {opt.state['Loss']:.4f}")


if __name__ == "__main__":
    main()
