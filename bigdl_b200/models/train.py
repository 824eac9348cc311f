"""The model-zoo training mains, one sub-command per reference ``Train`` object, with the reference's option names and
hyper-parameters:

    python -m bigdl_b200.models.train lenet        -f /data/mnist   [-b 12] [-r 0.05] [-e 15] [--checkpoint dir]
    python -m bigdl_b200.models.train autoencoder  -f /data/mnist   [-b 150] [-e 10]
    python -m bigdl_b200.models.train vgg          -f /data/cifar10 [-b 112] [-e 90]
    python -m bigdl_b200.models.train resnet       -f /data/cifar10 [--depth 20] [--shortcutType A] [--nEpochs 165]
    python -m bigdl_b200.models.train rnn          -f /data/text -s /tmp/rnn [--hidden 40] [--vocab 4000]
    python -m bigdl_b200.models.train inception | resnet-imagenet | vgg-imagenet   -f /data/imagenet-seq ...

(S/models/lenet/Train.scala:36-100, autoencoder/Train.scala:47-95, vgg/Train.scala:36-110, resnet/TrainCIFAR10.scala:
34-100, rnn/Train.scala:36-135; the three ImageNet mains — inception/Train.scala, resnet/TrainImageNet.scala,
vgg/TrainImageNet.scala — share ``examples/imagenet_train.py``'s loop and are forwarded to it.)

Every main resumes from ``--model`` / ``--state`` snapshots, writes checkpoints every epoch when ``--checkpoint`` is
given (``--overWrite`` keeps one copy), and runs on all GPUs when started under ``scripts/bigdl-b200-run.sh -n N``.
Each returns the Optimizer so tests can inspect the final state."""
from __future__ import annotations

import argparse
import os
import sys
import time


def _common(ap, folder_flag=("-f", "--folder"), batch=128, epochs=10, epoch_flag=("-e", "--maxEpoch")):
    ap.add_argument(*folder_flag, dest="folder", default="./", help="where the data is")
    ap.add_argument("-b", "--batchSize", type=int, default=batch)
    ap.add_argument("--model", dest="modelSnapshot", default=None, help="model snapshot location")
    ap.add_argument("--state", dest="stateSnapshot", default=None, help="state snapshot location")
    ap.add_argument("--checkpoint", default=None, help="where to cache the model")
    ap.add_argument(*epoch_flag, dest="maxEpoch", type=int, default=epochs)
    ap.add_argument("--overWrite", dest="overWriteCheckpoint", action="store_true")
    ap.add_argument("--maxIteration", type=int, default=0, help="stop after N iterations instead (smoke runs)")
    return ap


def _setup():
    from ..cli._common import init_distributed
    from ..utils.engine import Engine

    rank, world, dev = init_distributed()
    Engine.init()
    return rank, world, dev


def load_any(path: str):
    """A checkpoint written by ``Module.save`` / the optimizer (snapshot format) or a ``saveModule`` protobuf file."""
    from ..nn.module import Module

    try:
        return Module.load(path)
    except Exception:
        return Module.loadModule(path)


def _model(a, build, dev):
    m = load_any(a.modelSnapshot) if a.modelSnapshot else build()
    if getattr(a, "graphModel", False) and not a.modelSnapshot:
        m = m.toGraph()
    return m.to(dev)


def _optim(a, build):
    from ..optim.optim_method import OptimMethod

    return OptimMethod.load(a.stateSnapshot) if a.stateSnapshot else build()


def _finish(opt, a, val=None, methods=None, summary_name=None):
    from ..optim.trigger import Trigger

    if a.checkpoint:
        os.makedirs(a.checkpoint, exist_ok=True)
        opt.setCheckpoint(a.checkpoint, Trigger.everyEpoch())
    if a.overWriteCheckpoint:
        opt.overWriteCheckpoint()
    if val is not None:
        opt.setValidation(Trigger.everyEpoch(), val, methods)
    if getattr(a, "summaryPath", None) and summary_name:
        from ..visualization import TrainSummary, ValidationSummary

        stamp = time.strftime("%Y-%m-%d %H:%M:%S")
        opt.setTrainSummary(TrainSummary(a.summaryPath, f"{summary_name}-train-{stamp}"))
        opt.setValidationSummary(ValidationSummary(a.summaryPath, f"{summary_name}-val-{stamp}"))
    opt.setEndWhen(Trigger.maxIteration(a.maxIteration) if a.maxIteration else Trigger.maxEpoch(a.maxEpoch))
    opt.optimize()
    return opt


# ---------------------------------------------------------------------------------------------------- lenet
def lenet(argv=None):
    ap = _common(argparse.ArgumentParser("BigDL Lenet Train Example"), batch=12, epochs=15)
    ap.add_argument("-r", "--learningRate", type=float, default=0.05)
    ap.add_argument("-d", "--learningRateDecay", type=float, default=0.0)
    ap.add_argument("-g", "--graphModel", action="store_true")
    a = ap.parse_args(argv)
    from ..nn.criterion import ClassNLLCriterion
    from ..optim.optim_method import SGD
    from ..optim.optimizer import Optimizer
    from ..optim.validation import Loss, Top1Accuracy, Top5Accuracy
    from . import datasets as D
    from .lenet import LeNet5

    rank, world, dev = _setup()
    model = _model(a, lambda: LeNet5(10), dev)
    om = _optim(a, lambda: SGD(learningrate=a.learningRate, learningrate_decay=a.learningRateDecay))
    train = D.mnist_batches(a.folder, "train", a.batchSize, sc=world > 1 or None)
    val = D.mnist_batches(a.folder, "test", a.batchSize)
    opt = Optimizer(model=model, training_set=train, criterion=ClassNLLCriterion(), optim_method=om)
    return _finish(opt, a, val, [Top1Accuracy(), Top5Accuracy(), Loss(ClassNLLCriterion())])


# ---------------------------------------------------------------------------------------------------- autoencoder
def autoencoder(argv=None):
    ap = _common(argparse.ArgumentParser("BigDL Autoencoder on MNIST"), batch=150, epochs=10)
    ap.add_argument("-g", "--graphModel", action="store_true")
    a = ap.parse_args(argv)
    from ..nn.criterion import MSECriterion
    from ..optim.optim_method import Adagrad
    from ..optim.optimizer import Optimizer
    from . import datasets as D
    from .autoencoder import Autoencoder

    rank, world, dev = _setup()
    model = _model(a, lambda: Autoencoder(32), dev)
    om = _optim(a, lambda: Adagrad(learningrate=0.01, learningrate_decay=0.0, weightdecay=0.0005))
    train = D.mnist_batches(a.folder, "train", a.batchSize, sc=world > 1 or None).transform(D.toAutoencoderBatch())
    opt = Optimizer(model=model, training_set=train, criterion=MSECriterion(), optim_method=om)
    return _finish(opt, a)


# ---------------------------------------------------------------------------------------------------- vgg / cifar-10
def vgg(argv=None):
    ap = _common(argparse.ArgumentParser("BigDL Vgg on Cifar10 Example"), batch=112, epochs=90)
    ap.add_argument("-l", "--learningRate", type=float, default=0.01)
    ap.add_argument("--weightDecay", type=float, default=0.0005)
    ap.add_argument("--summary", dest="summaryPath", default=None)
    ap.add_argument("-g", "--graphModel", action="store_true")
    a = ap.parse_args(argv)
    from ..dataset import image as I
    from ..dataset.base import DataSet
    from ..nn.criterion import ClassNLLCriterion
    from ..optim.optim_method import SGD, EpochStep
    from ..optim.optimizer import Optimizer
    from ..optim.validation import Top1Accuracy
    from . import datasets as D
    from .vgg import VggForCifar10

    rank, world, dev = _setup()
    model = _model(a, lambda: VggForCifar10(10), dev)
    om = _optim(a, lambda: SGD(learningrate=a.learningRate, learningrate_decay=0.0, weightdecay=a.weightDecay,
                               momentum=0.9, dampening=0.0, nesterov=False, learningrate_schedule=EpochStep(25, 0.5)))
    train = (DataSet.array(D.load_cifar10_train(a.folder), world > 1) >> I.BytesToBGRImg()
             >> I.BGRImgNormalizer(*D.CIFAR10_TRAIN_MEAN, *D.CIFAR10_TRAIN_STD) >> I.BGRImgToBatch(a.batchSize))
    val = (DataSet.array(D.load_cifar10_test(a.folder)) >> I.BytesToBGRImg()
           >> I.BGRImgNormalizer(*D.CIFAR10_TEST_MEAN, *D.CIFAR10_TEST_STD) >> I.BGRImgToBatch(a.batchSize))
    opt = Optimizer(model=model, training_set=train, criterion=ClassNLLCriterion(), optim_method=om)
    return _finish(opt, a, val, [Top1Accuracy()], "vgg-on-cifar10")


# ---------------------------------------------------------------------------------------------------- resnet / cifar-10
def cifar10Decay(epoch: int) -> float:
    """lr = base · 0.1^decay: ÷10 from epoch 81, ÷100 from epoch 122 (TrainCIFAR10.scala:34-35)."""
    return 2.0 if epoch >= 122 else (1.0 if epoch >= 81 else 0.0)


def resnet(argv=None):
    ap = _common(argparse.ArgumentParser("BigDL ResNet Example"), batch=128, epochs=165, epoch_flag=("-e", "--nepochs", "--nEpochs"))
    ap.add_argument("--optnet", action="store_true", help="share gradInput buffers (a no-op here: autograd frees them)")
    ap.add_argument("--depth", type=int, default=20)
    ap.add_argument("--classes", type=int, default=10)
    ap.add_argument("--shortcutType", default="A", choices=["A", "B", "C"])
    ap.add_argument("--learningRate", type=float, default=0.1)
    ap.add_argument("--weightDecay", type=float, default=1e-4)
    ap.add_argument("--momentum", type=float, default=0.9)
    ap.add_argument("--dampening", type=float, default=0.0)
    ap.add_argument("--nesterov", type=lambda s: str(s).lower() in ("1", "true", "yes"), default=True)
    ap.add_argument("-g", "--graphModel", action="store_true")
    a = ap.parse_args(argv)
    from ..nn.criterion import CrossEntropyCriterion
    from ..optim.optim_method import SGD, EpochDecay
    from ..optim.optimizer import Optimizer
    from ..optim.validation import Top1Accuracy
    from . import datasets as D
    from .resnet import ResNet, model_init

    rank, world, dev = _setup()

    def build():
        m = ResNet(a.classes, depth=a.depth, shortcut_type=a.shortcutType, dataset="cifar10")
        model_init(m)
        return m

    model = _model(a, build, dev)
    om = _optim(a, lambda: SGD(learningrate=a.learningRate, learningrate_decay=0.0, weightdecay=a.weightDecay,
                               momentum=a.momentum, dampening=a.dampening, nesterov=a.nesterov,
                               learningrate_schedule=EpochDecay(cifar10Decay)))
    train = D.Cifar10DataSet.trainDataSet(a.folder, a.batchSize, 32, sc=world > 1 or None)
    val = D.Cifar10DataSet.valDataSet(a.folder, a.batchSize, 32)
    opt = Optimizer(model=model, training_set=train, criterion=CrossEntropyCriterion(), optim_method=om)
    return _finish(opt, a, val, [Top1Accuracy()])


# ---------------------------------------------------------------------------------------------------- rnn
def rnn(argv=None):
    ap = argparse.ArgumentParser("BigDL SimpleRNN Train Example")
    ap.add_argument("-f", "--dataFolder", default="./", help="folder with train.txt and val.txt")
    ap.add_argument("-s", "--saveFolder", default="./", help="where the dictionary is written")
    ap.add_argument("--model", dest="modelSnapshot", default=None)
    ap.add_argument("--state", dest="stateSnapshot", default=None)
    ap.add_argument("--checkpoint", default=None)
    ap.add_argument("-b", "--batchSize", type=int, default=128)
    ap.add_argument("-r", "--learningRate", type=float, default=0.1)
    ap.add_argument("-m", "--momentum", type=float, default=0.0)
    ap.add_argument("--weightDecay", type=float, default=0.0)
    ap.add_argument("--dampening", type=float, default=0.0)
    ap.add_argument("--hidden", dest="hiddenSize", type=int, default=40)
    ap.add_argument("--vocab", dest="vocabSize", type=int, default=4000)
    ap.add_argument("--bptt", type=int, default=4)
    ap.add_argument("-e", "--nEpochs", dest="maxEpoch", type=int, default=30)
    ap.add_argument("--sent", dest="sentFile", default=None)
    ap.add_argument("--token", dest="tokenFile", default=None)
    ap.add_argument("--overWrite", dest="overWriteCheckpoint", action="store_true")
    ap.add_argument("--maxIteration", type=int, default=0)
    a = ap.parse_args(argv)
    import torch

    from ..dataset import text as TX
    from ..dataset.base import DataSet, FixedLength, PaddingParam, SampleToMiniBatch
    from ..nn.criterion import CrossEntropyCriterion, TimeDistributedCriterion
    from ..optim.optim_method import SGD
    from ..optim.optimizer import Optimizer
    from ..optim.validation import Loss
    from . import datasets as D
    from .rnn import SimpleRNN

    rank, world, dev = _setup()
    tokens = D.SequencePreprocess(os.path.join(a.dataFolder, "train.txt"), None, a.sentFile, a.tokenFile)
    dictionary = TX.Dictionary(tokens, a.vocabSize)
    os.makedirs(a.saveFolder, exist_ok=True)
    if rank == 0:
        dictionary.save(a.saveFolder)
    valtokens = D.SequencePreprocess(os.path.join(a.dataFolder, "val.txt"), None, a.sentFile, a.tokenFile)
    max_train = max(len(t) for t in tokens)
    total = dictionary.getVocabSize() + 1
    start_idx, end_idx = dictionary.getIndex(TX.SentenceToken.start), dictionary.getIndex(TX.SentenceToken.end)
    pad_feature = torch.zeros(total)
    pad_feature[end_idx] = 1.0                                   # pad with the one-hot END token …
    pad_label = torch.tensor([float(start_idx) + 1.0])           # … and the START label, as the reference does
    fpad = PaddingParam([pad_feature], FixedLength([max_train]))
    lpad = PaddingParam([pad_label], FixedLength([max_train]))

    def pipeline(tok, distributed=False):
        return (DataSet.array(tok, distributed) >> TX.TextToLabeledSentence(dictionary)
                >> TX.LabeledSentenceToSample(total) >> SampleToMiniBatch(a.batchSize, fpad, lpad))

    def build():
        m = SimpleRNN(total, a.hiddenSize, total)
        m.reset()
        return m

    model = _model(a, build, dev)
    om = _optim(a, lambda: SGD(learningrate=a.learningRate, learningrate_decay=0.0, weightdecay=a.weightDecay,
                               momentum=a.momentum, dampening=a.dampening))
    crit = TimeDistributedCriterion(CrossEntropyCriterion(), True)
    opt = Optimizer(model=model, training_set=pipeline(tokens, world > 1), criterion=crit, optim_method=om)
    valset = pipeline([t for t in valtokens if len(t) <= max_train + 1])
    return _finish(opt, a, valset, [Loss(TimeDistributedCriterion(CrossEntropyCriterion(), True))])


# ---------------------------------------------------------------------------------------------------- ImageNet mains
def _imagenet(model_name):
    def run(argv=None):
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        sys.path.insert(0, os.path.join(root, "examples"))
        import imagenet_train

        return imagenet_train.main(["--model", model_name] + list(argv or []))

    return run


MAINS = {"lenet": lenet, "autoencoder": autoencoder, "vgg": vgg, "resnet": resnet, "resnet-cifar10": resnet, "rnn": rnn,
         "inception": _imagenet("inception"), "resnet-imagenet": _imagenet("resnet"), "vgg-imagenet": _imagenet("vgg")}


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] in ("-h", "--help") or argv[0] not in MAINS:
        print(__doc__)
        print("models:", ", ".join(sorted(MAINS)))
        return 2 if argv and argv[0] not in ("-h", "--help") else 0
    MAINS[argv[0]](argv[1:])
    return 0


if __name__ == "__main__":
    sys.exit(main())
