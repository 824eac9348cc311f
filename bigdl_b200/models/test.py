"""The model-zoo evaluation mains, one sub-command per reference ``Test`` object:

    python -m bigdl_b200.models.test lenet     -f /data/mnist    --model lenet.bigdl [-b 128]
    python -m bigdl_b200.models.test vgg       -f /data/cifar10  --model vgg.bigdl
    python -m bigdl_b200.models.test resnet    -f /data/cifar10  --model resnet20.bigdl
    python -m bigdl_b200.models.test resnet-imagenet -f /data/imagenet-seq/val --model resnet50.bigdl
    python -m bigdl_b200.models.test inception -f /data/imagenet-seq/val --model inception.bigdl
    python -m bigdl_b200.models.test rnn       -f /tmp/rnn --model rnn.bigdl [--words 10] [--evaluate false]

(S/models/lenet/Test.scala:34-60, vgg/Test.scala, resnet/Test.scala, resnet/TestImageNet.scala:36-58,
inception/Test.scala:36-70, rnn/Test.scala:36-135.) Each prints ``<method> is <result>`` lines like the reference and
returns the list of (result, method) pairs."""
from __future__ import annotations

import argparse
import os
import sys


def _parser(title, batch=128):
    ap = argparse.ArgumentParser(title)
    ap.add_argument("-f", "--folder", default="./")
    ap.add_argument("--model", required=True, help="model snapshot location")
    ap.add_argument("-b", "--batchSize", type=int, default=batch)
    return ap


def _load(a):
    from ..cli._common import init_distributed
    from ..utils.engine import Engine
    from .train import load_any

    rank, world, dev = init_distributed()
    Engine.init()
    return load_any(a.model).to(dev), rank, world


def _report(results, rank=0):
    if rank == 0:
        for res, method in results:
            print(f"{method} is {res}")
    return results


def lenet(argv=None):
    a = _parser("BigDL Lenet Test Example").parse_args(argv)
    from ..dataset import image as I
    from ..dataset.base import DataSet
    from ..optim.validation import Top1Accuracy
    from . import datasets as D

    model, rank, world = _load(a)
    files = [os.path.join(a.folder, n) for n in ("t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte")]
    data = (DataSet.array(D.load_mnist(*files), world > 1) >> I.BytesToGreyImg(28, 28)
            >> I.GreyImgNormalizer(D.MNIST_TEST_MEAN, D.MNIST_TEST_STD) >> I.GreyImgToSample())
    return _report(model.evaluate(data, [Top1Accuracy()], a.batchSize), rank)


def _cifar(title, mean, std):
    def run(argv=None):
        a = _parser(title).parse_args(argv)
        from ..dataset import image as I
        from ..dataset.base import DataSet
        from ..optim.validation import Top1Accuracy
        from . import datasets as D

        model, rank, world = _load(a)
        data = (DataSet.array(D.load_cifar10_test(a.folder), world > 1) >> I.BytesToBGRImg()
                >> I.BGRImgNormalizer(*mean, *std) >> I.BGRImgToSample())
        return _report(model.evaluate(data, [Top1Accuracy()], a.batchSize), rank)

    return run


def resnet_imagenet(argv=None):
    a = _parser("BigDL ResNet ImageNet Test Example").parse_args(argv)
    from ..optim.validation import Top1Accuracy, Top5Accuracy
    from . import datasets as D

    model, rank, world = _load(a)
    data = D.ImageNetDataSet.valDataSet(a.folder, a.batchSize, 224, sc=world > 1 or None)
    return _report(model.evaluate(data, [Top1Accuracy(), Top5Accuracy()]), rank)


def inception(argv=None):
    a = _parser("BigDL Inception Test Example").parse_args(argv)
    from ..dataset import image as I
    from ..dataset.seqfile import seq_file_dataset
    from ..optim.validation import Top1Accuracy, Top5Accuracy

    model, rank, world = _load(a)
    data = (seq_file_dataset(a.folder, 1000, world > 1) >> I.BytesToBGRImg() >> I.BGRImgCropper(224, 224, I.CropCenter)
            >> I.HFlip(0.5) >> I.BGRImgNormalizer(0.485, 0.456, 0.406, 0.229, 0.224, 0.225) >> I.BGRImgToSample())
    return _report(model.evaluate(data, [Top1Accuracy(), Top5Accuracy()], a.batchSize), rank)


def rnn(argv=None):
    ap = _parser("BigDL rnn Test Example", batch=4)
    ap.add_argument("--words", dest="numOfWords", type=int, default=10, help="number of words to write")
    ap.add_argument("--evaluate", type=lambda s: str(s).lower() in ("1", "true", "yes"), default=True)
    ap.add_argument("--sent", dest="sentFile", default=None)
    ap.add_argument("--token", dest="tokenFile", default=None)
    a = ap.parse_args(argv)
    import re

    import torch

    from ..dataset import text as TX
    from ..dataset.base import DataSet, FixedLength, PaddingParam, SampleToMiniBatch
    from ..nn.criterion import CrossEntropyCriterion, TimeDistributedCriterion
    from ..optim.validation import Loss
    from . import datasets as D

    vocab = TX.Dictionary(a.folder)
    model, rank, world = _load(a)
    total = vocab.getVocabSize() + 1
    if a.evaluate:
        tokens = D.SequencePreprocess(os.path.join(a.folder, "test.txt"), None, a.sentFile, a.tokenFile)
        max_len = max(len(t) for t in tokens)
        pad_feature = torch.zeros(total)
        pad_feature[vocab.getIndex(TX.SentenceToken.end)] = 1.0
        pad_label = torch.tensor([float(vocab.getIndex(TX.SentenceToken.start)) + 1.0])
        data = (DataSet.array(tokens) >> TX.TextToLabeledSentence(vocab) >> TX.LabeledSentenceToSample(total)
                >> SampleToMiniBatch(a.batchSize, PaddingParam([pad_feature], FixedLength([max_len])),
                                     PaddingParam([pad_label], FixedLength([max_len]))))
        return _report(model.evaluate(data, [Loss(TimeDistributedCriterion(CrossEntropyCriterion(), True))]), rank)
    # generation: feed each line of test.txt, append the arg-max word ``numOfWords`` times
    path = os.path.join(a.folder, "test.txt")
    if not os.path.exists(path):
        raise FileNotFoundError(f"test file {path} not exists!")
    with open(path, encoding="utf-8") as f:
        lines = [[w for w in re.split(r"\W+", line) if w] for line in f.read().splitlines() if line.strip()]
    dev = model.device()
    model.evaluate()
    out = []
    for words in lines:
        ids = [vocab.getIndex(w) for w in words]
        cur = torch.zeros(1, len(ids), total, device=dev)
        cur[0, torch.arange(len(ids)), torch.tensor(ids)] = 1.0
        for _ in range(a.numOfWords):
            with torch.no_grad():
                pred = model.forward(cur)[:, -1].argmax(-1)
            nxt = torch.zeros(1, 1, total, device=dev)
            nxt[0, 0, int(pred)] = 1.0
            cur = torch.cat([cur, nxt], 1)
        sent = [vocab.getWord(int(i)) for i in cur[0].argmax(-1).tolist()]
        out.append(sent)
        if rank == 0:
            print(" ".join(sent))
    return out


def _mains():
    from . import datasets as D

    return {"lenet": lenet, "vgg": _cifar("BigDL Vgg on Cifar10 Test Example", D.CIFAR10_TEST_MEAN, D.CIFAR10_TEST_STD),
            "resnet": _cifar("BigDL ResNet on Cifar10 Test Example", D.CIFAR10_TEST_MEAN, D.CIFAR10_TEST_STD),
            "resnet-imagenet": resnet_imagenet, "inception": inception, "rnn": rnn}


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    mains = _mains()
    if not argv or argv[0] not in mains:
        print(__doc__)
        print("models:", ", ".join(sorted(mains)))
        return 0 if (not argv or argv[0] in ("-h", "--help")) else 2
    mains[argv[0]](argv[1:])
    return 0


if __name__ == "__main__":
    sys.exit(main())
