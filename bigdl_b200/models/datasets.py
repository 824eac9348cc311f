"""The data pipelines the model-zoo mains build (S/models/{lenet,vgg,resnet,inception,rnn,autoencoder}/*): MNIST and
CIFAR-10 byte records, the CIFAR / ImageNet MiniBatch datasets of ``Cifar10DataSet`` / ``ImageNetDataSet``
(S/models/resnet/DataSet.scala:28-150), ``ImageNet2012`` / ``ImageNet2012Val`` (S/models/inception/ImageNet2012.scala),
the VGG ImageNet pipeline (S/models/vgg/Utils.scala:211-240), ``SequencePreprocess`` (S/models/rnn/Utils.scala:184-270)
and ``toAutoencoderBatch`` (S/models/autoencoder/Train.scala:31-45).

``sc`` arguments are accepted where the reference has a SparkContext overload: a non-None value (or an initialised
process group) makes the dataset rank-sharded, which is this framework's form of an RDD."""
from __future__ import annotations

import os
import struct

import numpy as np
import torch

from ..dataset import image as I
from ..dataset import mnist as _mnist
from ..dataset import text as TX
from ..dataset.base import DataSet, MiniBatch, Transformer
from ..utils import file as _file

# ------------------------------------------------------------------------------------------------ MNIST
MNIST_TRAIN_MEAN, MNIST_TRAIN_STD = 0.13066047740239506, 0.3081078
MNIST_TEST_MEAN, MNIST_TEST_STD = 0.13251460696903547, 0.31048024


def _read_bytes(path: str) -> bytes:
    if "://" in path or path.startswith("hdfs:"):
        return _file.File.readBytes(path)
    with open(path, "rb") as f:
        return f.read()


def load_mnist(featureFile: str, labelFile: str) -> list[I.ByteRecord]:
    """idx image + label files → ByteRecords of raw 28×28 bytes, labels shifted to 1..10
    (S/models/lenet/Utils.scala ``load``)."""
    fb, lb = _read_bytes(featureFile), _read_bytes(labelFile)
    magic_l, n_l = struct.unpack_from(">ii", lb, 0)
    magic_f, n_f, rows, cols = struct.unpack_from(">iiii", fb, 0)
    if magic_l != 2049 or magic_f != 2051:
        raise ValueError("not MNIST idx files (bad magic number)")
    if n_l != n_f:
        raise ValueError(f"{n_f} images but {n_l} labels")
    px = rows * cols
    return [I.ByteRecord(fb[16 + i * px:16 + (i + 1) * px], float(lb[8 + i]) + 1.0) for i in range(n_f)]


def mnist_batches(folder: str, kind: str, batchSize: int, sc=None):
    """BytesToGreyImg → GreyImgNormalizer → GreyImgToBatch over the train or t10k files of ``folder``."""
    img, lab = _mnist._FILES["train" if kind == "train" else "test"]
    mean, std = (MNIST_TRAIN_MEAN, MNIST_TRAIN_STD) if kind == "train" else (MNIST_TEST_MEAN, MNIST_TEST_STD)

    def find(name):
        for cand in (name, name + ".gz"):
            p = os.path.join(folder, cand)
            if os.path.exists(p):
                return p
        raise FileNotFoundError(os.path.join(folder, name))

    fi, fl = find(img), find(lab)
    if fi.endswith(".gz"):
        x, y = _mnist.read_data_sets(folder, "train" if kind == "train" else "test")
        recs = [I.ByteRecord(x[i].astype(np.uint8).tobytes(), float(y[i]) + 1.0) for i in range(len(y))]
    else:
        recs = load_mnist(fi, fl)
    return (DataSet.array(recs, sc is not None) >> I.BytesToGreyImg(28, 28) >> I.GreyImgNormalizer(mean, std)
            >> I.GreyImgToBatch(batchSize))


class toAutoencoderBatch(Transformer):
    """MiniBatch(x, y) → MiniBatch(x, x): the reconstruction target is the input."""

    def __call__(self, prev):
        for b in prev:
            x = b.getInput()
            yield MiniBatch(x, x)


# ------------------------------------------------------------------------------------------------ CIFAR-10
def load_cifar10_file(featureFile: str, result: list | None = None) -> list[I.ByteRecord]:
    """One CIFAR-10 binary batch (rows of 1 label byte + 3×32×32 planar RGB bytes) → ByteRecords with the 8-byte
    width/height header and interleaved BGR pixels, labels 1..10 (S/models/vgg/Utils.scala:166-206)."""
    result = [] if result is None else result
    raw = np.frombuffer(_read_bytes(featureFile), dtype=np.uint8)
    n = raw.size // 3073
    rows = raw[: n * 3073].reshape(n, 3073)
    bgr = rows[:, 1:].reshape(n, 3, 32, 32)[:, ::-1].transpose(0, 2, 3, 1)       # planar RGB → interleaved BGR
    header = struct.pack(">ii", 32, 32)
    bgr = np.ascontiguousarray(bgr)
    for i in range(n):
        result.append(I.ByteRecord(header + bgr[i].tobytes(), float(rows[i, 0]) + 1.0))
    return result


def load_cifar10_train(dataFile: str) -> list[I.ByteRecord]:
    out: list = []
    for i in range(1, 6):
        load_cifar10_file(os.path.join(dataFile, f"data_batch_{i}.bin"), out)
    return out


def load_cifar10_test(dataFile: str) -> list[I.ByteRecord]:
    return load_cifar10_file(os.path.join(dataFile, "test_batch.bin"))


CIFAR10_TRAIN_MEAN = (0.4913996898739353, 0.4821584196221302, 0.44653092422369434)
CIFAR10_TRAIN_STD = (0.24703223517429462, 0.2434851308749409, 0.26158784442034005)
CIFAR10_TEST_MEAN = (0.4942142913295297, 0.4851314002725445, 0.45040910258647154)
CIFAR10_TEST_STD = (0.2466525177466614, 0.2428922662655766, 0.26159238066790275)


class ResNetDataSet:
    """Interface of the two dataset objects below (S/models/resnet/DataSet.scala:28-38)."""

    def trainDataSet(self, path: str, batchSize: int, size: int = 32, sc=None):  # pragma: no cover - interface
        raise NotImplementedError

    def valDataSet(self, path: str, batchSize: int, size: int = 32, sc=None):  # pragma: no cover - interface
        raise NotImplementedError


class _Cifar10DataSet(ResNetDataSet):
    trainMean, trainStd, testMean, testStd = CIFAR10_TRAIN_MEAN, CIFAR10_TRAIN_STD, CIFAR10_TEST_MEAN, CIFAR10_TEST_STD

    def trainDataSet(self, path, batchSize, size=32, sc=None):
        """normalise → flip with probability 0.5 → zero-pad 4 and crop 32×32 at random → batch."""
        return (DataSet.array(load_cifar10_train(path), sc is not None) >> I.BytesToBGRImg()
                >> I.BGRImgNormalizer(*self.trainMean, *self.trainStd) >> I.HFlip(0.5)
                >> I.BGRImgRdmCropper(32, 32, 4) >> I.BGRImgToBatch(batchSize))

    def valDataSet(self, path, batchSize, size=32, sc=None):
        return (DataSet.array(load_cifar10_test(path), sc is not None) >> I.BytesToBGRImg()
                >> I.BGRImgNormalizer(*self.testMean, *self.testStd) >> I.BGRImgToBatch(batchSize))


Cifar10DataSet = _Cifar10DataSet()


# ------------------------------------------------------------------------------------------------ ImageNet
def _mt_batches(path, sc, class_num, image_size, batch_size, transformer):
    from ..transform.vision import image as V

    frame = V.SeqFileFolder.files_to_image_frame(path, sc, class_num)
    batcher = V.MTImageFeatureToBatch(image_size, image_size, batch_size, transformer, to_rgb=False)
    return DataSet.array(frame.features, sc is not None).transform(_Batcher(batcher))


class _Batcher(Transformer):
    def __init__(self, batcher):
        self.batcher = batcher

    def __call__(self, prev):
        return self.batcher(prev)


class _ImageNetDataSet(ResNetDataSet):
    """ResNet-50 ImageNet pipelines over SequenceFiles (S/models/resnet/DataSet.scala:88-150): training = random
    aspect / area crop to 224 + flip; validation = resize 256 + centre crop; both scaled by 1/128 around the channel
    means (104, 117, 123)."""

    trainMean, trainStd = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    testMean, testStd = trainMean, trainStd

    def trainDataSet(self, path, batchSize, size=224, sc=None):
        from ..transform.vision import image as V

        t = V.Pipeline([V.PixelBytesToMat(), V.RandomAlterAspect(), V.RandomCropper(224, 224, True, "random"),
                        V.ChannelScaledNormalizer(104, 117, 123, 0.0078125), V.MatToTensor()])
        return _mt_batches(path, sc, 1000, size, batchSize, t)

    def valDataSet(self, path, batchSize, size=224, sc=None):
        from ..transform.vision import image as V

        t = V.Pipeline([V.PixelBytesToMat(), V.RandomResize(256, 256), V.CenterCrop(224, 224),
                        V.ChannelScaledNormalizer(104, 117, 123, 0.0078125), V.MatToTensor()])
        return _mt_batches(path, sc, 1000, size, batchSize, t)


ImageNetDataSet = _ImageNetDataSet()


class ImageNet2012:
    """Inception training data: resize 256×256, random 224 crop with mirroring, subtract (123, 117, 104)
    (S/models/inception/ImageNet2012.scala:28-66)."""

    crop, mirror = "random", True

    def __new__(cls, path, sc=None, imageSize: int = 224, batchSize: int = 32, nodeNumber: int = 1,
                coresPerNode: int = 1, classNumber: int = 1000):
        from ..transform.vision import image as V

        t = V.Pipeline([V.PixelBytesToMat(), V.Resize(256, 256), V.RandomCropper(224, 224, cls.mirror, cls.crop),
                        V.ChannelNormalize(123, 117, 104), V.MatToTensor()])
        return _mt_batches(path, sc, classNumber, imageSize, batchSize, t)

    @classmethod
    def rdd(cls, path, batchSize, sc=None, imageSize: int = 224):
        """The ImageFrame form: crop → flip(0.5) → normalise to the torchvision statistics → Sample → MiniBatch."""
        from ..transform.vision import image as V

        crop = V.RandomCrop(imageSize, imageSize) if cls.crop == "random" else V.CenterCrop(imageSize, imageSize)
        t = V.Pipeline([V.PixelBytesToMat(), crop, V.RandomTransformer(V.HFlip(), 0.5),
                        V.ChannelNormalize(0.485, 0.456, 0.406, 0.229, 0.224, 0.225), V.MatToTensor(),
                        V.ImageFrameToSample(target_keys=["label"])])
        frame = V.SeqFileFolder.files_to_image_frame(path, sc, 1000)
        feats = DataSet.array(frame.features, sc is not None).transform(_Each(t))
        return feats.transform(_Batcher(V.ImageFeatureToMiniBatch(batchSize)))


class ImageNet2012Val(ImageNet2012):
    """Validation twin: centre crop, no mirroring (ImageNet2012.scala:68-108)."""
    crop, mirror = "center", False


class _Each(Transformer):
    def __init__(self, ft):
        self.ft = ft

    def __call__(self, prev):
        return (self.ft.transform(f.clone()) for f in prev)


def vgg_imagenet_train(path, sc=None, imageSize: int = 224, batchSize: int = 32):
    """S/models/vgg/Utils.scala:226-240."""
    from ..transform.vision import image as V

    t = V.Pipeline([V.PixelBytesToMat(), V.RandomResize(256, 256), V.RandomCropper(224, 224, True, "random"),
                    V.ChannelScaledNormalizer(104, 117, 124, 1), V.MatToTensor()])
    return _mt_batches(path, sc, 1000, imageSize, batchSize, t)


def vgg_imagenet_val(path, sc=None, imageSize: int = 224, batchSize: int = 32):
    """S/models/vgg/Utils.scala:211-224."""
    from ..transform.vision import image as V

    t = V.Pipeline([V.PixelBytesToMat(), V.RandomResize(256, 256), V.RandomCropper(224, 224, False, "center"),
                    V.ChannelScaledNormalizer(104, 117, 124, 1), V.MatToTensor()])
    return _mt_batches(path, sc, 1000, imageSize, batchSize, t)


# ------------------------------------------------------------------------------------------------ text
class SequencePreprocess:
    """``SequencePreprocess(fileName, sc, sentBin, tokenBin)`` → token lists with start / end markers;
    ``SequencePreprocess.ptb(dir, vocabSize)`` → (train, valid, test) 1-based word-id streams + Dictionary of the PTB
    files; ``reader(raw, numSteps)`` cuts a stream into windows of numSteps + 1 that overlap by one word."""

    def __new__(cls, fileName: str, sc=None, sentBin=None, tokenBin=None, vocabSize: int | None = None):
        if os.path.isdir(fileName) or vocabSize is not None:
            return cls.ptb(fileName, vocabSize if vocabSize is not None else sc)
        if not os.path.exists(fileName):
            raise FileNotFoundError(f"data file {fileName} not exists!")
        with open(fileName, encoding="utf-8") as f:
            lines = f.read().splitlines()
        sentences = [s for para in TX.SentenceSplitter(sentBin)(iter(lines)) for s in para]
        padded = TX.SentenceBiPadding()(iter(sentences))
        return list(TX.SentenceTokenizer(tokenBin)(padded))

    @staticmethod
    def readWords(fileName: str) -> list[str]:
        out = []
        with open(fileName, encoding="utf-8") as f:
            for line in f.read().splitlines():
                out.extend(line.split(" "))
                out.append("<eos>")
        return out

    @staticmethod
    def fileToWordIdx(fileName: str, dictionary) -> list[float]:
        return [float(dictionary.getIndex(w)) + 1.0 for w in SequencePreprocess.readWords(fileName)]

    @staticmethod
    def ptb(fileDirect: str, vocabSize: int):
        paths = [os.path.join(fileDirect, f"ptb.{k}.txt") for k in ("train", "valid", "test")]
        dictionary = TX.Dictionary(SequencePreprocess.readWords(paths[0]), vocabSize - 1)
        tr, va, te = (SequencePreprocess.fileToWordIdx(p, dictionary) for p in paths)
        return tr, va, te, dictionary

    @staticmethod
    def reader(rawData, numSteps: int):
        out, offset = [], 0
        while offset <= len(rawData) - 1 - numSteps:
            out.append(list(rawData[offset:offset + numSteps + 1]))
            offset += numSteps
        return out
