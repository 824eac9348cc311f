"""ImageFrame inference / validation pipeline (reference: P/examples/imageframe/inception_validation.py and
S/example/imageclassification, S/example/dlframes/imageInference): images on disk → ``ImageFrame.read`` → Resize /
CenterCrop / ChannelNormalize / MatToTensor / ImageFrameToSample → ``model.predictImage`` → top-1 accuracy.

No ImageNet here: the script writes a small folder of synthetic PNGs whose class is the dominant colour channel, fits a
tiny CNN on them (same transforms) and validates through the ImageFrame path.

    python examples/image_validation.py [--folder DIR] [-n 120]
"""
from __future__ import annotations

import argparse
import os
import sys
import tempfile

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def write_images(folder: str, n: int, seed: int = 0):
    import cv2

    rs = np.random.RandomState(seed)
    labels = {}
    for i in range(n):
        c = int(rs.randint(3))
        img = rs.randint(0, 90, size=(40, 48, 3)).astype(np.uint8)
        img[..., c] = np.clip(img[..., c].astype(np.int32) + 120 + rs.randint(0, 30), 0, 255).astype(np.uint8)  # BGR
        name = f"img_{i:04d}_c{c}.png"
        cv2.imwrite(os.path.join(folder, name), img)
        labels[name] = c + 1
    return labels


def pipeline():
    from bigdl_b200.transform.vision.image import (CenterCrop, ChannelNormalize, ImageFrameToSample, MatToTensor,
                                                   Pipeline, Resize)

    return Pipeline([Resize(36, 36), CenterCrop(32, 32), ChannelNormalize(100.0, 100.0, 100.0, 60.0, 60.0, 60.0),
                     MatToTensor(), ImageFrameToSample()])


def build_model():
    from bigdl_b200.nn import layer as L

    m = L.Sequential()
    m.add(L.SpatialConvolution(3, 8, 3, 3, 2, 2, 1, 1)).add(L.ReLU())
    m.add(L.SpatialAveragePooling(16, 16, 16, 16)).add(L.View(8).setNumInputDims(3))
    m.add(L.Linear(8, 3)).add(L.LogSoftMax())
    return m


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--folder", default=None)
    ap.add_argument("-n", "--numImages", type=int, default=120)
    ap.add_argument("-e", "--maxEpoch", type=int, default=30)
    a = ap.parse_args(argv)
    from bigdl_b200.dataset.base import DataSet, Sample
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import Trigger
    from bigdl_b200.transform.vision.image import ImageFrame
    from bigdl_b200.utils.engine import Engine
    from bigdl_b200.utils.random_generator import RNG

    Engine.init(device="cpu")
    RNG.setSeed(4)
    tmp = None
    folder = a.folder
    if folder is None:
        tmp = tempfile.TemporaryDirectory()
        folder = tmp.name
    labels = write_images(folder, a.numImages)

    frame = ImageFrame.read(folder).transform(pipeline())
    feats = list(frame.features) if hasattr(frame, "features") else list(frame)
    names = [os.path.basename(f["uri"]) for f in feats]
    xs = [torch.as_tensor(np.asarray(f["imageTensor"])) for f in feats]
    ys = [labels[n] for n in names]
    n_train = int(0.7 * len(xs))
    train = DataSet.array([Sample(x, torch.tensor([float(y)])) for x, y in zip(xs[:n_train], ys[:n_train])])
    model = build_model()
    Optimizer(model, train, ClassNLLCriterion(), Trigger.maxEpoch(a.maxEpoch), 16, SGD(0.2, momentum=0.9)).optimize()

    # validation goes through the image-frame path: predictImage writes a "predict" entry into every ImageFeature
    val = ImageFrame(feats[n_train:]) if not hasattr(ImageFrame, "array") else type(frame)(feats[n_train:])
    model.evaluate()
    model.predictImage(val, batchPerPartition=8)
    preds = [int(np.asarray(f["predict"]).argmax()) + 1 for f in (val.features if hasattr(val, "features") else val)]
    acc = float(np.mean([p == y for p, y in zip(preds, ys[n_train:])]))
    print(f"ImageFrame validation: {len(preds)} images, top-1 accuracy {acc:.3f}")
    if tmp is not None:
        tmp.cleanup()
    return acc


if __name__ == "__main__":
    main()
