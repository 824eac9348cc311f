"""DataFrame image pipelines (reference: S/example/dlframes/imageInference/ImageInference.scala and
imageTransferLearning/ImageTransferLearning.scala).

* inference: DLImageReader.readImages -> DLImageTransformer(resize, normalise, to tensor) -> DLClassifierModel.transform
* transfer learning: a frozen feature extractor produces an "embedding" column (DLModel), a DLClassifier trains a new
  logistic-regression head on it, and the two are chained for prediction.

    python examples/dlframes_image.py           # synthetic two-class images (bright vs dark blobs)
"""
from __future__ import annotations

import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def write_images(folder: str, n: int = 40, seed: int = 0):
    import cv2

    rs = np.random.RandomState(seed)
    labels = {}
    for i in range(n):
        cls = i % 2
        img = rs.randint(0, 60, (40, 40, 3)).astype(np.uint8)
        if cls:
            cv2.circle(img, (int(rs.randint(12, 28)), int(rs.randint(12, 28))), 9, (230, 230, 230), -1)
        path = os.path.join(folder, f"img{i:03d}.png")
        cv2.imwrite(path, img)
        labels[path] = float(cls + 1)
    return labels


def to_tensor_column(df, col, out, size=32):
    df[out] = [np.frombuffer(r["data"], dtype=np.float32).reshape(size, size, 3).transpose(2, 0, 1).copy() for r in df[col]]
    return df


def feature_extractor():
    from bigdl_b200.nn import layer as L

    return (L.Sequential()
            .add(L.SpatialConvolution(3, 8, 5, 5, 2, 2)).add(L.ReLU())
            .add(L.SpatialMaxPooling(2, 2, 2, 2))
            .add(L.SpatialConvolution(8, 8, 3, 3)).add(L.ReLU())
            .add(L.SpatialAveragePooling(5, 5, 5, 5)).add(L.View(8)))


def main(argv=None):
    from bigdl_b200.dlframes import DLClassifier, DLImageReader, DLImageTransformer
    from bigdl_b200.dlframes.dl_classifier import DLClassifierModel, DLModel
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.transform.vision import image as V
    from bigdl_b200.utils.engine import Engine

    Engine.init(device="cpu")
    folder = tempfile.mkdtemp()
    labels = write_images(folder)
    # ---- shared preprocessing (ImageInference.scala:44-52)
    df = DLImageReader.readImages(folder)
    pre = DLImageTransformer(V.Resize(32, 32) >> V.ChannelNormalize(60, 60, 60, 60, 60, 60)).setOutputCol("features")
    df = to_tensor_column(pre.transform(df), "features", "tensor")
    df["label"] = [labels[r["origin"]] for r in df["image"]]
    # ---- transfer learning (ImageTransferLearning.scala:47-71): frozen extractor -> embedding column -> new head
    extractor = feature_extractor().evaluate()
    emb = DLModel(extractor, [3, 32, 32]).setFeaturesCol("tensor").setPredictionCol("embedding").setBatchSize(8)
    df = emb.transform(df)
    train, val = df.iloc[:30].copy(), df.iloc[30:].copy()
    head = L.Sequential().add(L.Linear(8, 2)).add(L.LogSoftMax())
    clf = DLClassifier(head, ClassNLLCriterion(), [8]).setFeaturesCol("embedding").setLabelCol("label") \
        .setBatchSize(10).setMaxEpoch(60).setLearningRate(0.5)
    head_model = clf.fit(train)
    acc = float((head_model.transform(val)["prediction"] == val["label"]).mean())
    print(f"transfer-learning head accuracy on held-out images: {acc:.3f}")
    # ---- inference with the chained model (ImageInference.scala:54-60)
    full = L.Sequential().add(extractor).add(head)
    out = DLClassifierModel(full, [3, 32, 32]).setFeaturesCol("tensor").setBatchSize(8).transform(val)
    same = float((out["prediction"] == head_model.transform(val)["prediction"]).mean())
    print("chained model predictions:", out["prediction"].tolist()[:6], f"(agreement with the two-stage pipeline {same:.2f})")
    return acc, same


if __name__ == "__main__":
    main()
