"""DataFrame pipelines (reference: S/example/MLPipeline/{DLClassifierLogisticRegression,DLEstimatorMultiLabelLR}.scala,
S/example/dlframes/imageInference, S/example/udfpredictor): fit a DLClassifier / DLEstimator on a DataFrame, transform a
DataFrame with the fitted model, and apply a model as a column-wise UDF over an image DataFrame."""
from __future__ import annotations

import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def logistic_regression():
    import pandas as pd

    from bigdl_b200.dlframes import DLClassifier
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.criterion import ClassNLLCriterion

    rs = np.random.RandomState(0)
    x = rs.randn(200, 2).astype(np.float32)
    df = pd.DataFrame({"features": list(x), "label": ((x[:, 0] + x[:, 1] > 0) + 1).astype(float)})
    model = L.Sequential().add(L.Linear(2, 2)).add(L.LogSoftMax())
    clf = DLClassifier(model, ClassNLLCriterion(), [2]).setBatchSize(20).setMaxEpoch(30).setLearningRate(0.5)
    out = clf.fit(df).transform(df)
    acc = float((out["prediction"] == df["label"]).mean())
    print(f"DLClassifier logistic regression accuracy: {acc:.3f}")
    return acc


def multi_label_regression():
    import pandas as pd

    from bigdl_b200.dlframes import DLEstimator
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn.criterion import MSECriterion

    rs = np.random.RandomState(1)
    x = rs.randn(200, 2).astype(np.float32)
    y = np.stack([x[:, 0] * 2 + 1, x[:, 1] - x[:, 0]], 1).astype(np.float32)
    df = pd.DataFrame({"features": list(x), "label": list(y)})
    est = DLEstimator(L.Sequential().add(L.Linear(2, 2)), MSECriterion(), [2], [2]).setBatchSize(20).setMaxEpoch(40) \
        .setLearningRate(0.1)
    out = est.fit(df).transform(df)
    err = float(np.abs(np.stack(out["prediction"].to_list()) - y).mean())
    print(f"DLEstimator multi-output regression mean abs error: {err:.4f}")
    return err


def image_udf():
    import tempfile

    import cv2
    import torch

    from bigdl_b200.dlframes import DLImageReader, DLImageTransformer
    from bigdl_b200.dlframes.dl_classifier import DLClassifierModel
    from bigdl_b200.nn import layer as L
    from bigdl_b200.transform.vision import image as V

    d = tempfile.mkdtemp()
    rs = np.random.RandomState(2)
    for i in range(4):
        cv2.imwrite(os.path.join(d, f"{i}.png"), rs.randint(0, 256, (40, 48, 3), dtype=np.uint8))
    df = DLImageReader.readImages(d)
    df = DLImageTransformer(V.Resize(32, 32) >> V.ChannelNormalize(120, 120, 120, 60, 60, 60)).setOutputCol("features").transform(df)
    df["tensor"] = [np.frombuffer(r["data"], dtype=np.float32).reshape(32, 32, 3).transpose(2, 0, 1) for r in df["features"]]
    net = L.Sequential().add(L.SpatialConvolution(3, 4, 3, 3)).add(L.ReLU()).add(L.View(4 * 30 * 30)).add(L.Linear(3600, 5)) \
        .add(L.LogSoftMax())
    out = DLClassifierModel(net, [3, 32, 32]).setFeaturesCol("tensor").setBatchSize(2).transform(df)
    print("image DataFrame predictions:", out["prediction"].tolist())
    return out


if __name__ == "__main__":
    from bigdl_b200.utils.engine import Engine

    Engine.init(device="cpu")
    logistic_regression()
    multi_label_regression()
    image_udf()
