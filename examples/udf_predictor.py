"""Text classifier as a DataFrame UDF (reference: S/example/udfpredictor/{DataframePredictor,Utils}.scala — train or load
the text CNN, wrap `predict` as a UDF, apply it to a text column and filter / group on the predicted category).

    python examples/udf_predictor.py            # synthetic corpus; prints per-category counts and the UDF accuracy
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def make_udf(model, w2v, seq_len, embed):
    """text -> 1-based category, the `textClassifier` UDF of Utils.scala:150-175: tokenise, shape to seq_len, embed,
    forward one sample."""
    model.evaluate()
    zero = np.zeros(embed, dtype=np.float32)

    def classify(text: str) -> int:
        toks = text.lower().split()[:seq_len]
        vecs = [w2v.get(t, zero) for t in toks] + [zero] * (seq_len - len(toks))
        x = torch.from_numpy(np.stack(vecs)).unsqueeze(0)
        with torch.no_grad():
            return int(model.forward(x).reshape(-1).argmax()) + 1
    return classify


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--maxEpoch", type=int, default=6)
    ap.add_argument("--batchSize", type=int, default=32)
    ap.add_argument("--modelPath", default=None, help="a saved .bigdl text classifier; trained here when absent")
    a = ap.parse_args(argv)
    import pandas as pd
    import text_classification as tc
    from bigdl_b200.dataset.base import DataSet, Sample
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.nn.module import Module
    from bigdl_b200.optim.optim_method import Adagrad
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import Trigger
    from bigdl_b200.utils.engine import Engine

    Engine.init(device="cpu")
    seq_len, embed = 60, 50
    texts, w2v, classes = tc.synthetic(n=500)
    if a.modelPath and os.path.exists(a.modelPath):
        model = Module.loadModule(a.modelPath)
    else:
        model = tc.build_model(classes, embed, seq_len)
        zero = np.zeros(embed, dtype=np.float32)
        samples = []
        for t, label in texts[:400]:
            toks = t.split()[:seq_len]
            vecs = [w2v.get(w, zero) for w in toks] + [zero] * (seq_len - len(toks))
            samples.append(Sample(torch.from_numpy(np.stack(vecs)), torch.tensor([float(label)])))
        Optimizer(model=model, training_set=DataSet.array(samples), criterion=ClassNLLCriterion(),
                  optim_method=Adagrad(learningrate=0.05), end_trigger=Trigger.maxEpoch(a.maxEpoch),
                  batch_size=a.batchSize).optimize()
        if a.modelPath:
            model.saveModule(a.modelPath, overWrite=True)
    udf = make_udf(model, w2v, seq_len, embed)
    df = pd.DataFrame({"filename": [f"doc{i}.txt" for i in range(100)], "text": [t for t, _ in texts[400:]],
                       "label": [l for _, l in texts[400:]]})
    df["textLabel"] = df["text"].map(udf)                       # df.withColumn("textLabel", classifierUDF($"text"))
    print(df.groupby("textLabel").size().to_string())           # the reference's "group by predicted type" query
    hits = df[df["textLabel"] == 1][["filename", "textLabel"]]  # and its filter query
    print(f"{len(hits)} documents predicted as category 1")
    acc = float((df["textLabel"] == df["label"]).mean())
    print(f"UDF accuracy on held-out documents: {acc:.3f}")
    return acc


if __name__ == "__main__":
    main()
