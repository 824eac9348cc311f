"""CNN text classifier over word embeddings (reference: S/example/textclassification/TextClassifier.scala and
P/models/textclassifier/textclassifier.py — tokens → GloVe vectors [seq, embed] → TemporalConvolution/ReLU/
TemporalMaxPooling stack → Linear → LogSoftMax, Adagrad).

Uses 20-Newsgroups + GloVe from --baseDir when present (P/dataset/news20.py layout); otherwise a synthetic corpus
where each class has its own vocabulary bias and embeddings are random, so the example runs offline.

    python examples/text_classification.py [--baseDir /tmp/news20] [-b 32] [-e 3]
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def synthetic(n=600, classes=4, vocab=200, length=60, seed=0):
    rs = np.random.RandomState(seed)
    prefs = np.random.RandomState(7).dirichlet(np.full(vocab, 0.1), size=classes)      # per-class word distribution
    texts = []
    for _ in range(n):
        c = rs.randint(classes)
        words = rs.choice(vocab, size=length, p=prefs[c])
        texts.append((" ".join(f"tok{w}" for w in words), c + 1))
    w2v = {f"tok{i}": np.random.RandomState(100 + i).randn(50).astype(np.float32) for i in range(vocab)}
    return texts, w2v, classes


def build_model(class_num, embed, seq_len):
    from bigdl_b200.nn import layer as L

    m = L.Sequential()
    m.add(L.TemporalConvolution(embed, 128, 5)).add(L.ReLU())
    m.add(L.TemporalMaxPooling(5, 5))
    m.add(L.TemporalConvolution(128, 128, 5)).add(L.ReLU())
    pooled = ((seq_len - 4) // 5 - 4)
    m.add(L.TemporalMaxPooling(pooled, pooled))
    m.add(L.Reshape([128]))
    m.add(L.Linear(128, 100)).add(L.Dropout(0.2)).add(L.ReLU())
    m.add(L.Linear(100, class_num)).add(L.LogSoftMax())
    return m


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--baseDir", default=None)
    ap.add_argument("-b", "--batchSize", type=int, default=32)
    ap.add_argument("-e", "--maxEpoch", type=int, default=4)
    ap.add_argument("--maxSequenceLength", type=int, default=60)
    ap.add_argument("--embeddingDim", type=int, default=50)
    ap.add_argument("--trainingSplit", type=float, default=0.8)
    ap.add_argument("--learningRate", type=float, default=0.05)
    a = ap.parse_args(argv)
    from bigdl_b200.dataset.base import DataSet, Sample
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.optim.optim_method import Adagrad
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import Trigger
    from bigdl_b200.optim.validation import Top1Accuracy
    from bigdl_b200.utils.engine import Engine

    dev = "cuda" if torch.cuda.is_available() else "cpu"
    Engine.init(device=dev)
    if a.baseDir and os.path.isdir(os.path.join(a.baseDir, "20news-18828")):
        from bigdl_b200.dataset import news20

        texts = news20.get_news20(a.baseDir)
        w2v = {k: np.asarray(v, dtype=np.float32) for k, v in news20.get_glove_w2v(a.baseDir, a.embeddingDim).items()}
        classes = news20.CLASS_NUM
    else:
        texts, w2v, classes = synthetic()
        a.embeddingDim = 50
    zero = np.zeros(a.embeddingDim, dtype=np.float32)

    def vectorise(text):
        toks = [t for t in text.lower().split() if t in w2v][:a.maxSequenceLength]
        vecs = [w2v[t] for t in toks] + [zero] * (a.maxSequenceLength - len(toks))
        return np.stack(vecs)

    samples = [Sample(vectorise(t), np.array([float(l)], dtype=np.float32)) for t, l in texts]
    rs = np.random.RandomState(1)
    rs.shuffle(samples)
    n_train = int(len(samples) * a.trainingSplit)
    train, val = DataSet.array(samples[:n_train]), DataSet.array(samples[n_train:])
    model = build_model(classes, a.embeddingDim, a.maxSequenceLength).to(dev)
    opt = Optimizer(model=model, training_set=train, criterion=ClassNLLCriterion(),
                    optim_method=Adagrad(learningrate=a.learningRate, learningrate_decay=0.001),
                    end_trigger=Trigger.maxEpoch(a.maxEpoch), batch_size=a.batchSize)
    opt.setValidation(Trigger.everyEpoch(), val, [Top1Accuracy()], a.batchSize)
    opt.optimize()
    acc = model.evaluate(val, [Top1Accuracy()], a.batchSize)[0][0].result()[0]
    print(f"validation Top1Accuracy {acc:.4f} over {classes} classes")
    return acc


if __name__ == "__main__":
    main()
