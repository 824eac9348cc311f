"""Word-level LSTM language model (reference: S/example/languagemodel/PTBWordLM.scala + PTBModel.scala — embedding →
stacked LSTM → TimeDistributed(Linear) trained with TimeDistributedCriterion(CrossEntropy), Adagrad; and
S/models/rnn/Train.scala for the simple RNN variant).

Reads a plain-text corpus (one sentence per line) from --dataFolder/{train,valid}.txt when present (the PTB layout),
otherwise generates a small synthetic corpus from a random first-order Markov chain so the example runs offline.

    python examples/language_model.py [--dataFolder /tmp/ptb] [--hiddenSize 64] [--numSteps 20] [-e 2]
"""
from __future__ import annotations

import argparse
import math
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def synthetic_corpus(n_sent=600, vocab=40, length=21, seed=0):
    trans = np.random.RandomState(99).dirichlet(np.full(vocab, 0.05), size=vocab)   # peaky, shared by train/valid
    rs = np.random.RandomState(seed)
    sents = []
    for _ in range(n_sent):
        w = [rs.randint(vocab)]
        for _ in range(length - 1):
            w.append(rs.choice(vocab, p=trans[w[-1]]))
        sents.append([f"w{t}" for t in w])
    return sents


def load_corpus(folder, name):
    from bigdl_b200.dataset.text import SentenceTokenizer, read_sentences

    path = os.path.join(folder or "", f"{name}.txt")
    if folder and os.path.exists(path):
        return list(SentenceTokenizer()(iter(read_sentences(path))))
    return synthetic_corpus(seed=0 if name == "train" else 1, n_sent=600 if name == "train" else 100)


def build_model(vocab, embed, hidden, layers, keep_prob):
    from bigdl_b200.nn import layer as L

    m = L.Sequential().add(L.LookupTable(vocab, embed))
    if keep_prob < 1:
        m.add(L.Dropout(1 - keep_prob))
    n_in = embed
    for _ in range(layers):
        m.add(L.Recurrent().add(L.LSTM(n_in, hidden)))
        n_in = hidden
    return m.add(L.TimeDistributed(L.Linear(hidden, vocab)))


def to_samples(sentences, dictionary, num_steps):
    from bigdl_b200.dataset.base import Sample

    stream = [dictionary.getIndex(w) + 1 for s in sentences for w in s]              # 1-based ids
    out = []
    for i in range(0, len(stream) - num_steps - 1, num_steps):
        x = np.asarray(stream[i:i + num_steps], dtype=np.float32)
        y = np.asarray(stream[i + 1:i + num_steps + 1], dtype=np.float32)
        out.append(Sample(x, y))
    return out


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("-f", "--dataFolder", default=None)
    ap.add_argument("-b", "--batchSize", type=int, default=20)
    ap.add_argument("-e", "--nEpochs", type=int, default=3)
    ap.add_argument("--vocabSize", type=int, default=10000)
    ap.add_argument("--hiddenSize", type=int, default=64)
    ap.add_argument("--numLayers", type=int, default=1)
    ap.add_argument("--numSteps", type=int, default=20)
    ap.add_argument("--keepProb", type=float, default=1.0)
    ap.add_argument("--learningRate", type=float, default=0.1)
    a = ap.parse_args(argv)
    from bigdl_b200.dataset.base import DataSet
    from bigdl_b200.dataset.text import Dictionary
    from bigdl_b200.nn.criterion import CrossEntropyCriterion, TimeDistributedCriterion
    from bigdl_b200.optim.optim_method import Adagrad
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import Trigger
    from bigdl_b200.optim.validation import Loss
    from bigdl_b200.utils.engine import Engine

    dev = "cuda" if torch.cuda.is_available() else "cpu"
    Engine.init(device=dev)
    train, valid = load_corpus(a.dataFolder, "train"), load_corpus(a.dataFolder, "valid")
    dictionary = Dictionary(train, a.vocabSize)
    vocab = dictionary.getVocabSize() + 1                                         # +1: the out-of-vocabulary bucket
    model = build_model(vocab, a.hiddenSize, a.hiddenSize, a.numLayers, a.keepProb).to(dev)
    crit = TimeDistributedCriterion(CrossEntropyCriterion(), size_average=True)
    opt = Optimizer(model=model, training_set=DataSet.array(to_samples(train, dictionary, a.numSteps)), criterion=crit,
                    optim_method=Adagrad(learningrate=a.learningRate), end_trigger=Trigger.maxEpoch(a.nEpochs),
                    batch_size=a.batchSize)
    val_set = DataSet.array(to_samples(valid, dictionary, a.numSteps))
    opt.setValidation(Trigger.everyEpoch(), val_set, [Loss(crit)], a.batchSize)
    opt.optimize()
    loss = model.evaluate(val_set, [Loss(crit)], a.batchSize)[0][0].result()[0]
    print(f"validation loss {loss:.4f}  perplexity {math.exp(loss):.2f}  (uniform guess would be {vocab})")
    return math.exp(loss), vocab


if __name__ == "__main__":
    main()
