"""Sentiment classification with a constituency Tree-LSTM (reference: S/example/treeLSTMSentiment/{Train,TreeSentiment,
Utils}.scala — Stanford Sentiment Treebank: GloVe embeddings → BinaryTreeLSTM → per-node classifier, Adagrad,
TreeNNAccuracy on the root).

The treebank is not bundled; without --baseDir the example builds a synthetic corpus (words carry a polarity, the label
of a sentence is the sign of their sum, "not" flips the word that follows) with left-branching parse trees in the
reference's encoding: one row per tree node = (left child, right child, leaf word index), 1-based, 0 = none.

    python examples/tree_lstm_sentiment.py [-b 16] [-e 6] [--hiddenSize 32]
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

VOCAB, NOT = 40, 40          # word ids 1..39 carry a polarity, id 40 is the negation word


def left_branching_tree(n_words: int) -> np.ndarray:
    """2n-1 nodes: internal nodes 1..n-1 (1 = root), leaves n..2n-1 (leaf of word k is node n-1+k)."""
    rows = np.zeros((2 * n_words - 1, 3), dtype=np.float32)
    for i in range(1, n_words):                       # internal node i spans words 1..n-i+1
        right_word = n_words - i + 1
        left = i + 1 if i < n_words - 1 else n_words  # the last internal node joins the leaves of words 1 and 2
        rows[i - 1] = (left, n_words - 1 + right_word, 0)
    for k in range(1, n_words + 1):
        rows[n_words - 1 + k - 1] = (0, 0, k)
    return rows


def synthetic(n=600, length=7, seed=0):
    rs = np.random.RandomState(seed)
    polarity = np.random.RandomState(3).choice([-1.0, 1.0], size=VOCAB + 1)
    polarity[NOT] = 0.0
    xs, ys = [], []
    tree = left_branching_tree(length)
    while len(xs) < n:
        words = rs.randint(1, VOCAB + 1, size=length)
        score, flip = 0.0, 1.0
        for w in words:
            if w == NOT:
                flip = -1.0
                continue
            score += flip * polarity[w]
            flip = 1.0
        if score == 0:
            continue
        xs.append(words.astype(np.float32))
        ys.append(1.0 if score < 0 else 2.0)
    return np.stack(xs), np.asarray(ys, dtype=np.float32), tree


def build_model(embed: int, hidden: int, classes: int = 2):
    from bigdl_b200.nn import layer as L

    m = L.Sequential()
    m.add(L.ParallelTable().add(L.LookupTable(VOCAB, embed)).add(L.Identity()))
    m.add(L.BinaryTreeLSTM(embed, hidden))
    m.add(L.Select(2, 1))                              # hidden state of the root node
    m.add(L.Linear(hidden, classes)).add(L.LogSoftMax())
    return m


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("-b", "--batchSize", type=int, default=16)
    ap.add_argument("-e", "--maxEpoch", type=int, default=6)
    ap.add_argument("--hiddenSize", type=int, default=32)
    ap.add_argument("--embeddingDim", type=int, default=16)
    ap.add_argument("--learningRate", type=float, default=0.1)
    a = ap.parse_args(argv)
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.optim.optim_method import Adagrad
    from bigdl_b200.utils.engine import Engine
    from bigdl_b200.utils.random_generator import RNG
    from bigdl_b200.utils.table import Table

    Engine.init(device="cpu")                          # recursive, data-dependent control flow: a host-side model
    RNG.setSeed(1)
    torch.manual_seed(1)
    x, y, tree = synthetic()
    n_train = int(0.8 * len(x))
    model, crit, om = build_model(a.embeddingDim, a.hiddenSize), ClassNLLCriterion(), Adagrad(a.learningRate)
    w, g = model.getParameters()
    trees = torch.from_numpy(np.broadcast_to(tree, (a.batchSize,) + tree.shape).copy())

    def batch(i0, i1):
        return Table(torch.from_numpy(x[i0:i1]), trees[:i1 - i0]), torch.from_numpy(y[i0:i1])

    for epoch in range(a.maxEpoch):
        model.training()
        perm = np.random.RandomState(epoch).permutation(n_train)
        x[:n_train], y[:n_train] = x[perm], y[perm]
        total = 0.0
        for i0 in range(0, n_train - a.batchSize + 1, a.batchSize):
            inp, tgt = batch(i0, i0 + a.batchSize)
            model.zeroGradParameters()
            out = model.forward(inp)
            total += float(crit.forward(out, tgt))
            model.backward(inp, crit.backward(out, tgt))
            om.update(w, g)
        print(f"epoch {epoch + 1}: loss {total / (n_train // a.batchSize):.4f}")
    model.evaluate()
    correct = count = 0
    for i0 in range(n_train, len(x) - a.batchSize + 1, a.batchSize):
        inp, tgt = batch(i0, i0 + a.batchSize)
        pred = model.forward(inp).argmax(dim=1) + 1
        correct += int((pred == tgt.long()).sum())
        count += tgt.numel()
    acc = correct / max(count, 1)
    print(f"root-node accuracy: {acc:.3f}")
    return acc


if __name__ == "__main__":
    main()
