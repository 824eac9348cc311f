"""Keras-style sentiment LSTM (reference: P/examples/keras/imdb_cnn_lstm.py — Embedding → (Conv1D →) LSTM → Dense(1,
sigmoid), binary cross-entropy). The IMDB archive is not bundled: the reviews are synthetic token sequences whose label
is decided by which of two "sentiment" vocab ranges dominates, with distractor tokens in between.

    python examples/keras_imdb_lstm.py [-b 32] [-e 4]
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

VOCAB, MAXLEN = 200, 30


def synthetic(n, seed):
    rs = np.random.RandomState(seed)
    x = rs.randint(21, VOCAB + 1, size=(n, MAXLEN))              # neutral tokens 21..200
    y = rs.randint(0, 2, size=n)
    for i in range(n):
        k = rs.randint(4, 9)
        pos = rs.choice(MAXLEN, size=k, replace=False)
        lo, hi = (1, 10) if y[i] == 1 else (11, 20)              # 1..10 positive words, 11..20 negative words
        x[i, pos] = rs.randint(lo, hi + 1, size=k)
    return x.astype(np.float32), y.astype(np.float32).reshape(-1, 1)


def build_model():
    import bigdl_b200.nn.keras as K

    m = K.Sequential()
    m.add(K.Embedding(VOCAB + 1, 24, input_shape=(MAXLEN,)))
    m.add(K.Convolution1D(32, 3, activation="relu"))
    m.add(K.MaxPooling1D(2))
    m.add(K.LSTM(32))
    m.add(K.Dense(1, activation="sigmoid"))
    return m


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("-b", "--batchSize", type=int, default=32)
    ap.add_argument("-e", "--nbEpoch", type=int, default=4)
    a = ap.parse_args(argv)
    from bigdl_b200.utils.engine import Engine
    from bigdl_b200.utils.random_generator import RNG

    Engine.init(device="cpu")
    RNG.setSeed(2)
    xtr, ytr = synthetic(1024, 0)
    xte, yte = synthetic(256, 1)
    model = build_model()
    model.compile(optimizer="adam", loss="binary_crossentropy", metrics=["accuracy"])
    model.fit(xtr, ytr, batch_size=a.batchSize, nb_epoch=a.nbEpoch, validation_data=(xte, yte))
    loss, acc = model.evaluate(xte, yte, batch_size=a.batchSize)
    print(f"test loss {loss:.4f}  accuracy {acc:.4f}")
    return acc


if __name__ == "__main__":
    main()
