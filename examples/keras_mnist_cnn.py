"""Keras-style API example (reference: P/examples/keras/mnist_cnn.py and S/example/keras/LeNet.scala): a small CNN built
with bigdl_b200.nn.keras layers, ``compile`` / ``fit`` / ``evaluate``. MNIST idx files are read from --folder when
present; otherwise the synthetic digits of examples/lenet_local.py are used."""
from __future__ import annotations

import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def build_model(class_num=10):
    import bigdl_b200.nn.keras as K

    m = K.Sequential()
    m.add(K.Convolution2D(16, 3, 3, activation="relu", input_shape=(1, 28, 28)))
    m.add(K.MaxPooling2D())
    m.add(K.Convolution2D(32, 3, 3, activation="relu"))
    m.add(K.MaxPooling2D())
    m.add(K.Flatten())
    m.add(K.Dense(64, activation="relu"))
    m.add(K.Dropout(0.25))
    m.add(K.Dense(class_num, activation="softmax"))
    return m


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("-f", "--folder", default="/tmp/mnist")
    ap.add_argument("-b", "--batchSize", type=int, default=64)
    ap.add_argument("-e", "--nbEpoch", type=int, default=3)
    a = ap.parse_args(argv)
    from lenet_local import load

    from bigdl_b200.utils.engine import Engine

    Engine.init(device="cpu")
    xtr, ytr = load(a.folder, "train", 2048)
    xte, yte = load(a.folder, "test", 512)
    model = build_model()
    model.summary()
    model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    model.fit(xtr, ytr, batch_size=a.batchSize, nb_epoch=a.nbEpoch, validation_data=(xte, yte))
    loss, acc = model.evaluate(xte, yte, batch_size=a.batchSize)
    print(f"test loss {loss:.4f}  accuracy {acc:.4f}")
    return acc


if __name__ == "__main__":
    main()
