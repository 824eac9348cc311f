"""MNIST idx-file reader (P/dataset/mnist.py:30-130; Scala twin S/models/lenet/Utils.scala load). Files are looked up in
``train_dir`` (gz or plain); there is no network in this build, so a missing file is an error that names the expected
path instead of triggering a download."""
from __future__ import annotations

import gzip
import os
import struct

import numpy as np

TRAIN_MEAN, TRAIN_STD = 0.13066047740239506 * 255, 0.3081078 * 255
TEST_MEAN, TEST_STD = 0.13251460696903547 * 255, 0.31048024 * 255
_FILES = {"train": ("train-images-idx3-ubyte", "train-labels-idx1-ubyte"),
          "test": ("t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte")}


def _open(path):
    for p in (path, path + ".gz"):
        if os.path.exists(p):
            return gzip.open(p, "rb") if p.endswith(".gz") else open(p, "rb")
    raise FileNotFoundError(f"MNIST file not found: {path}[.gz] (no network access: place the idx files there)")


def extract_images(f) -> np.ndarray:
    magic, n, rows, cols = struct.unpack(">IIII", f.read(16))
    if magic != 2051:
        raise ValueError(f"Invalid magic number {magic} in MNIST image file")
    return np.frombuffer(f.read(n * rows * cols), dtype=np.uint8).reshape(n, rows, cols, 1)


def extract_labels(f) -> np.ndarray:
    magic, n = struct.unpack(">II", f.read(8))
    if magic != 2049:
        raise ValueError(f"Invalid magic number {magic} in MNIST label file")
    return np.frombuffer(f.read(n), dtype=np.uint8)


def read_data_sets(train_dir: str, data_type: str = "train"):
    """→ (images uint8 [N,28,28,1], labels uint8 [N]); labels are 0-based here and the callers add 1, as in the
    reference's lenet5.py."""
    img, lab = _FILES[data_type]
    with _open(os.path.join(train_dir, img)) as f:
        images = extract_images(f)
    with _open(os.path.join(train_dir, lab)) as f:
        labels = extract_labels(f)
    return images, labels


def write_idx(path: str, images: np.ndarray | None = None, labels: np.ndarray | None = None):
    """Inverse of the readers (used by tests and synthetic-data generators)."""
    with open(path, "wb") as f:
        if images is not None:
            n, r, c = images.shape[:3]
            f.write(struct.pack(">IIII", 2051, n, r, c) + np.ascontiguousarray(images, dtype=np.uint8).tobytes())
        else:
            f.write(struct.pack(">II", 2049, len(labels)) + np.ascontiguousarray(labels, dtype=np.uint8).tobytes())
