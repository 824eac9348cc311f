"""Feature normalisation helpers of the Python API (P/dataset/transformer.py): ``normalizer(data, mean, std)``."""
import numpy as np


def normalizer(data, mean, std):
    return (np.asarray(data, dtype=np.float32) - mean) / std
