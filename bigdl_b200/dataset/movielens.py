"""MovieLens-1M ratings loader (P/dataset/movielens.py): ``read_data_sets(dir)`` → int array [N,3] of
(user, item, rating); ``get_id_pairs`` / ``get_id_ratings`` helpers."""
from __future__ import annotations

import os

import numpy as np


def read_data_sets(data_dir: str) -> np.ndarray:
    path = os.path.join(data_dir, "ml-1m", "ratings.dat")
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} not found (no network access: extract ml-1m.zip there)")
    rows = []
    with open(path) as f:
        for line in f:
            u, i, r, _ = line.strip().split("::")
            rows.append((int(u), int(i), int(r)))
    return np.asarray(rows, dtype=np.int64)


def get_id_pairs(data_dir: str) -> np.ndarray:
    return read_data_sets(data_dir)[:, 0:2]


def get_id_ratings(data_dir: str) -> np.ndarray:
    return read_data_sets(data_dir)[:, 0:3]
