"""20-Newsgroups + GloVe loaders (P/dataset/news20.py:30-110): ``get_news20(dir)`` → [(text, 1-based label)],
``get_glove_w2v(dir, dim)`` → {word: vector}. Offline: the extracted folders must already exist under ``source_dir``."""
from __future__ import annotations

import os

CLASS_NUM = 20


def get_news20(source_dir: str = "/tmp/news20/"):
    root = os.path.join(source_dir, "20news-18828")
    if not os.path.isdir(root):
        raise FileNotFoundError(f"{root} not found (no network access: extract 20news-18828 there)")
    texts, label_id = [], 0
    for name in sorted(os.listdir(root)):
        path = os.path.join(root, name)
        label_id += 1
        if os.path.isdir(path):
            for fname in sorted(os.listdir(path)):
                if fname.isdigit():
                    with open(os.path.join(path, fname), encoding="latin-1") as f:
                        texts.append((f.read(), label_id))
    return texts


def get_glove_w2v(source_dir: str = "/tmp/news20/", dim: int = 100):
    path = os.path.join(source_dir, "glove.6B", f"glove.6B.{dim}d.txt")
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} not found (no network access)")
    out = {}
    with open(path, encoding="utf-8") as f:
        for line in f:
            parts = line.rstrip().split(" ")
            out[parts[0]] = [float(v) for v in parts[1:]]
    return out
