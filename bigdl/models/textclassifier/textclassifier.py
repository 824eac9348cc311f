"""pyspark/bigdl/models/textclassifier/textclassifier.py: the CNN / LSTM / GRU news20 text classifier helpers."""
import re

import numpy as np


def text_to_words(review_text):
    letters_only = re.sub("[^a-zA-Z]", " ", review_text)
    return letters_only.lower().split()


def analyze_texts(texts):
    """[(text, label)] -> [(word, (1-based frequency rank, count))], most frequent first."""
    counts = {}
    for text, _ in texts:
        for w in text_to_words(text):
            counts[w] = counts.get(w, 0) + 1
    ordered = sorted(counts.items(), key=lambda kv: -kv[1])
    return [(w, (i + 1, c)) for i, (w, c) in enumerate(ordered)]


def pad(l, fill_value, width):
    return l[:width] if len(l) >= width else l + [fill_value] * (width - len(l))


def to_vec(token, b_w2v, embedding_dim):
    return b_w2v[token] if token in b_w2v else pad([], 0, embedding_dim)


def to_sample(vectors, label, embedding_dim):
    from bigdl_b200.dataset.base import Sample
    import torch
    flat = np.asarray(vectors, dtype=np.float32).reshape(-1, embedding_dim)
    return Sample(torch.from_numpy(flat), torch.tensor([float(label)]))


def build_model(class_num, model_type="cnn", embedding_dim=50, sequence_len=50, p=0.0):
    from bigdl_b200.nn import layer as L
    model = L.Sequential()
    if model_type.lower() == "cnn":
        model.add(L.TemporalConvolution(embedding_dim, 256, 5)).add(L.ReLU()).add(L.TemporalMaxPooling(sequence_len - 5 + 1))
        model.add(L.Squeeze(2))
    elif model_type.lower() == "lstm":
        model.add(L.Recurrent().add(L.LSTM(embedding_dim, 256, p))).add(L.Select(2, -1))
    elif model_type.lower() == "gru":
        model.add(L.Recurrent().add(L.GRU(embedding_dim, 256, p))).add(L.Select(2, -1))
    else:
        raise ValueError("model can only be cnn, lstm, or gru")
    model.add(L.Linear(256, 128)).add(L.Dropout(0.2)).add(L.ReLU()).add(L.Linear(128, class_num)).add(L.LogSoftMax())
    return model
