"""Text pipeline (S/dataset/text/*): sentence splitting / tokenising, start-end padding, frequency ``Dictionary``,
sentences → (data, label) index sequences shifted by one (language-model targets) → one-hot or index Samples.
The reference wraps OpenNLP models for splitting/tokenising and falls back to simple regex rules when no model file is
given (SentenceSplitter.scala:56-70, SentenceTokenizer.scala:53-68); the regex rules are what is implemented here."""
from __future__ import annotations

import os
import random
import re

import numpy as np
import torch

from .base import Sample, Transformer


class SentenceToken:
    """Start / end markers added to a sentence for language modelling (S/dataset/text/utils/Types.scala:25-28)."""
    start = "SENTENCESTART"
    end = "SENTENCEEND"


class Sentence:
    """A token-index sequence without a label (S/dataset/text/Types.scala ``Sentence``)."""

    def __init__(self, data=None):
        self._data = list(data) if data is not None else []

    def data(self): return self._data
    def dataLength(self): return len(self._data)
    def getData(self, i): return self._data[i]


class LabeledSentence(Sentence):
    def __init__(self, data=None, label=None):
        self._data = list(data) if data is not None else []
        self._label = list(label) if label is not None else []

    def copyToData(self, storage, offset: int = 0):
        storage[offset:offset + len(self._data)] = self._data
        return storage

    def copyToLabel(self, storage, offset: int = 0):
        storage[offset:offset + len(self._label)] = self._label
        return storage

    def clone(self):
        return LabeledSentence(self._data, self._label)

    def data(self): return self._data
    def label(self): return self._label
    def dataLength(self): return len(self._data)
    def labelLength(self): return len(self._label)
    def getData(self, i): return self._data[i]
    def getLabel(self, i): return self._label[i]

    def copy(self, data, label):
        self._data, self._label = list(data), list(label)
        return self


class SentenceSplitter(Transformer):
    """Paragraph → sentences on [.!?] boundaries (the reference's no-model fallback splits on '.' only)."""

    def __init__(self, sent_file: str | None = None):
        self.sent_file = sent_file

    def __call__(self, it):
        for text in it:
            yield [s.strip() for s in re.split(r"(?<=[.!?])\s+", text.strip()) if s.strip()]


class SentenceTokenizer(Transformer):
    """Sentence → tokens. No-model fallback: whitespace split (SentenceTokenizer.scala:62-66)."""

    def __init__(self, token_file: str | None = None):
        self.token_file = token_file

    def __call__(self, it):
        for s in it:
            yield s.split()


class SentenceBiPadding(Transformer):
    """Wrap each sentence in SENTENCESTART … SENTENCEEND (SentenceBiPadding.scala:28-43)."""

    def __init__(self, start: str | None = None, end: str | None = None):
        self.start, self.end = start or SentenceToken.start, end or SentenceToken.end

    def __call__(self, it):
        for s in it:
            yield f"{self.start} {s} {self.end}"


class Dictionary:
    """Vocabulary of the ``vocab_size`` most frequent words; everything else is "discarded" and maps to index
    ``vocabSize`` (Dictionary.scala:30-215). Indices are 0-based here as in the reference (the +1 happens in
    LabeledSentenceToSample)."""

    def __init__(self, source=None, vocab_size: int | None = None):
        self._vocab, self._discard, self._w2i, self._i2w = [], [], {}, {}
        if isinstance(source, str) and os.path.isdir(source):
            self._load(source)
        elif source is not None:
            counts: dict[str, int] = {}
            for item in source:
                for w in ([item] if isinstance(item, str) else item):
                    counts[w] = counts.get(w, 0) + 1
            freq = sorted(counts.items(), key=lambda kv: kv[1])          # ascending, stable — as sortBy(_._2)
            n = min(vocab_size if vocab_size is not None else len(freq), len(freq))
            self._vocab = [w for w, _ in freq[len(freq) - n:]]
            self._discard = [w for w, _ in freq[:len(freq) - n]]
            self._w2i = {w: i for i, w in enumerate(self._vocab)}
            self._i2w = {i: w for w, i in self._w2i.items()}

    def _load(self, d: str):
        for line in open(os.path.join(d, "dictionary.txt"), encoding="utf-8"):
            line = line.rstrip("\n")
            if not line:
                continue
            w, i = line.rsplit("->", 1)
            self._w2i[w.rstrip(" ")] = int(i.strip())
        self._i2w = {i: w for w, i in self._w2i.items()}
        self._vocab = list(self._w2i)
        self._discard = [l.rstrip("\n") for l in open(os.path.join(d, "discard.txt"), encoding="utf-8") if l.strip()]

    def getVocabSize(self): return len(self._vocab)
    def getDiscardSize(self): return len(self._discard)
    def word2Index(self): return dict(self._w2i)
    def index2Word(self): return dict(self._i2w)
    def vocabulary(self): return list(self._vocab)
    def discardVocab(self): return list(self._discard)

    def getIndex(self, word: str) -> int:
        return self._w2i.get(word, len(self._vocab))

    def getWord(self, index) -> str:
        i = int(index)
        if i in self._i2w:
            return self._i2w[i]
        if self._discard:
            return random.choice(self._discard)
        return self._i2w[random.randrange(len(self._vocab))]

    def save(self, folder: str):
        if not os.path.isdir(folder):
            raise ValueError("Dictionary: saveFolder should be a directory.")
        with open(os.path.join(folder, "dictionary.txt"), "w", encoding="utf-8") as f:
            f.write("\n".join(f"{w} -> {i}" for w, i in self._w2i.items()))
        with open(os.path.join(folder, "discard.txt"), "w", encoding="utf-8") as f:
            f.write("\n".join(self._discard))


class TextToLabeledSentence(Transformer):
    """tokens → indices; data = all but last, label = all but first (TextToLabeledSentence.scala:46-66)."""

    def __init__(self, dictionary: Dictionary):
        self.dictionary = dictionary

    def __call__(self, it):
        for sent in it:
            idx = [float(self.dictionary.getIndex(w)) for w in sent]
            yield LabeledSentence(idx[:-1], idx[1:])


class TextToSentenceWithSteps(Transformer):
    def __init__(self, num_steps: int):
        self.n = num_steps

    def __call__(self, it):
        for sent in it:
            if len(sent) < self.n + 1:
                raise ValueError(f"input sentence length should be numSteps + 1, sentence.length = {len(sent)}, "
                                 f"numSteps = {self.n}")
            yield LabeledSentence(sent[:self.n], sent[1:self.n + 1])


class LabeledSentenceToSample(Transformer):
    """(LabeledSentenceToSample.scala:74-150) one-hot features [len, vocab] with END-token padding and 1-based labels
    padded with the START token, or raw index features when ``one_hot`` is False."""

    def __init__(self, vocab_length: int, fix_data_length: int | None = None, fix_label_length: int | None = None,
                 one_hot: bool = True):
        self.vocab, self.fd, self.fl, self.one_hot = vocab_length, fix_data_length, fix_label_length, one_hot

    def __call__(self, it):
        for s in it:
            dl = self.fd if self.fd is not None else s.dataLength()
            ll = self.fl if self.fl is not None else s.labelLength()
            if self.one_hot:
                feat = torch.zeros(dl, self.vocab)
                start = s.getData(0)
                end = 0 if ll == 1 else int(s.getLabel(s.labelLength() - 1))
                for i in range(s.dataLength()):
                    feat[i, int(s.getData(i))] = 1.0
                for i in range(s.dataLength(), dl):
                    feat[i, end] = 1.0
                label = torch.zeros(ll)
                for i in range(s.labelLength()):
                    label[i] = s.getLabel(i) + 1.0
                for i in range(s.labelLength(), ll):
                    label[i] = start + 1.0
            else:
                feat = torch.zeros(dl)
                label = torch.zeros(ll)
                feat[:min(dl, s.dataLength())] = torch.tensor(s.data()[:dl])
                label[:min(ll, s.labelLength())] = torch.tensor(s.label()[:ll])
            yield Sample(feat, label)


def read_sentences(path: str) -> list[str]:
    with open(path, encoding="utf-8") as f:
        return [l.strip() for l in f if l.strip()]


from .base import PaddingParam  # noqa: E402,F401  (text pipelines pad variable-length sentences with it)
