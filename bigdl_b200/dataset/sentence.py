"""Sentence helpers of the Python API (P/dataset/sentence.py): read a text file into sentences, pad with start/end
tokens, tokenise."""
from __future__ import annotations

import re


def read_localfile(file_name: str) -> list[str]:
    with open(file_name, encoding="utf-8") as f:
        return [l.strip() for l in f if l.strip()]


def sentences_split(line: str) -> list[str]:
    return [s.strip() for s in re.split(r"(?<=[.!?])\s+", line) if s.strip()]


def sentences_bipadding(sent: str) -> str:
    return "SENTENCESTART " + sent + " SENTENCEEND"


def sentence_tokenizer(sentences: str) -> list[str]:
    return [t for t in re.findall(r"[A-Za-z0-9']+|[^\sA-Za-z0-9]", sentences)]
