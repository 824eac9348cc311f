"""TFRecord files and tf.train.Example records (reference: S/utils/tf/{TFRecordWriter,TFRecordIterator,
FixedLengthRecordReader,TFRecordInputFormat,TFRecordOutputFormat}.scala). Framing and masked CRC32C are done by the C++
host library (csrc/host/checksum.cpp)."""
from __future__ import annotations

import struct

import numpy as np

from ..utils import native


class TFRecordWriter:
    def __init__(self, path: str):
        self._f = open(path, "wb")

    def write(self, record: bytes):
        self._f.write(native.tfrecord_frame(record))

    def close(self):
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class TFRecordIterator:
    """Iterate the payloads of a TFRecord file; ``verify=False`` skips the CRC checks like the reference iterator."""

    def __init__(self, path: str, verify: bool = True):
        with open(path, "rb") as f:
            self._buf = f.read()
        self._verify = verify

    def __iter__(self):
        return native.tfrecord_iter(self._buf, self._verify)


class FixedLengthRecordReader:
    """Fixed-size records with optional header/footer bytes and a hop between records (CIFAR-10 binary layout)."""

    def __init__(self, path: str, footer_bytes: int = 0, header_bytes: int = 0, hop_bytes: int = 0,
                 record_bytes: int = 1):
        self.path, self.footer, self.header, self.hop, self.record = path, footer_bytes, header_bytes, hop_bytes, record_bytes

    def __iter__(self):
        with open(self.path, "rb") as f:
            data = f.read()
        end = len(data) - self.footer
        pos = self.header
        step = self.record + self.hop
        while pos + self.record <= end:
            yield data[pos:pos + self.record]
            pos += step


# ---- tf.train.Example encoding (decoding lives in nn/tf.py parse_example)
def _ld(field: int, payload: bytes) -> bytes:
    from ..visualization.proto import _varint

    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def encode_example(features: dict) -> bytes:
    """{name: bytes | str | list[float] | list[int]} → serialized tf.train.Example."""
    from ..visualization.proto import _varint

    entries = b""
    for name, v in features.items():
        if isinstance(v, (bytes, str)):
            v = [v]
        v = list(v)
        if v and isinstance(v[0], (bytes, str)):
            items = b"".join(_ld(1, x if isinstance(x, bytes) else x.encode()) for x in v)
            feat = _ld(1, items)
        elif v and isinstance(v[0], (float, np.floating)):
            feat = _ld(2, _ld(1, struct.pack(f"<{len(v)}f", *v)))
        else:
            feat = _ld(3, _ld(1, b"".join(_varint(int(x)) for x in v)))
        entries += _ld(1, _ld(1, name.encode()) + _ld(2, feat))
    return _ld(1, entries)
