"""Hadoop SequenceFile (version 6, uncompressed, Text/Text records) reader and writer, and the ImageNet seq-file
dataset built on it.

Reference: S/dataset/image/BGRImgToLocalSeqFile.scala:36-76 (key = "name\\nlabel" or "label"; value = 4-byte BE width,
4-byte BE height, then BGR bytes), S/dataset/image/LocalSeqFileToBytes.scala, S/dataset/DataSet.scala:486-560
(``SeqFileFolder``: ``readLabel/readName`` split the key on "\\n"), S/models/utils/ImageNetSeqFileGenerator.scala.
The reference uses hadoop-common for the container format; this is a from-scratch implementation of the on-disk layout
(header "SEQ\\x06", class names, 16-byte sync marker, [recordLen][keyLen][key][value] records with a sync escape of −1
every ≥2000 bytes) so files written by either side can be read by the other."""
from __future__ import annotations

import glob
import os
import struct

import numpy as np

_SYNC_INTERVAL = 2000
_TEXT = b"org.apache.hadoop.io.Text"


def _write_vint(v: int) -> bytes:
    if -112 <= v <= 127:
        return struct.pack("b", v)
    ln = -112
    if v < 0:
        v = ~v
        ln = -120
    tmp = v
    while tmp != 0:
        tmp >>= 8
        ln -= 1
    out = struct.pack("b", ln)
    n = -(ln + 120) if ln < -120 else -(ln + 112)
    for i in range(n, 0, -1):
        out += bytes([(v >> (8 * (i - 1))) & 0xFF])
    return out


def _read_vint(buf, pos: int):
    first = struct.unpack_from("b", buf, pos)[0]
    if first >= -112:
        return first, pos + 1
    n = (-119 - first) if first < -120 else (-111 - first)
    v = 0
    for i in range(n - 1):
        v = (v << 8) | buf[pos + 1 + i]
    if first < -120:
        v = ~v
    return v, pos + n


def _text(b: bytes) -> bytes:
    return _write_vint(len(b)) + b


class SequenceFileWriter:
    def __init__(self, path: str, key_class: bytes = _TEXT, value_class: bytes = _TEXT, sync: bytes | None = None):
        self._f = open(path, "wb")
        self._sync = sync or os.urandom(16)
        self._f.write(b"SEQ\x06" + _text(key_class) + _text(value_class) + b"\x00\x00" + struct.pack(">i", 0) + self._sync)
        self._last_sync = self._f.tell()

    def append(self, key: bytes, value: bytes):
        if self._f.tell() >= self._last_sync + _SYNC_INTERVAL:
            self._f.write(struct.pack(">i", -1) + self._sync)
            self._last_sync = self._f.tell()
        k, v = _text(key), _text(value)
        self._f.write(struct.pack(">ii", len(k) + len(v), len(k)) + k + v)

    def close(self):
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def read_sequence_file(path: str):
    """Yield (key bytes, value bytes) of every record (Text payloads, vint length prefix stripped)."""
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:3] != b"SEQ":
        raise ValueError(f"{path} is not a SequenceFile")
    if buf[3] != 6:
        raise ValueError(f"unsupported SequenceFile version {buf[3]}")
    pos = 4
    n, pos = _read_vint(buf, pos)
    pos += n
    n, pos = _read_vint(buf, pos)
    pos += n
    compressed, block = buf[pos], buf[pos + 1]
    pos += 2
    if compressed or block:
        raise ValueError("compressed SequenceFiles are not supported")
    meta = struct.unpack_from(">i", buf, pos)[0]
    pos += 4
    for _ in range(meta * 2):
        n, pos = _read_vint(buf, pos)
        pos += n
    sync = buf[pos:pos + 16]
    pos += 16
    end = len(buf)
    while pos < end:
        rec_len = struct.unpack_from(">i", buf, pos)[0]
        pos += 4
        if rec_len == -1:
            if buf[pos:pos + 16] != sync:
                raise ValueError("corrupt SequenceFile: bad sync marker")
            pos += 16
            continue
        key_len = struct.unpack_from(">i", buf, pos)[0]
        pos += 4
        kn, kp = _read_vint(buf, pos)
        key = bytes(buf[kp:kp + kn])
        vpos = pos + key_len
        vn, vp = _read_vint(buf, vpos)
        value = bytes(buf[vp:vp + vn])
        pos += rec_len
        yield key, value


def read_label(key: bytes) -> str:
    parts = key.decode("utf-8").split("\n")
    return parts[0] if len(parts) == 1 else parts[1]


def read_name(key: bytes) -> str:
    parts = key.decode("utf-8").split("\n")
    if len(parts) < 2:
        raise ValueError("key in seq file only contains label, no name")
    return parts[0]


def encode_image_value(bgr: np.ndarray) -> bytes:
    h, w = bgr.shape[:2]
    return struct.pack(">ii", w, h) + np.ascontiguousarray(bgr, dtype=np.uint8).tobytes()


def decode_image_value(value: bytes) -> np.ndarray:
    w, h = struct.unpack_from(">ii", value, 0)
    return np.frombuffer(value, dtype=np.uint8, offset=8, count=w * h * 3).reshape(h, w, 3)


def write_image_seq_files(images, base_name: str, block_size: int = 12800, has_name: bool = False) -> list[str]:
    """``images``: iterable of (bgr uint8 HxWx3, label[, name]). Writes ``base_name_<i>.seq`` blocks."""
    files, idx, w, n = [], 0, None, 0
    for item in images:
        if w is None:
            fn = f"{base_name}_{idx}.seq"
            files.append(fn)
            w = SequenceFileWriter(fn)
        img, label = item[0], int(item[1])
        key = (f"{item[2]}\n{label}" if has_name else f"{label}").encode()
        w.append(key, encode_image_value(img))
        n += 1
        if n == block_size:
            w.close()
            w, n, idx = None, 0, idx + 1
    if w is not None:
        w.close()
    return files


def seq_file_paths(folder: str) -> list[str]:
    return sorted(glob.glob(os.path.join(folder, "*.seq")))


def seq_file_dataset(path: str, class_num: int = 1000, distributed: bool = True):
    """DataSet of ``ByteRecord``s (raw value bytes + float label) over every ``*.seq`` under ``path``; labels above
    ``class_num`` are filtered like SeqFileFolder.files(url, sc, classNum)."""
    from .base import DataSet
    from .image import ByteRecord

    recs = []
    for fn in seq_file_paths(path):
        for k, v in read_sequence_file(fn):
            lab = float(read_label(k))
            if lab <= class_num:
                recs.append(ByteRecord(v, lab))
    return DataSet.array(recs, distributed)
