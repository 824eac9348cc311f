"""Minimal protobuf wire encoder/decoder for the TensorBoard messages (Event, Summary, Summary.Value, HistogramProto).
Field numbers follow tensorflow/core/util/event.proto and framework/summary.proto, the same messages the reference
builds with generated Java classes (S/visualization/tensorboard/{EventWriter,Summary}.scala)."""
from __future__ import annotations

import struct


def _varint(v: int) -> bytes:
    v &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(field: int, wt: int) -> bytes:
    return _varint((field << 3) | wt)


def f_varint(field, v): return _key(field, 0) + _varint(int(v))
def f_double(field, v): return _key(field, 1) + struct.pack("<d", float(v))
def f_float(field, v): return _key(field, 5) + struct.pack("<f", float(v))
def f_bytes(field, b): return _key(field, 2) + _varint(len(b)) + bytes(b)
def f_string(field, s): return f_bytes(field, s.encode("utf-8"))
def f_packed_double(field, vals): return f_bytes(field, b"".join(struct.pack("<d", float(v)) for v in vals))


def histogram(mn, mx, num, s, sq, limits, counts) -> bytes:
    return (f_double(1, mn) + f_double(2, mx) + f_double(3, num) + f_double(4, s) + f_double(5, sq) +
            f_packed_double(6, limits) + f_packed_double(7, counts))


def summary_scalar(tag: str, value: float) -> bytes:
    return f_bytes(1, f_string(1, tag) + f_float(2, value))


def summary_histogram(tag: str, histo: bytes) -> bytes:
    return f_bytes(1, f_string(1, tag) + f_bytes(5, histo))


def event(wall_time: float, step: int | None = None, file_version: str | None = None, summary: bytes | None = None,
          graph_def: bytes | None = None) -> bytes:
    out = f_double(1, wall_time)
    if step is not None:
        out += f_varint(2, step)
    if file_version is not None:
        out += f_string(3, file_version)
    if graph_def is not None:
        out += f_bytes(4, graph_def)
    if summary is not None:
        out += f_bytes(5, summary)
    return out


# ------------------------------------------------------------------------------------------------ decoding
def parse(buf: bytes):
    """Yield (field, wire_type, value) triples of one message."""
    i, n = 0, len(buf)
    while i < n:
        key, i = _read_varint(buf, i)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _read_varint(buf, i)
        elif wt == 1:
            v = struct.unpack_from("<d", buf, i)[0]
            i += 8
        elif wt == 5:
            v = struct.unpack_from("<f", buf, i)[0]
            i += 4
        elif wt == 2:
            ln, i = _read_varint(buf, i)
            v = buf[i:i + ln]
            i += ln
        else:
            raise ValueError(f"unsupported wire type {wt}")
        yield field, wt, v


def _read_varint(buf, i):
    shift = v = 0
    while True:
        b = buf[i]
        i += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, i
        shift += 7


def decode_event(buf: bytes) -> dict:
    ev = {"values": []}
    for f, wt, v in parse(buf):
        if f == 1:
            ev["wall_time"] = v
        elif f == 2:
            ev["step"] = v
        elif f == 3:
            ev["file_version"] = v.decode()
        elif f == 4:
            ev["graph_def"] = v
        elif f == 5:
            for f2, _, val in parse(v):
                if f2 == 1:
                    item = {}
                    for f3, _, x in parse(val):
                        if f3 == 1:
                            item["tag"] = x.decode()
                        elif f3 == 2:
                            item["simple_value"] = x
                        elif f3 == 5:
                            item["histo"] = {k: y for k, _, y in parse(x) if k <= 5}
                    ev["values"].append(item)
    return ev
