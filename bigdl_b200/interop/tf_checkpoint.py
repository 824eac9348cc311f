"""TensorFlow checkpoints without TensorFlow.

The reference's ``scripts/export_tf_checkpoint.py`` / ``P/util/tf_utils.py`` (dump_model, export_checkpoint,
merge_checkpoint, get_path) drive a live TensorFlow session to pull variables out of a checkpoint and the GraphDef out
of a ``.meta`` file. This module does the same from the FILES:

* a V2 checkpoint ("tensor bundle": ``<prefix>.index`` + ``<prefix>.data-0000N-of-0000M``) — the index is a
  LevelDB-format sorted string table whose values are ``BundleEntryProto``s, the data files hold the raw little-endian
  tensor bytes; ``read_checkpoint`` / ``write_checkpoint`` implement both directions (the writer exists so checkpoints
  can be produced here for tests and for handing weights back to TensorFlow users);
* ``<prefix>.meta`` (a ``MetaGraphDef``) or a SavedModel folder (``saved_model.pb`` + ``variables/``) — the GraphDef is
  cut out of the wire format;
* ``freeze`` turns ``VariableV2`` / ``VarHandleOp`` nodes into ``Const`` nodes holding the checkpoint values and keeps
  only what the requested outputs need (TensorFlow's ``freeze_graph``).

Table format: blocks of prefix-compressed (shared, non_shared, value_len, key_delta, value) entries followed by the
restart offsets and their count; each block is followed by a 1-byte compression tag and a masked CRC-32C; the 48-byte
footer holds the metaindex and index block handles and the magic number 0xdb4775248b80fb57."""
from __future__ import annotations

import os
import struct

import numpy as np

from ..utils.native import crc32c
from ..visualization.proto import _read_varint, _varint, f_bytes, f_varint
from . import tf_pb as P

_MAGIC = 0xDB4775248B80FB57
_MASK_DELTA = 0xA282EAD8
_RESTART_INTERVAL = 16
_BLOCK_SIZE = 4096

_NP_OF = dict(P.NP_OF)
_DT_OF_NP = dict(P.DT_OF_NP)


def parse(buf: bytes):
    """(field, wire type, value) of one message; fixed-width fields come back as raw bytes."""
    i, n = 0, len(buf)
    while i < n:
        key, i = _read_varint(buf, i)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _read_varint(buf, i)
        elif wt == 1:
            v, i = buf[i:i + 8], i + 8
        elif wt == 5:
            v, i = buf[i:i + 4], i + 4
        elif wt == 2:
            ln, i = _read_varint(buf, i)
            v, i = buf[i:i + ln], i + ln
        else:
            raise ValueError(f"unsupported wire type {wt}")
        yield field, wt, v


def _mask(crc: int) -> int:
    return (((crc >> 15) | (crc << 17)) + _MASK_DELTA) & 0xFFFFFFFF


def _unmask(m: int) -> int:
    rot = (m - _MASK_DELTA) & 0xFFFFFFFF
    return ((rot >> 17) | (rot << 15)) & 0xFFFFFFFF


# ------------------------------------------------------------------------------------------------ sorted string table
def _read_block(buf: bytes, offset: int, size: int, verify: bool = True) -> bytes:
    body, tag = buf[offset:offset + size], buf[offset + size]
    if verify:
        stored = struct.unpack_from("<I", buf, offset + size + 1)[0]
        if _unmask(stored) != crc32c(buf[offset:offset + size + 1]):
            raise ValueError("checkpoint index: block checksum mismatch")
    if tag == 0:
        return body
    if tag == 1:
        return _snappy_decompress(body)
    raise ValueError(f"checkpoint index: unknown block compression {tag}")


def _snappy_decompress(data: bytes) -> bytes:
    """Raw Snappy block format (literal / copy tags) — TensorFlow only uses it for the index when built with snappy."""
    n, i = _read_varint(data, 0)
    out = bytearray()
    while i < len(data):
        tag = data[i]
        i += 1
        kind = tag & 3
        if kind == 0:
            ln = tag >> 2
            if ln >= 60:
                nb = ln - 59
                ln = int.from_bytes(data[i:i + nb], "little")
                i += nb
            ln += 1
            out += data[i:i + ln]
            i += ln
            continue
        if kind == 1:
            ln = ((tag >> 2) & 7) + 4
            off = ((tag >> 5) << 8) | data[i]
            i += 1
        elif kind == 2:
            ln = (tag >> 2) + 1
            off = int.from_bytes(data[i:i + 2], "little")
            i += 2
        else:
            ln = (tag >> 2) + 1
            off = int.from_bytes(data[i:i + 4], "little")
            i += 4
        if off == 0 or off > len(out):
            raise ValueError("corrupt snappy data")
        for _ in range(ln):               # overlapping copies are allowed
            out.append(out[-off])
    if len(out) != n:
        raise ValueError("corrupt snappy data: length mismatch")
    return bytes(out)


def _block_entries(block: bytes):
    n_restarts = struct.unpack_from("<I", block, len(block) - 4)[0]
    end = len(block) - 4 - 4 * n_restarts
    pos, key = 0, b""
    while pos < end:
        shared, pos = _read_varint(block, pos)
        non_shared, pos = _read_varint(block, pos)
        vlen, pos = _read_varint(block, pos)
        key = key[:shared] + block[pos:pos + non_shared]
        pos += non_shared
        yield key, block[pos:pos + vlen]
        pos += vlen


def _handle(buf: bytes, pos: int):
    off, pos = _read_varint(buf, pos)
    size, pos = _read_varint(buf, pos)
    return off, size, pos


def read_table(path: str, verify: bool = True) -> list[tuple[bytes, bytes]]:
    """All (key, value) pairs of a LevelDB-format table file, in key order."""
    with open(path, "rb") as f:
        buf = f.read()
    if len(buf) < 48 or struct.unpack_from("<Q", buf, len(buf) - 8)[0] != _MAGIC:
        raise ValueError(f"{path} is not a TensorFlow checkpoint index (bad magic number)")
    footer = buf[-48:]
    _, _, p = _handle(footer, 0)                       # metaindex: unused by tensor bundles
    ioff, isize, _ = _handle(footer, p)
    out = []
    for _, hv in _block_entries(_read_block(buf, ioff, isize, verify)):
        doff, dsize, _ = _handle(hv, 0)
        out.extend(_block_entries(_read_block(buf, doff, dsize, verify)))
    return out


class _BlockBuilder:
    def __init__(self):
        self.buf, self.restarts, self.count, self.last = bytearray(), [0], 0, b""

    def add(self, key: bytes, value: bytes):
        shared = 0
        if self.count < _RESTART_INTERVAL:
            m = min(len(self.last), len(key))
            while shared < m and self.last[shared] == key[shared]:
                shared += 1
        else:
            self.restarts.append(len(self.buf))
            self.count = 0
        self.buf += _varint(shared) + _varint(len(key) - shared) + _varint(len(value)) + key[shared:] + value
        self.last = key
        self.count += 1

    def finish(self) -> bytes:
        return bytes(self.buf) + b"".join(struct.pack("<I", r) for r in self.restarts) + struct.pack("<I", len(self.restarts))

    def size(self) -> int:
        return len(self.buf) + 4 * len(self.restarts) + 4

    def empty(self) -> bool:
        return not self.buf


def write_table(path: str, items):
    """Write sorted (key, value) byte pairs as an uncompressed table file."""
    items = sorted(items, key=lambda kv: kv[0])
    out = bytearray()
    index = _BlockBuilder()

    def emit(block: bytes):
        off = len(out)
        out.extend(block)
        out.append(0)                                                   # kNoCompression
        out.extend(struct.pack("<I", _mask(crc32c(block + b"\x00"))))
        return off, len(block)

    cur = _BlockBuilder()
    for i, (k, v) in enumerate(items):
        cur.add(k, v)
        if cur.size() >= _BLOCK_SIZE or i == len(items) - 1:
            off, size = emit(cur.finish())
            index.add(cur.last, _varint(off) + _varint(size))           # the last key is a valid separator
            cur = _BlockBuilder()
    moff, msize = emit(_BlockBuilder().finish())                        # empty metaindex block
    ioff, isize = emit(index.finish())
    footer = _varint(moff) + _varint(msize) + _varint(ioff) + _varint(isize)
    out.extend(footer + b"\x00" * (40 - len(footer)) + struct.pack("<Q", _MAGIC))
    with open(path, "wb") as f:
        f.write(bytes(out))


# ------------------------------------------------------------------------------------------------ tensor bundle
def _shape_of(buf: bytes) -> list[int]:
    dims = []
    for f, _, v in parse(buf):
        if f == 2:                                   # TensorShapeProto.dim
            size = 0
            for f2, _, v2 in parse(v):
                if f2 == 1:
                    size = v2 - (1 << 64) if v2 >= 1 << 63 else v2
            dims.append(int(size))
    return dims


def _entry(buf: bytes) -> dict:
    e = {"dtype": 0, "shape": [], "shard_id": 0, "offset": 0, "size": 0, "crc32c": 0, "slices": 0}
    for f, wt, v in parse(buf):
        if f == 1:
            e["dtype"] = int(v)
        elif f == 2:
            e["shape"] = _shape_of(v)
        elif f == 3:
            e["shard_id"] = int(v)
        elif f == 4:
            e["offset"] = int(v)
        elif f == 5:
            e["size"] = int(v)
        elif f == 6:
            e["crc32c"] = struct.unpack("<I", v)[0] if isinstance(v, (bytes, bytearray)) else int(v)
        elif f == 7:
            e["slices"] += 1
    return e


def _data_file(prefix: str, shard: int, num_shards: int) -> str:
    return f"{prefix}.data-{shard:05d}-of-{num_shards:05d}"


def resolve_prefix(path: str) -> str:
    """Accepts a checkpoint prefix, its ``.index`` / ``.meta`` / ``.data-…`` file, a folder holding a ``checkpoint``
    state file or a SavedModel folder."""
    if os.path.isdir(path):
        if os.path.exists(os.path.join(path, "variables", "variables.index")):
            return os.path.join(path, "variables", "variables")
        state = os.path.join(path, "checkpoint")
        if os.path.exists(state):
            for line in open(state, encoding="utf-8"):
                if line.startswith("model_checkpoint_path:"):
                    name = line.split(":", 1)[1].strip().strip('"')
                    return name if os.path.isabs(name) else os.path.join(path, name)
        idx = sorted(f for f in os.listdir(path) if f.endswith(".index"))
        if idx:
            return os.path.join(path, idx[-1][:-6])
        raise FileNotFoundError(f"no checkpoint found in {path}")
    for suffix in (".index", ".meta"):
        if path.endswith(suffix):
            return path[:-len(suffix)]
    if ".data-" in os.path.basename(path):
        return path[:path.rindex(".data-")]
    return path


def list_variables(prefix: str) -> list[tuple[str, list[int], np.dtype | None]]:
    """(name, shape, dtype) of every tensor of the checkpoint (``tf.train.list_variables``)."""
    prefix = resolve_prefix(prefix)
    out = []
    for k, v in read_table(prefix + ".index"):
        if k:
            e = _entry(v)
            dt = _NP_OF.get(e["dtype"])
            out.append((k.decode(), e["shape"], np.dtype(dt) if dt is not None else (np.dtype(object) if e["dtype"] == 7 else None)))
    return out


def read_checkpoint(prefix: str, names=None, verify: bool = True) -> dict[str, np.ndarray]:
    """{variable name: array} of a V2 checkpoint. ``names`` restricts the tensors read; checksums of the index blocks
    and of every tensor are verified unless ``verify`` is False."""
    prefix = resolve_prefix(prefix)
    table = read_table(prefix + ".index", verify)
    num_shards, little = 1, True
    for k, v in table:
        if k == b"":
            for f, _, x in parse(v):
                if f == 1:
                    num_shards = int(x)
                elif f == 2:
                    little = int(x) == 0
    if not little:
        raise ValueError("big-endian checkpoints are not supported")
    wanted = set(names) if names is not None else None
    files: dict[int, bytes] = {}
    out: dict[str, np.ndarray] = {}
    for k, v in table:
        if not k:
            continue
        name = k.decode()
        if wanted is not None and name not in wanted:
            continue
        e = _entry(v)
        if e["slices"]:
            raise ValueError(f"{name}: partitioned (sliced) variables are not supported")
        if e["shard_id"] not in files:
            with open(_data_file(prefix, e["shard_id"], num_shards), "rb") as f:
                files[e["shard_id"]] = f.read()
        raw = files[e["shard_id"]][e["offset"]:e["offset"] + e["size"]]
        if len(raw) != e["size"]:
            raise ValueError(f"{name}: data file is truncated")
        if e["dtype"] == 7:
            out[name] = _decode_strings(raw, e["shape"], e["crc32c"], verify, name)
            continue
        if verify and _unmask(e["crc32c"]) != crc32c(raw):
            raise ValueError(f"{name}: tensor checksum mismatch")
        if e["dtype"] == 14:                                           # bfloat16 → float32
            arr = (np.frombuffer(raw, "<u2").astype(np.uint32) << 16).view(np.float32)
        else:
            dt = _NP_OF.get(e["dtype"])
            if dt is None:
                raise ValueError(f"{name}: unsupported dtype enum {e['dtype']}")
            arr = np.frombuffer(raw, np.dtype(dt).newbyteorder("<"))
        out[name] = arr.reshape(e["shape"]).copy()
    return out


def _decode_strings(raw: bytes, shape, stored_crc, verify, name):
    n = int(np.prod(shape)) if shape else 1
    pos, lens = 0, []
    for _ in range(n):
        ln, pos = _read_varint(raw, pos)
        lens.append(ln)
    lengths_crc = struct.unpack_from("<I", raw, pos)[0]
    header_end = pos
    pos += 4
    vals = []
    for ln in lens:
        vals.append(raw[pos:pos + ln])
        pos += ln
    if verify:
        # the entry checksum covers the lengths (as fixed-width integers) followed by the string bytes
        for width in ("<I", "<Q"):
            c = 0
            for ln in lens:
                c = crc32c(struct.pack(width, ln), c)
            if _unmask(lengths_crc) == c:
                break
        else:
            raise ValueError(f"{name}: string-length checksum mismatch")
        del header_end
    return np.array(vals, dtype=object).reshape(shape)


def write_checkpoint(prefix: str, tensors: dict) -> str:
    """Write {name: array} as a single-shard V2 checkpoint (``prefix.index`` + ``prefix.data-00000-of-00001``) plus the
    ``checkpoint`` state file next to it; returns the prefix."""
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    header = f_varint(1, 1) + f_varint(2, 0) + f_bytes(3, f_varint(1, 1))         # 1 shard, little endian, version 1
    items = [(b"", header)]
    data = bytearray()
    for name in sorted(tensors):
        arr = tensors[name].detach().cpu().numpy() if hasattr(tensors[name], "detach") else np.asarray(tensors[name])
        if not arr.flags.c_contiguous:
            arr = arr.copy(order="C")         # (np.ascontiguousarray would turn a scalar into shape [1])
        if arr.dtype not in _DT_OF_NP:
            raise ValueError(f"{name}: dtype {arr.dtype} cannot be stored in a checkpoint")
        raw = arr.astype(arr.dtype.newbyteorder("<"), copy=False).tobytes()
        shape = b"".join(f_bytes(2, f_varint(1, int(d))) for d in arr.shape)
        entry = f_varint(1, _DT_OF_NP[arr.dtype]) + f_bytes(2, shape)
        if len(data):
            entry += f_varint(4, len(data))
        entry += f_varint(5, len(raw)) + bytes([(6 << 3) | 5]) + struct.pack("<I", _mask(crc32c(raw)))
        items.append((name.encode(), entry))
        data += raw
    with open(_data_file(prefix, 0, 1), "wb") as f:
        f.write(bytes(data))
    write_table(prefix + ".index", items)
    base = os.path.basename(prefix)
    with open(os.path.join(os.path.dirname(os.path.abspath(prefix)), "checkpoint"), "w", encoding="utf-8") as f:
        f.write(f'model_checkpoint_path: "{base}"\nall_model_checkpoint_paths: "{base}"\n')
    return prefix


# ------------------------------------------------------------------------------------------------ graphs
def read_graph_def(path: str):
    """GraphDef from a frozen ``.pb``, a ``.meta`` MetaGraphDef, a ``saved_model.pb`` or a SavedModel folder."""
    if os.path.isdir(path):
        path = os.path.join(path, "saved_model.pb")
    with open(path, "rb") as f:
        buf = f.read()

    def graph_of_meta(meta: bytes):
        for f_, wt, v in parse(meta):
            if f_ == 2 and wt == 2:                    # MetaGraphDef.graph_def
                return v
        return None

    raw = None
    base = os.path.basename(path)
    if base.endswith(".meta"):
        raw = graph_of_meta(buf)
    elif base == "saved_model.pb":
        for f_, wt, v in parse(buf):
            if f_ == 2 and wt == 2:                    # SavedModel.meta_graphs
                raw = graph_of_meta(v)
                break
    gd = P.GraphDef()
    gd.ParseFromString(raw if raw is not None else buf)
    if not len(gd.node) and raw is None:               # a .meta file under another name
        raw = graph_of_meta(buf)
        if raw is not None:
            gd.ParseFromString(raw)
    return gd


def get_path(output_name: str, graph_def) -> list[str]:
    """Names of every node ``output_name`` depends on, inputs first (``tf_utils.get_path``)."""
    by_name = {n.name: n for n in graph_def.node}
    order, seen = [], set()

    def visit(name):
        name = name.lstrip("^").split(":")[0]
        if name in seen or name not in by_name:
            return
        seen.add(name)
        for i in by_name[name].input:
            visit(i)
        order.append(name)
    visit(output_name)
    return order


_VARIABLE_OPS = ("VariableV2", "Variable", "VarHandleOp")


def freeze(graph_def, variables: dict, output_node_names):
    """GraphDef with every variable the outputs depend on turned into a ``Const`` of its checkpoint value; reads of
    resource variables (``ReadVariableOp``) become ``Identity``. Nodes the outputs do not need (savers, initialisers,
    optimiser slots) are dropped."""
    by_name = {n.name: n for n in graph_def.node}
    keep: list[str] = []
    seen = set()
    for o in output_node_names:
        for n in get_path(o, graph_def):
            if n not in seen:
                seen.add(n)
                keep.append(n)
    out = P.GraphDef()
    out.versions.CopyFrom(graph_def.versions)
    for name in keep:
        src = by_name[name]
        if src.op in _VARIABLE_OPS:
            if name not in variables:
                raise KeyError(f"variable {name} is not in the checkpoint")
            arr = np.asarray(variables[name])
            n = out.node.add()
            n.name, n.op = name, "Const"
            n.attr["dtype"].type = _DT_OF_NP[arr.dtype]
            n.attr["value"].tensor.CopyFrom(P.numpy_to_tensor(arr))
        elif src.op == "ReadVariableOp":
            n = out.node.add()
            n.name, n.op = name, "Identity"
            n.input.extend(i for i in src.input if not i.startswith("^"))
            if "dtype" in src.attr:
                n.attr["T"].type = src.attr["dtype"].type
        else:
            n = out.node.add()
            n.CopyFrom(src)
            del n.input[:]
            # control edges to dropped nodes (initialisers) must not dangle
            n.input.extend(i for i in src.input if i.lstrip("^").split(":")[0] in seen)
    return out


def dump_model(path: str, graph=None, ckpt_file: str | None = None, bin_name: str = "model.bin"):
    """Write ``path/model.pb`` (the GraphDef) and ``path/model.bin`` (the variables, loadable as ``binFile`` by
    ``Module.loadTF``) from a ``.meta`` / SavedModel / GraphDef and a checkpoint. Returns (pb path, bin path)."""
    os.makedirs(path, exist_ok=True)
    gd = read_graph_def(graph) if isinstance(graph, str) else graph
    variables = read_checkpoint(ckpt_file) if ckpt_file else {}
    pb, binf = os.path.join(path, "model.pb"), os.path.join(path, bin_name)
    if gd is not None:
        with open(pb, "wb") as f:
            f.write(gd.SerializeToString())
    with open(binf, "wb") as f:
        np.savez(f, **{k: np.asarray(v, dtype=np.float32) if np.asarray(v).dtype.kind == "f" else np.asarray(v)
                       for k, v in variables.items() if np.asarray(v).dtype != object})
    return pb, binf
