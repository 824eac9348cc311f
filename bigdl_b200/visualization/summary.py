"""TensorBoard summaries (reference: S/visualization/{Summary,TrainSummary,ValidationSummary}.scala and
S/visualization/tensorboard/{FileWriter,EventWriter,RecordWriter,FileReader}.scala).

``TrainSummary(log_dir, app_name)`` writes ``<log_dir>/<app_name>/train/bigdl.tfevents.<ts>.<host>``; the optimizer
logs Loss / Throughput / LearningRate every iteration and Parameters (histograms) on a user-set trigger
(TrainSummary.scala:33-88). Events are queued and written by a background thread (EventWriter.scala:34-78); records are
TFRecord-framed with masked CRC32C computed by the C++ host library (csrc/host/checksum.cpp).
"""
from __future__ import annotations

import math
import os
import queue
import socket
import threading
import time

import numpy as np

from ..utils import native
from . import proto


def _bucket_limits():
    """TensorFlow's default histogram buckets: ±1e-12 * 1.1^k up to 1e20 (Summary.scala:127-146)."""
    pos, v = [], 1e-12
    while v < 1e20:
        pos.append(v)
        v *= 1.1
    return [-x for x in reversed(pos)] + [0.0] + pos + [float("inf")]  # last bucket = DBL_MAX in the reference


_LIMITS = np.array(_bucket_limits()[:-1] + [np.finfo(np.float64).max])


def histogram_proto(values) -> bytes:
    v = np.asarray(values, dtype=np.float64).reshape(-1)
    if v.size == 0:
        return proto.histogram(0, 0, 0, 0, 0, [], [])
    idx = np.searchsorted(_LIMITS, v, side="left")
    idx = np.minimum(idx, len(_LIMITS) - 1)
    counts = np.bincount(idx, minlength=len(_LIMITS)).astype(np.float64)
    # keep non-empty buckets plus one empty neighbour on each side, as TensorFlow's encoder does
    limits, bucket = [], []
    nz = np.nonzero(counts)[0]
    if nz.size:
        lo, hi = max(int(nz[0]) - 1, 0), min(int(nz[-1]) + 1, len(_LIMITS) - 1)
        limits, bucket = _LIMITS[lo:hi + 1].tolist(), counts[lo:hi + 1].tolist()
    return proto.histogram(v.min(), v.max(), v.size, v.sum(), float((v * v).sum()), limits, bucket)


class RecordWriter:
    def __init__(self, path: str):
        self._f = open(path, "ab")

    def write(self, payload: bytes):
        self._f.write(native.tfrecord_frame(payload))

    def flush(self):
        self._f.flush()

    def close(self):
        self._f.close()


class EventWriter(threading.Thread):
    """Background writer draining a queue of serialised events (EventWriter.scala:34-78)."""

    def __init__(self, log_dir: str, flush_millis: int = 1000):
        super().__init__(daemon=True)
        os.makedirs(log_dir, exist_ok=True)
        self.path = os.path.join(log_dir, f"bigdl.tfevents.{int(time.time())}.{socket.gethostname()}")
        self._w = RecordWriter(self.path)
        self._q: queue.Queue = queue.Queue()
        self._flush_s = flush_millis / 1000.0
        self._halt = threading.Event()
        self._w.write(proto.event(time.time(), file_version="brain.Event:2"))
        self._w.flush()
        self.start()

    def add(self, ev: bytes):
        self._q.put(ev)

    def run(self):
        while not self._halt.is_set() or not self._q.empty():
            try:
                ev = self._q.get(timeout=self._flush_s)
                self._w.write(ev)
                self._q.task_done()
            except queue.Empty:
                pass
            if self._q.empty():
                self._w.flush()

    def flush(self):
        self._q.join()
        self._w.flush()

    def close(self):
        self._q.join()
        self._halt.set()
        self.join(timeout=5)
        self._w.flush()
        self._w.close()


class FileWriter:
    def __init__(self, log_dir: str, flush_millis: int = 1000):
        self.log_dir = log_dir
        self._ew = EventWriter(log_dir, flush_millis)

    def addSummary(self, summary: bytes, global_step: int):
        self._ew.add(proto.event(time.time(), step=int(global_step), summary=summary))
        return self

    def addGraphDef(self, graph_def: bytes):
        self._ew.add(proto.event(time.time(), graph_def=graph_def))
        return self

    def flush(self):
        self._ew.flush()
        return self

    def close(self):
        self._ew.close()


class FileReader:
    """Read scalars back from event files (tensorboard/FileReader.scala:31-106)."""

    @staticmethod
    def list_files(folder: str):
        out = []
        for root, _, files in os.walk(folder):
            out += [os.path.join(root, f) for f in files if "tfevents" in f]
        return sorted(out)

    @staticmethod
    def readScalar(folder_or_file: str, tag: str):
        files = [folder_or_file] if os.path.isfile(folder_or_file) else FileReader.list_files(folder_or_file)
        rows = []
        for fn in files:
            with open(fn, "rb") as f:
                data = f.read()
            for rec in native.tfrecord_iter(data):
                ev = proto.decode_event(rec)
                for v in ev["values"]:
                    if v.get("tag") == tag and "simple_value" in v:
                        rows.append((int(ev.get("step", 0)), float(v["simple_value"]), float(ev.get("wall_time", 0))))
        rows.sort(key=lambda r: r[0])
        return rows

    read_scalar = readScalar


class Summary:
    """Base class (visualization/Summary.scala:32-110)."""

    def __init__(self, log_dir: str, app_name: str, sub: str):
        self.log_dir, self.app_name = log_dir, app_name
        self.folder = os.path.join(log_dir, app_name, sub)
        self.writer = FileWriter(self.folder)
        self._triggers: dict = {}

    # builders of the serialized ``Summary`` messages (the companion object's Summary.scalar / Summary.histogram,
    # Summary.scala:95,108) for callers that drive a FileWriter themselves
    @staticmethod
    def scalar(tag: str, value: float) -> bytes:
        return proto.summary_scalar(tag, float(value))

    @staticmethod
    def histogram(tag: str, values) -> bytes:
        import torch

        if isinstance(values, torch.Tensor):
            values = values.detach().float().cpu().numpy()
        return proto.summary_histogram(tag, histogram_proto(values))

    def addScalar(self, tag: str, value: float, step: int):
        v = float(value)
        self.writer.addSummary(proto.summary_scalar(tag, v), step)
        return self

    add_scalar = addScalar

    def addHistogram(self, tag: str, values, step: int):
        import torch

        if isinstance(values, torch.Tensor):
            values = values.detach().float().cpu().numpy()
        self.writer.addSummary(proto.summary_histogram(tag, histogram_proto(values)), step)
        return self

    add_histogram = addHistogram

    def readScalar(self, tag: str):
        self.writer.flush()
        return FileReader.readScalar(self.folder, tag)

    read_scalar = readScalar

    def close(self):
        self.writer.close()

    # trigger plumbing shared with the optimizer
    def has_trigger(self, tag: str) -> bool:
        return tag in self._triggers

    def should_log(self, tag: str, state) -> bool:
        t = self._triggers.get(tag)
        return bool(t(state)) if t is not None else False


class TrainSummary(Summary):
    """Loss / Throughput / LearningRate every iteration by default; "Parameters" only when a trigger is set
    (TrainSummary.scala:33-88)."""

    _ALLOWED = ("Loss", "Throughput", "LearningRate", "Parameters", "gradientNorm2")

    def __init__(self, log_dir: str, app_name: str):
        super().__init__(log_dir, app_name, "train")
        from ..optim.trigger import Trigger

        every = Trigger.severalIteration(1)
        self._triggers = {"Loss": every, "Throughput": every, "LearningRate": every}

    def setSummaryTrigger(self, tag: str, trigger):
        if tag not in self._ALLOWED:
            raise ValueError(f"TrainSummary: only support {', '.join(self._ALLOWED[:4])}")
        self._triggers[tag] = trigger
        return self

    set_summary_trigger = setSummaryTrigger

    def getSummaryTrigger(self, tag: str):
        return self._triggers.get(tag)


class ValidationSummary(Summary):
    def __init__(self, log_dir: str, app_name: str):
        super().__init__(log_dir, app_name, "validation")

    def should_log(self, tag, state):
        return True


def write_graph_topology(graph, log_path: str):
    """Dump the model DAG as a TensorFlow GraphDef event so TensorBoard's graph tab shows it
    (S/nn/Graph.scala saveGraphTopology:200-232: one NodeDef per module with its inputs)."""
    nodes = b""
    for n in graph.exec_order:
        m = n.element
        nd = proto.f_string(1, m.getName()) + proto.f_string(2, type(m).__name__)
        for p in n.prev:
            nd += proto.f_string(3, p.element.getName())
        nodes += proto.f_bytes(1, nd)
    w = FileWriter(log_path)
    w.addGraphDef(nodes)
    w.close()
    return graph
