"""Run TensorFlow graphs for training / prediction (reference: S/utils/tf/Session.scala:40-739 — ``BigDLSessionImpl``
builds a BigDL model from the sub-graph that produces the requested endpoints and trains it with the ordinary
Optimizer; queue/reader input pipelines of the TF graph are replaced by a DataSet).

Two ways to feed a graph: an explicit DataSet (placeholders are fed from the samples' features in order), or — as in the
reference, S/utils/tf/Session.scala:112-420 ``handleLocalDequeue`` / ``handleReaderNode`` — the graph's own queue input
pipeline: the ``QueueDequeue*`` node the endpoints hang off is traced back through its queue to the ``QueueEnqueue*``
ops, whose component sub-graphs are evaluated here (constants, or records pulled by ``ReaderReadV2`` from the TFRecord /
fixed-length files a filename queue names, decoded by ``ParseExample`` / ``DecodeRaw`` ...) into a DataSet; the last
dequeued component is the label, the others are the model inputs."""
from __future__ import annotations

import numpy as np
import torch

from . import tf_pb as P
from .tensorflow import load_tf


class Session:
    def __init__(self, graph, bin_file: str | None = None):
        """``graph``: path to a frozen GraphDef, its bytes, or a parsed GraphDef message."""
        if isinstance(graph, str):
            with open(graph, "rb") as f:
                graph = f.read()
        if isinstance(graph, (bytes, bytearray)):
            gd = P.GraphDef()
            gd.ParseFromString(bytes(graph))
        else:
            gd = graph
        self.graph_def, self.bin_file = gd, bin_file
        self._models: dict = {}

    def placeholders(self) -> list[str]:
        return [n.name for n in self.graph_def.node if n.op in ("Placeholder", "PlaceholderV2")]

    def _model(self, endpoints, inputs=None):
        ins = list(inputs) if inputs else self._needed_placeholders(endpoints)
        key = (tuple(endpoints), tuple(ins))
        if key not in self._models:
            model = load_tf(self.graph_def.SerializeToString(), ins, list(endpoints), binFile=self.bin_file)
            # variables are shared by name across every sub-graph built from this session (Session.scala Context):
            # a model built after training starts from the trained values
            store = self.__dict__.setdefault("_variables", {})
            for m, _, p in model.leaf_parameters():
                prev = store.get(m.getName())
                if prev is not None and prev.shape == p.shape:
                    with torch.no_grad():
                        p.copy_(prev)
                    m._mark_weights_dirty()
                store[m.getName()] = p
            self._models[key] = model
        else:
            store = self.__dict__.setdefault("_variables", {})
            for m, _, p in self._models[key].leaf_parameters():
                latest = store.get(m.getName())
                if latest is not None and latest is not p and latest.shape == p.shape:
                    with torch.no_grad():
                        p.copy_(latest)
                    m._mark_weights_dirty()
        return self._models[key]

    def _needed_placeholders(self, endpoints):
        by_name = {n.name: n for n in self.graph_def.node}
        seen, stack, found = set(), [e.split(":")[0] for e in endpoints], []
        while stack:
            name = stack.pop()
            if name in seen:
                continue
            seen.add(name)
            nd = by_name[name]
            if nd.op in ("Placeholder", "PlaceholderV2"):
                found.append(name)
            stack.extend(i.lstrip("^").split(":")[0] for i in nd.input)
        order = {n: i for i, n in enumerate(self.placeholders())}
        return sorted(found, key=lambda n: order[n])

    # ------------------------------------------------------------------ the graph's own input pipeline
    _DEQUEUE = ("QueueDequeueV2", "QueueDequeueManyV2", "QueueDequeueUpToV2")
    _ENQUEUE = ("QueueEnqueueV2", "QueueEnqueueManyV2")
    _READERS = {"TFRecordReaderV2": "tfrecord", "FixedLengthRecordReaderV2": "fixed", "TextLineReaderV2": "text"}

    def _find_dequeue(self, endpoints):
        by_name = {n.name: n for n in self.graph_def.node}
        seen, stack = set(), [e.split(":")[0] for e in endpoints]
        while stack:
            name = stack.pop()
            if name in seen:
                continue
            seen.add(name)
            nd = by_name[name]
            if nd.op in self._DEQUEUE:
                return nd
            stack.extend(i.lstrip("^").split(":")[0] for i in nd.input)
        return None

    def _enqueues_of(self, queue_name):
        return [n for n in self.graph_def.node if n.op in self._ENQUEUE and n.input[0].split(":")[0] == queue_name]

    def _eval(self, names, feed=None):
        """Evaluate graph tensors (``node`` or ``node:k``) that depend only on constants and ``feed`` {tensor: value}."""
        from ..utils.table import Table

        feed = feed or {}
        per_node: dict = {}
        for key, val in feed.items():                       # "node:k" feeds output k of a multi-output node
            node, _, k = key.partition(":")
            per_node.setdefault(node, {})[int(k) if k else 0] = val
        ins = list(per_node)
        m = load_tf(self.graph_def.SerializeToString(), ins, list(names), binFile=self.bin_file).evaluate()
        vals = []
        for node in ins:
            outs = per_node[node]
            if list(outs) == [0]:
                vals.append(outs[0])
            else:
                vals.append(Table(*[outs.get(i, torch.zeros(0)) for i in range(max(outs) + 1)]))
        if not ins:
            out = m.updateOutput(Table())
        elif len(vals) == 1:
            out = m.updateOutput(vals[0])
        else:
            out = m.updateOutput(Table(*vals))
        if isinstance(out, torch.Tensor) or not hasattr(out, "to_list"):
            return [out]
        return out.to_list()

    def _records(self, read_node):
        """Records behind a ``ReaderReadV2`` node: every file of its filename queue, through the reader's format."""
        from .tfrecord import FixedLengthRecordReader, TFRecordIterator

        by_name = {n.name: n for n in self.graph_def.node}
        reader = by_name[read_node.input[0].split(":")[0]]
        fq = read_node.input[1].split(":")[0]
        files = []
        for enq in self._enqueues_of(fq):
            for comp in enq.input[1:]:
                if comp.startswith("^"):
                    continue
                for v in self._eval([comp]):
                    arr = np.asarray(v.cpu() if isinstance(v, torch.Tensor) else v, dtype=object).reshape(-1)
                    files.extend(x.decode() if isinstance(x, bytes) else str(x) for x in arr)
        kind = self._READERS.get(reader.op)
        if kind is None:
            raise ValueError(f"unsupported reader {reader.op}")
        for f in files:
            if kind == "tfrecord":
                yield from TFRecordIterator(f)
            elif kind == "fixed":
                a = reader.attr
                yield from FixedLengthRecordReader(f, int(a["record_bytes"].i), int(a["header_bytes"].i),
                                                   int(a["footer_bytes"].i))
            else:
                with open(f, "rb") as fh:
                    for line in fh:
                        yield line.rstrip(b"\n")

    def queue_dataset(self, endpoints):
        """(model input tensor names, list of component tuples) for the queue the ``endpoints`` read from."""
        deq = self._find_dequeue(endpoints)
        if deq is None:
            raise ValueError("the endpoints do not depend on a QueueDequeue op; pass a DataSet")
        by_name = {n.name: n for n in self.graph_def.node}
        queue = deq.input[0].split(":")[0]
        ncomp = len(by_name[queue].attr["component_types"].list.type) or 1
        rows = []
        for enq in self._enqueues_of(queue):
            comps = [c for c in enq.input[1:] if not c.startswith("^")]
            many = enq.op == "QueueEnqueueManyV2"
            # does a reader feed this enqueue?
            read = None
            seen, stack = set(), [c.split(":")[0] for c in comps]
            while stack and read is None:
                nm = stack.pop()
                if nm in seen:
                    continue
                seen.add(nm)
                nd = by_name[nm]
                if nd.op == "ReaderReadV2":
                    read = nd
                stack.extend(i.lstrip("^").split(":")[0] for i in nd.input)
            if read is None:
                vals = self._eval(comps)
                if many:
                    n = len(vals[0])
                    rows.extend(tuple(v[i] for v in vals) for i in range(n))
                else:
                    rows.append(tuple(vals))
            else:
                for rec in self._records(read):
                    vals = self._eval(comps, {read.name + ":1": np.array([rec], dtype=object) if many else rec})
                    if many:
                        n = len(vals[0])
                        rows.extend(tuple(v[i] for v in vals) for i in range(n))
                    else:
                        rows.append(tuple(vals))
        names = [deq.name if k == 0 else f"{deq.name}:{k}" for k in range(ncomp)]
        return names, rows

    def train(self, outputs, dataSet=None, optMethod=None, criterion=None, endWhen=None, batch_size: int = 32,
              inputs=None):
        """Train the sub-graph producing ``outputs`` (Session.train :74-110). Returns the trained model. With
        ``dataSet=None`` the data comes from the graph's queue pipeline (see the module docstring)."""
        from ..optim.optimizer import Optimizer

        if dataSet is None:
            from ..dataset.base import DataSet, Sample

            names, rows = self.queue_dataset(outputs)
            to_t = lambda v: v if isinstance(v, torch.Tensor) else torch.as_tensor(np.asarray(v))      # noqa: E731
            samples = [Sample([to_t(v).float() for v in r[:-1]], to_t(r[-1]).float()) for r in rows]
            dataSet = DataSet.array(samples)
            inputs = names[:-1]
        model = self._model(outputs, inputs)
        model.training()
        opt = Optimizer(model=model, training_set=dataSet, criterion=criterion, optim_method=optMethod,
                        end_trigger=endWhen, batch_size=batch_size)
        opt.optimize()
        store = self.__dict__.setdefault("_variables", {})
        for m, _, p in model.leaf_parameters():
            store[m.getName()] = p
        return model

    def predict(self, endpoints, dataSet, batch_size: int = 32, inputs=None):
        from ..optim.predictor import Predictor

        model = self._model(endpoints, inputs).evaluate()
        return Predictor(model, batch_size=batch_size).predict(dataSet)

    def run(self, endpoints, feed: dict):
        """Evaluate endpoints for one feed dict {placeholder: array} (the TF ``session.run`` shape of the API)."""
        names = list(feed)
        model = self._model(endpoints, names).evaluate()
        vals = [torch.as_tensor(np.asarray(feed[n])) for n in names]
        out = model.forward(vals[0] if len(vals) == 1 else vals)
        return out

    def saveParameters(self, path: str, endpoints=None):
        """Dump the trainable variables of the (trained) sub-graph as an .npz usable as ``binFile``."""
        if endpoints is None:
            if not self._models:
                raise ValueError("nothing has been built yet: call train/predict first or pass endpoints")
            model = next(iter(self._models.values()))
        else:
            model = self._model(endpoints)
        arrays = {m.getName(): p.detach().cpu().numpy() for m, _, p in model.leaf_parameters()}
        np.savez(path, **arrays)
        return path
