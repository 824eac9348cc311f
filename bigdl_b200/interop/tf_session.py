"""Run TensorFlow graphs for training / prediction (reference: S/utils/tf/Session.scala:40-739 — ``BigDLSessionImpl``
builds a BigDL model from the sub-graph that produces the requested endpoints and trains it with the ordinary
Optimizer; queue/reader input pipelines of the TF graph are replaced by a DataSet).

Here the input pipeline is always an explicit DataSet (the TF queue runners have no counterpart without a TF runtime);
placeholders of the graph are fed from the samples' features in order."""
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

    def train(self, outputs, dataSet, optMethod, criterion, endWhen, batch_size: int = 32, inputs=None):
        """Train the sub-graph producing ``outputs`` (Session.train :74-110). Returns the trained model."""
        from ..optim.optimizer import Optimizer

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
