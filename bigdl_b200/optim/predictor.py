"""Predictor / LocalPredictor / PredictionService (S/optim/Predictor.scala:35-257, LocalPredictor.scala,
PredictionService.scala:56-354)."""
from __future__ import annotations

import queue
import threading

import numpy as np
import torch

from ..dataset.base import AbstractDataSet, MiniBatch, Sample
from ..utils.table import Table


def _iter_batches(data, batch_size):
    if isinstance(data, AbstractDataSet):
        data = data.data(train=False)
    if isinstance(data, (np.ndarray, torch.Tensor)):
        t = torch.as_tensor(data)
        bs = batch_size if batch_size and batch_size > 0 else len(t)
        for i in range(0, len(t), bs):
            yield MiniBatch(t[i:i + bs])
        return
    buf = []
    bs = batch_size if batch_size and batch_size > 0 else 32
    for item in data:
        if isinstance(item, MiniBatch):
            yield item
        else:
            buf.append(item if isinstance(item, Sample) else Sample(item))
            if len(buf) == bs:
                yield MiniBatch.from_samples(buf)
                buf = []
    if buf:
        yield MiniBatch.from_samples(buf)


class Predictor:
    def __init__(self, model, batch_size: int = -1, feature_padding=None, batch_per_partition: int = 4):
        self.model, self.batch_size = model, batch_size

    def _outputs(self, data):
        model = self.model
        was = model.isTraining()
        model.evaluate()
        model._refresh_compute_weights()
        dev = model.device()
        outs = []
        with torch.no_grad():
            for b in _iter_batches(data, self.batch_size):
                if dev.type == "cuda":
                    b = b.to(dev)
                o = model._call(b.getInput())
                if isinstance(o, Table):
                    rows = list(zip(*[list(t.float().cpu()) for t in o.to_list()]))
                    outs.extend(Table(*r) for r in rows)
                else:
                    outs.extend(list(o.float().cpu()))
        if was:
            model.training()
        return outs

    def predict(self, data):
        return self._outputs(data)

    def predictClass(self, data):
        res = []
        for o in self._outputs(data):
            o = o.reshape(-1)
            res.append(int(o.argmax()) + 1 if o.numel() > 1 else int(o.item() >= 0.5))
        return res

    def predictImage(self, image_frame, output_layer=None, predict_key="predict"):
        feats = getattr(image_frame, "features", None)
        feats = feats() if callable(feats) else (list(image_frame) if feats is None else list(feats))

        def sample_of(f):
            if hasattr(f, "get_sample"):
                return f.get_sample()
            try:
                return f["sample"]          # written by ImageFrameToSample
            except (KeyError, TypeError, IndexError):
                return f

        samples = [sample_of(f) for f in feats]
        outs = self._outputs(samples)
        for f, o in zip(feats, outs):
            if hasattr(f, "__setitem__"):
                f[predict_key] = o
        return image_frame


LocalPredictor = Predictor


class PredictionService:
    """Thread-safe serving front end: a pool of ``num_threads`` model replicas that share weights
    (S/optim/PredictionService.scala:56-110). ``predict`` accepts a tensor/table activity or serialized bytes."""

    def __init__(self, model, num_threads: int = 1):
        model.evaluate()
        self._pool: "queue.Queue" = queue.Queue()
        self._pool.put(model)
        for _ in range(max(num_threads, 1) - 1):
            clone = model.cloneModule()
            clone.evaluate()
            # replicas share the weight storage of the original
            for p, q in zip(model.torch_parameters(), clone.torch_parameters()):
                q.data = p.data
            self._pool.put(clone)

    @staticmethod
    def apply(model, num_threads: int = 1):
        return PredictionService(model, num_threads)

    def predict(self, request):
        if isinstance(request, (bytes, bytearray)):
            from ..serialization.activity_io import deserialize_activity, serialize_activity

            try:
                act = deserialize_activity(bytes(request))
            except Exception as e:  # noqa: BLE001 - errors travel back as data, like the reference
                return serialize_activity(torch.tensor([float("nan")]), error=str(e))
            return serialize_activity(self.predict(act))
        m = self._pool.get()
        try:
            m._refresh_compute_weights()
            with torch.no_grad():
                from ..nn.module import _as_activity, _detach

                x = _as_activity(request, m.device() if m.torch_parameters() else None)
                out = m._call(x)
                return _detach(out) if not isinstance(out, torch.Tensor) else out.detach().clone()
        except Exception as e:  # noqa: BLE001
            return torch.tensor([float("nan")])
        finally:
            m.clearState()
            self._pool.put(m)
