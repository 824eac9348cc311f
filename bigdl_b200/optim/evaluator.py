"""Evaluator / Validator: run ValidationMethods over a dataset (S/optim/Evaluator.scala:51-85, Validator.scala,
DistriValidator.scala, LocalValidator.scala, AbstractOptimizer.validate :93-191)."""
from __future__ import annotations

import torch

from ..dataset.base import AbstractDataSet, DataSet, MiniBatch, Sample, SampleToMiniBatch
from .validation import all_reduce_result


def _batches(dataset, batch_size):
    if isinstance(dataset, AbstractDataSet):
        it = dataset.data(train=False)
    else:
        it = iter(dataset)
    buf = []
    for item in it:
        if isinstance(item, MiniBatch):
            yield item
        else:
            buf.append(item)
            if len(buf) == (batch_size or 32):
                yield MiniBatch.from_samples(buf)
                buf = []
    if buf:
        yield MiniBatch.from_samples(buf)


def run_validation(model, dataset, methods, batch_size: int | None = None, distributed: bool = False):
    was_training = model.isTraining()
    model.evaluate()
    device = model.device()
    model._refresh_compute_weights()
    totals = [None] * len(methods)
    with torch.no_grad():
        for batch in _batches(dataset, batch_size):
            if device.type == "cuda":
                batch = batch.to(device)
            out = model._call(batch.getInput())
            tgt = batch.getTarget()
            for i, m in enumerate(methods):
                r = m(out, tgt)
                totals[i] = r if totals[i] is None else totals[i] + r
    if distributed:
        totals = [all_reduce_result(t) if t is not None else t for t in totals]
    if was_training:
        model.training()
    return totals


class Evaluator:
    def __init__(self, model):
        self.model = model

    def test(self, dataset, vMethods=None, batchSize: int | None = None, batch_size: int | None = None,
             val_methods=None):
        methods = vMethods if vMethods is not None else val_methods
        if isinstance(methods, int):            # Python API order: evaluate(val_rdd, batch_size, val_methods)
            methods, batchSize = batchSize, methods
        if isinstance(dataset, (list, tuple)) and dataset and isinstance(dataset[0], Sample):
            dataset = DataSet.array(list(dataset))
        res = run_validation(self.model, dataset, methods, batchSize or batch_size,
                             distributed=isinstance(dataset, AbstractDataSet) and dataset.is_distributed())
        return list(zip(res, methods))


class Validator(Evaluator):
    """Reference alias kept for API parity (S/optim/Validator.scala)."""

    def __init__(self, model, dataset=None):
        super().__init__(model)
        self.dataset = dataset

    def test(self, vMethods, dataset=None, **kw):
        return super().test(dataset if dataset is not None else self.dataset, vMethods, **kw)


LocalValidator = Validator
DistriValidator = Validator
