"""Run a compiled Keras-1.2.2 model on this framework (P/keras/backend.py: ``with_bigdl_backend(kmodel)`` →
``KerasModelWrapper`` with ``fit / evaluate / predict``).

The Keras model is duck-typed — ``to_json()``, ``layers[i].name`` / ``.get_weights()``, and after ``compile`` the
attributes ``loss``, ``optimizer``, ``metrics`` — so the wrapper works with a real Keras 1.2.2 object and needs no Keras
import itself. The definition becomes a ``bigdl_b200.nn.keras`` model, the weights are converted once
(``WeightLoader``), and training runs through the ordinary ``Optimizer`` (local or distributed by world size; the
reference distinguishes ``is_distributed`` because its local mode runs without Spark)."""
from __future__ import annotations

import numpy as np
import torch

from .keras_converter import DefinitionLoader, WeightLoader
from .keras_helpers import OptimConverter


def _unsupported(what):
    raise Exception(f"Unsupported: {what}")


class KerasModelWrapper:
    def __init__(self, kmodel):
        self.bmodel = DefinitionLoader.from_kmodel(kmodel)
        WeightLoader.load_weights_from_kmodel(self.bmodel, kmodel)
        loss = getattr(kmodel, "loss", None)
        opt = getattr(kmodel, "optimizer", None)
        metrics = getattr(kmodel, "metrics", None)
        self.criterion = OptimConverter.to_bigdl_criterion(loss) if loss else None
        self.optim_method = OptimConverter.to_bigdl_optim_method(opt) if opt else None
        self.metrics = OptimConverter.to_bigdl_metrics(metrics) if metrics else None

    # ------------------------------------------------------------------ helpers
    @staticmethod
    def _samples(x, y=None):
        from ..dataset.base import Sample

        xs = x if isinstance(x, (list, tuple)) else [x]
        n = len(xs[0])
        out = []
        for i in range(n):
            feats = [torch.as_tensor(np.asarray(a[i]), dtype=torch.float32) for a in xs]
            label = None if y is None else torch.as_tensor(np.asarray(y[i]), dtype=torch.float32).reshape(-1)
            out.append(Sample(feats if len(feats) > 1 else feats[0], label))
        return out

    def _module(self):
        m = self.bmodel
        return m.to_model() if hasattr(m, "to_model") else getattr(m, "model", m)

    # ------------------------------------------------------------------ Keras API
    def evaluate(self, x, y=None, batch_size=32, sample_weight=None, is_distributed=False):
        if sample_weight:
            _unsupported("sample_weight")
        if not self.metrics:
            raise Exception("No Metrics found.")
        from ..dataset.base import DataSet
        from ..optim.evaluator import Evaluator

        ds = x if y is None and not isinstance(x, (np.ndarray, list, tuple)) else DataSet.array(self._samples(x, y))
        results = Evaluator(self._module()).test(ds, self.metrics, batch_size)
        return [r[0].result()[0] if hasattr(r[0], "result") else r[0] for r in results]

    def predict(self, x, batch_size=None, verbose=None, is_distributed=False):
        if verbose:
            _unsupported("verbose")
        from ..dataset.base import DataSet
        from ..optim.predictor import Predictor

        ds = DataSet.array(self._samples(x)) if isinstance(x, (np.ndarray, list, tuple)) else x
        outs = Predictor(self._module(), batch_size=batch_size or 32).predict(ds)
        return np.stack([np.asarray(o.detach().cpu() if hasattr(o, "detach") else o) for o in outs])

    def fit(self, x, y=None, batch_size=32, nb_epoch=10, verbose=1, callbacks=None, validation_split=0.,
            validation_data=None, shuffle=True, class_weight=None, sample_weight=None, initial_epoch=0,
            is_distributed=False):
        for name, v in (("callbacks", callbacks), ("class_weight", class_weight), ("sample_weight", sample_weight)):
            if v:
                _unsupported(name)
        if validation_split != 0.:
            _unsupported("validation_split")
        if initial_epoch != 0:
            _unsupported("initial_epoch")
        if self.criterion is None or self.optim_method is None:
            raise Exception("compile the Keras model (loss and optimizer) before fit")
        from ..dataset.base import DataSet
        from ..optim.optimizer import Optimizer
        from ..optim.trigger import Trigger

        ds = DataSet.array(self._samples(x, y)) if isinstance(x, (np.ndarray, list, tuple)) else x
        opt = Optimizer(model=self._module(), training_set=ds, criterion=self.criterion, optim_method=self.optim_method,
                        end_trigger=Trigger.maxEpoch(nb_epoch), batch_size=batch_size)
        if validation_data is not None and self.metrics:
            vx, vy = validation_data
            opt.setValidation(Trigger.everyEpoch(), DataSet.array(self._samples(vx, vy)), self.metrics, batch_size)
        opt.optimize()
        return self


def with_bigdl_backend(kmodel):
    """Replace a Keras model's execution engine: returns a wrapper with the same fit / evaluate / predict calls."""
    return KerasModelWrapper(kmodel)
