"""Keras-style ``Sequential`` / ``Model`` with ``compile / fit / evaluate / predict`` (S/nn/keras/Topology.scala:38-330,
P/nn/keras/topology.py:25-250). Training goes through the ordinary Optimizer (Local or Distri, picked from the
process group), so everything the optimizers provide — fused GPU steps, CUDA graphs, checkpoints, summaries — applies."""
from __future__ import annotations

import numpy as np
import torch

from .. import criterion as C
from ..module import Graph, Node
from ...dataset.base import DataSet, Sample
from ...optim import optim_method as OM
from ...optim import validation as VM
from ...optim.trigger import Trigger
from .layers import KerasLayer


def _to_optim(o):
    if not isinstance(o, str):
        return o
    o = o.lower()
    table = {"adagrad": lambda: OM.Adagrad(learningrate=0.01), "sgd": lambda: OM.SGD(learningrate=0.01),
             "adam": OM.Adam, "rmsprop": lambda: OM.RMSprop(learningrate=0.001, decayrate=0.9),
             "adadelta": lambda: OM.Adadelta(decayrate=0.95, epsilon=1e-8), "adamax": lambda: OM.Adamax(epsilon=1e-8)}
    if o not in table:
        raise TypeError(f"Unsupported optimizer: {o}")
    return table[o]()


def _to_loss(l):
    if not isinstance(l, str):
        return l
    l = l.lower()
    table = {"categorical_crossentropy": C.CategoricalCrossEntropy, "mse": C.MSECriterion,
             "mean_squared_error": C.MSECriterion, "binary_crossentropy": C.BCECriterion, "mae": C.AbsCriterion,
             "mean_absolute_error": C.AbsCriterion, "hinge": C.MarginCriterion,
             "mean_absolute_percentage_error": C.MeanAbsolutePercentageCriterion,
             "mape": C.MeanAbsolutePercentageCriterion,
             "mean_squared_logarithmic_error": C.MeanSquaredLogarithmicCriterion,
             "msle": C.MeanSquaredLogarithmicCriterion, "squared_hinge": lambda: C.MarginCriterion(squared=True),
             "sparse_categorical_crossentropy": lambda: C.ClassNLLCriterion(logProbAsInput=False),
             "kullback_leibler_divergence": C.KullbackLeiblerDivergenceCriterion,
             "kld": C.KullbackLeiblerDivergenceCriterion, "poisson": C.PoissonCriterion,
             "cosine_proximity": C.CosineProximityCriterion, "cosine": C.CosineProximityCriterion}
    if l not in table:
        raise TypeError(f"Unsupported loss: {l}")
    return table[l]()


def _to_metrics(ms):
    out = []
    for m in ms or []:
        if isinstance(m, str):
            if m.lower() != "accuracy":
                raise TypeError(f"Unsupported metrics: {m}")
            out.append(VM.Top1Accuracy())
        else:
            out.append(m)
    return out


class KerasModel:
    """Mixin holding the training configuration."""

    def compile(self, optimizer, loss, metrics=None):
        self._optim, self._loss, self._metrics = _to_optim(optimizer), _to_loss(loss), _to_metrics(metrics)
        return self

    def _dataset(self, x, y=None):
        if isinstance(x, np.ndarray) or isinstance(x, torch.Tensor):
            x = np.asarray(x)
            samples = [Sample(x[i], None if y is None else np.asarray(y)[i]) for i in range(len(x))]
            return DataSet.array(samples)
        if isinstance(x, (list, tuple)):
            return DataSet.array(list(x))
        return x

    def fit(self, x, y=None, batch_size=32, nb_epoch=10, validation_data=None, distributed=True):
        if getattr(self, "_optim", None) is None:
            raise RuntimeError("compile must be called before fit")
        from ...optim.optimizer import Optimizer

        ds = self._dataset(x, y)
        opt = Optimizer(model=self, training_set=ds, criterion=self._loss, optim_method=self._optim,
                        end_trigger=Trigger.maxEpoch(nb_epoch), batch_size=batch_size)
        if validation_data is not None and self._metrics:
            vx, vy = validation_data if isinstance(validation_data, tuple) else (validation_data, None)
            opt.setValidation(Trigger.everyEpoch(), self._dataset(vx, vy), self._metrics, batch_size)
        for attr in ("_train_summary", "_val_summary"):
            s = getattr(self, attr, None)
            if s is not None:
                (opt.setTrainSummary if attr == "_train_summary" else opt.setValidationSummary)(s)
        ck = getattr(self, "_checkpoint", None)
        if ck is not None:
            opt.setCheckpoint(ck[0], Trigger.everyEpoch(), ck[1])
        opt.optimize()
        self._last_optimizer = opt
        return self

    def evaluate(self, x=None, y=None, batch_size=32):
        if x is None:
            return super().evaluate()              # no data: the plain "switch to inference mode" of every Module
        if getattr(self, "_loss", None) is None:
            raise RuntimeError("compile must be called before evaluate")
        from ...optim.evaluator import Evaluator

        methods = [VM.Loss(self._loss)] + list(self._metrics or [])
        res = Evaluator(self).test(self._dataset(x, y), methods, batch_size)
        return [r[0].result()[0] for r in res]

    def predict(self, x, batch_size=32, distributed=True):
        from ...optim.predictor import Predictor

        return Predictor(self, batch_size=batch_size).predict(self._dataset(x))

    def predict_classes(self, x, batch_size=32, zero_based_label=True):
        from ...optim.predictor import Predictor

        cls = Predictor(self, batch_size=batch_size).predictClass(self._dataset(x))
        return [int(c) - 1 for c in cls] if zero_based_label else [int(c) for c in cls]

    def set_tensorboard(self, log_dir, app_name):
        from ...visualization import TrainSummary, ValidationSummary

        self._train_summary, self._val_summary = TrainSummary(log_dir, app_name), ValidationSummary(log_dir, app_name)
        return self

    def set_checkpoint(self, path, over_write=True):
        self._checkpoint = (path, over_write)
        return self

    def summary(self):
        lines = []
        for m in getattr(self, "modules", []):
            try:
                shp = m.getOutputShape() if isinstance(m, KerasLayer) else "-"
            except RuntimeError:
                shp = "?"
            n = sum(p.numel() for _, _, p in m.leaf_parameters())
            lines.append(f"{m.getName():32s} {type(m).__name__:24s} {str(shp):24s} {n}")
        text = "\n".join(lines)
        print(text)
        return text


class _FromValue:
    @staticmethod
    def from_jvalue(jvalue, bigdl_type="float"):
        """Python API hook that wrapped a JVM object; the Python object is the model here."""
        return jvalue


class Sequential(_FromValue, KerasModel, KerasLayer):
    """Keras Sequential: the first layer must carry ``input_shape``; every ``add`` builds the new layer against the
    running output shape (Topology.scala:269-326)."""

    def __init__(self, name=None):
        KerasLayer.__init__(self, None, name)
        self.modules = []
        self.labor = None
        self._cur_shape = None

    def add(self, layer):
        if not isinstance(layer, KerasLayer):
            raise TypeError(f"Sequential(Keras) only accepts Keras-style layers, got {type(layer).__name__}")
        if not self.modules:
            if layer._input_shape is None and not isinstance(layer, Sequential):
                raise ValueError("The first layer should explicitly declare inputShape")
            self._input_shape = layer._input_shape
            self._cur_shape = layer.build(layer._input_shape)
        else:
            self._cur_shape = layer.build(self._cur_shape)
        self.modules.append(layer)
        self._output_shape = self._cur_shape
        return self

    def isBuilt(self):
        return bool(self.modules)

    def build(self, input_shape):
        return self._output_shape

    def updateOutput(self, input):
        x = input
        for m in self.modules:
            x = m._call(x)
        return x


class Model(_FromValue, KerasModel, Graph):
    """Functional model over ``Input(shape=…)`` nodes (Topology.scala:165-265)."""

    def __init__(self, input, output, name=None):
        ins = list(input) if isinstance(input, (list, tuple)) else [input]
        outs = list(output) if isinstance(output, (list, tuple)) else [output]
        Graph.__init__(self, ins, outs)
        if name:
            self.setName(name)
        self._input_shape = [n._keras_shape for n in ins]
        self._output_shape = [getattr(n, "_keras_shape", None) for n in outs]

    def isKerasStyle(self):
        return True

    def getOutputShape(self):
        s = [tuple([None] + list(x)) for x in self._output_shape]
        return s[0] if len(s) == 1 else s

    def getInputShape(self):
        s = [tuple([None] + list(x)) for x in self._input_shape]
        return s[0] if len(s) == 1 else s

    get_output_shape, get_input_shape = getOutputShape, getInputShape
