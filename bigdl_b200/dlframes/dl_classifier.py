from __future__ import annotations

import numpy as np
import torch

from ..dataset.base import DataSet, Sample
from ..optim.optim_method import SGD
from ..optim.optimizer import Optimizer
from ..optim.predictor import Predictor
from ..optim.trigger import Trigger


class _Params:
    def __init__(self):
        self.featuresCol, self.labelCol, self.predictionCol = "features", "label", "prediction"
        self.batchSize, self.maxEpoch, self.learningRate = 1, 50, 1e-3
        self.optimMethod = None

    def setFeaturesCol(self, v): self.featuresCol = v; return self
    def setLabelCol(self, v): self.labelCol = v; return self
    def setPredictionCol(self, v): self.predictionCol = v; return self
    def setBatchSize(self, v): self.batchSize = int(v); return self
    def setMaxEpoch(self, v): self.maxEpoch = int(v); return self
    def setLearningRate(self, v): self.learningRate = float(v); return self
    def setOptimMethod(self, v): self.optimMethod = v; return self
    def getFeaturesCol(self): return self.featuresCol
    def getLabelCol(self): return self.labelCol
    def getPredictionCol(self): return self.predictionCol
    def getBatchSize(self): return self.batchSize
    def getMaxEpoch(self): return self.maxEpoch
    def getLearningRate(self): return self.learningRate


def _column(df, name):
    col = df[name]
    return list(col.values) if hasattr(col, "values") else list(col)


def _to_tensor(v, size):
    a = np.asarray(v, dtype=np.float32).reshape(-1)
    return torch.from_numpy(a.reshape(size) if size else a)


class DLEstimator(_Params):
    """``DLEstimator(model, criterion, feature_size, label_size)``; ``fit(df)`` trains and returns a ``DLModel``
    (DLEstimator.scala:60-300: features/labels may be scalars, arrays or vectors and are reshaped to the given sizes)."""

    def __init__(self, model, criterion, feature_size, label_size):
        super().__init__()
        self.model, self.criterion = model, criterion
        self.featureSize, self.labelSize = list(feature_size), list(label_size)

    def _samples(self, df):
        feats, labels = _column(df, self.featuresCol), _column(df, self.labelCol)
        return [Sample(_to_tensor(f, self.featureSize), _to_tensor(l, self.labelSize)) for f, l in zip(feats, labels)]

    def fit(self, df):
        om = self.optimMethod or SGD(learningrate=self.learningRate)
        opt = Optimizer(model=self.model, training_set=DataSet.array(self._samples(df)), criterion=self.criterion,
                        optim_method=om, end_trigger=Trigger.maxEpoch(self.maxEpoch), batch_size=self.batchSize)
        opt.optimize()
        return self._wrap(self.model)

    def _wrap(self, model):
        m = DLModel(model, self.featureSize)
        m.featuresCol, m.predictionCol, m.batchSize = self.featuresCol, self.predictionCol, self.batchSize
        return m


class DLModel(_Params):
    def __init__(self, model, featureSize):
        super().__init__()
        self.model, self.featureSize = model, list(featureSize)

    @classmethod
    def of(cls, jvalue, feature_preprocessing=None, bigdl_type="float"):
        """Python API constructor from an existing transformer (P/dlframes/dl_classifier.py ``DLModel.of``)."""
        return jvalue if isinstance(jvalue, DLModel) else cls(jvalue, feature_preprocessing or [])

    def setFeatureSize(self, v): self.featureSize = list(v); return self
    def getFeatureSize(self): return self.featureSize

    def _predict(self, df):
        samples = [Sample(_to_tensor(f, self.featureSize)) for f in _column(df, self.featuresCol)]
        return Predictor(self.model, batch_size=max(self.batchSize, 1)).predict(DataSet.array(samples))

    def _post(self, outputs):
        return [np.asarray(o, dtype=np.float32).reshape(-1).tolist() for o in outputs]

    def transform(self, df):
        preds = self._post(self._predict(df))
        try:
            out = df.copy()
            out[self.predictionCol] = preds
            return out
        except AttributeError:
            out = dict(df)
            out[self.predictionCol] = preds
            return out


class DLClassifier(DLEstimator):
    """Classification flavour: scalar 1-based labels, prediction = argmax + 1 (DLClassifier.scala:38-90)."""

    def __init__(self, model, criterion, feature_size):
        super().__init__(model, criterion, feature_size, [1])

    def _wrap(self, model):
        m = DLClassifierModel(model, self.featureSize)
        m.featuresCol, m.predictionCol, m.batchSize = self.featuresCol, self.predictionCol, self.batchSize
        return m


class DLClassifierModel(DLModel):
    def _post(self, outputs):
        return [float(np.asarray(o).reshape(-1).argmax() + 1) for o in outputs]
