from bigdl_b200.dlframes.dl_classifier import DLClassifier, DLClassifierModel, DLEstimator, DLModel  # noqa: F401


class _Param:
    _name, _default = "", None

    def _get(self):
        return getattr(self, self._name, self._default)


class HasBatchSize:
    def setBatchSize(self, val): self.batchSize = int(val); return self
    def getBatchSize(self): return getattr(self, "batchSize", 1)


class HasMaxEpoch:
    def setMaxEpoch(self, val): self.maxEpoch = int(val); return self
    def getMaxEpoch(self): return getattr(self, "maxEpoch", 100)


class HasFeatureSize:
    def setFeatureSize(self, val): self.featureSize = list(val); return self
    def getFeatureSize(self): return getattr(self, "featureSize", None)


class HasLearningRate:
    def setLearningRate(self, val): self.learningRate = float(val); return self
    def getLearningRate(self): return getattr(self, "learningRate", 1e-3)
