"""ML-pipeline style estimators over DataFrames (S/dlframes/{DLEstimator,DLClassifier,DLImageReader,
DLImageTransformer}.scala, spark/dl/src/main/scala/org/apache/spark/ml/*, P/dlframes/*).

The reference plugs into Spark ML (Estimator.fit(DataFrame) → Model.transform(DataFrame)). Here the DataFrame is a
pandas DataFrame (or any mapping of column name → sequence); the parameter plumbing (setFeaturesCol, setLabelCol,
setPredictionCol, setBatchSize, setMaxEpoch, setLearningRate, …) and the fit/transform contract are the same, and
training runs through the framework's Optimizer."""
from .dl_classifier import DLClassifier, DLClassifierModel, DLEstimator, DLModel
from .dl_image import DLImageReader, DLImageTransformer

__all__ = ["DLEstimator", "DLModel", "DLClassifier", "DLClassifierModel", "DLImageReader", "DLImageTransformer"]
