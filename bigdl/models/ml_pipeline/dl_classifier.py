from bigdl_b200.dlframes.dl_classifier import DLClassifier, DLClassifierModel, DLEstimator, DLModel  # noqa: F401
