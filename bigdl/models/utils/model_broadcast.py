"""pyspark/bigdl/models/utils/model_broadcast.py: ``broadcast_model(sc, layer)`` -> an object with ``.value``."""
from bigdl_b200.parallel.model_broadcast import ModelBroadcast  # noqa: F401


def broadcast_model(sc, layer):
    """``sc`` (a SparkContext in the reference) is ignored: ranks are torch.distributed processes."""
    return ModelBroadcast().broadcast(layer)
