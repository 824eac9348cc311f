"""P/nn/layer.py surface: ``Layer`` (base), ``Container``, ``Model`` (graph), ``Sequential`` and every layer class."""
from bigdl_b200.nn.layer import *  # noqa: F401,F403
from bigdl_b200.nn.layer import Graph, Module, Node
from bigdl_b200.nn.module import Container, Sequential  # noqa: F401

Layer = Module
Model = Graph
SharedStaticUtils = Module


def _py_aliases(cls):
    """snake_case methods of the Python API on top of the camelCase reference names."""
    pairs = {"zero_grad_parameters": "zeroGradParameters", "update_parameters": "updateParameters",
             "get_weights": "getWeights", "set_weights": "setWeights", "set_name": "setName", "name": "getName",
             "reset": "reset", "saveModel": "saveModule", "save_tensorflow": "saveTF", "set_seed": None,
             "is_training": "isTraining", "unfreeze": "unFreeze", "set_running_mean": None, "set_running_std": None}
    for py, ref in pairs.items():
        if ref and hasattr(cls, ref) and not hasattr(cls, py):
            setattr(cls, py, getattr(cls, ref))


_py_aliases(Module)
if not hasattr(Module, "loadModel"):
    Module.loadModel = staticmethod(Module.loadModule)

from bigdl_b200.nn.onnx import Gemm  # noqa: F401,E402  (pyspark/bigdl/nn/layer.py also defines Gemm)
