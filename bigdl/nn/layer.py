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


# ---- the rest of pyspark/bigdl/nn/layer.py ``Layer`` / ``Model`` method surface
import numpy as _np  # noqa: E402
import torch as _torch  # noqa: E402


def _set_running(name):
    def setter(self, arr):
        """Only for BatchNormalization layers (Layer.set_running_mean / set_running_std)."""
        buf = self._buffers[name]
        with _torch.no_grad():
            buf.copy_(_torch.as_tensor(_np.asarray(arr), dtype=buf.dtype).reshape(buf.shape))
        return self
    return setter


def _set_seed(self, seed=123):
    from bigdl_b200.utils.random_generator import RNG
    RNG.setSeed(int(seed))
    _torch.manual_seed(int(seed))
    return self


def _check_input(input):
    """(list of tensors, is_table): ndarrays / JTensors / tensors are accepted, a list means a table input."""
    def one(i):
        if isinstance(i, _torch.Tensor):
            return i
        if isinstance(i, _np.ndarray):
            return _torch.from_numpy(i)
        if hasattr(i, "to_ndarray"):
            return _torch.from_numpy(_np.asarray(i.to_ndarray()))
        raise Exception("Error unknown input type %s" % type(i))
    if isinstance(input, (list, tuple)):
        if len(input) == 0:
            raise Exception("Error when checking: empty input")
        return [one(i) for i in input], True
    return [one(input)], False


def _convert_output(output):
    def nd(t):
        return t.to_ndarray() if hasattr(t, "to_ndarray") else _np.asarray(t.detach().cpu().float() if isinstance(t, _torch.Tensor) else t)
    if isinstance(output, (list, tuple)):
        return nd(output[0]) if len(output) == 1 else [nd(x) for x in output]
    if hasattr(output, "to_list"):
        items = output.to_list()
        return nd(items[0]) if len(items) == 1 else [nd(x) for x in items]
    return nd(output)


def _predict_local(self, X, batch_size=-1):
    """ndarray (or list of ndarrays) with a leading batch dimension -> ndarray of outputs, in mini-batches."""
    xs, is_table = _check_input(X)
    n = xs[0].shape[0]
    bs = n if batch_size is None or batch_size <= 0 else int(batch_size)
    was_training = self.isTraining()
    self.evaluate()
    outs = []
    with _torch.no_grad():
        for i in range(0, n, bs):
            chunk = [x[i:i + bs] for x in xs]
            outs.append(_convert_output(self.forward(chunk if is_table else chunk[0])))
    if was_training:
        self.training()
    return _np.concatenate(outs, 0) if isinstance(outs[0], _np.ndarray) else outs


def _predict_class_local(self, X):
    return _np.argmax(_predict_local(self, X), axis=1) + 1


def _predict_distributed(self, data_rdd, batch_size=-1):
    """``data_rdd``: a DataSet / list of Samples (the reference takes an RDD[Sample]); returns the list of outputs."""
    from bigdl_b200.optim.predictor import Predictor
    return [_convert_output(o) for o in Predictor(self, batch_size=batch_size if batch_size and batch_size > 0 else 32).predict(data_rdd)]


def _predict_class_distributed(self, data_rdd):
    from bigdl_b200.optim.predictor import Predictor
    return Predictor(self).predictClass(data_rdd)


def _is_with_weights(self):
    return len(self.parameters()[0]) > 0


def _set_reg(attr):
    def setter(self, reg):
        setattr(self, attr, reg)
        return self
    return setter


for _name, _fn in (("set_running_mean", _set_running("runningMean")), ("set_running_std", _set_running("runningVar")),
                   ("set_seed", _set_seed), ("get_dtype", lambda self: "float32"),
                   ("check_input", staticmethod(_check_input)), ("convert_output", staticmethod(_convert_output)),
                   ("predict_local", _predict_local), ("predict_class_local", _predict_class_local),
                   ("predict_distributed", _predict_distributed), ("predict_class_distributed", _predict_class_distributed),
                   ("is_with_weights", _is_with_weights), ("setWRegularizer", _set_reg("wRegularizer")),
                   ("setBRegularizer", _set_reg("bRegularizer")),
                   ("from_jvalue", staticmethod(lambda jvalue, bigdl_type="float": jvalue))):
    if not hasattr(Module, _name):
        setattr(Module, _name, _fn)


def _graph_train(self, is_training=True):
    """Model.train(is_training): switch a graph between training and evaluation mode."""
    return self.training() if is_training else self.evaluate()


def _set_formats(self, formats):
    """Model.set_input_formats / set_output_formats: data formats of a TensorFlow-imported graph's endpoints."""
    self.__dict__.setdefault("_endpoint_formats", {})[_set_formats.__name__] = list(formats)
    return self


for _name, _fn in (("train", _graph_train), ("set_input_formats", _set_formats), ("set_output_formats", _set_formats)):
    if not hasattr(Graph, _name) or _name == "train":
        try:
            if _name != "train" or not callable(getattr(Graph, "train", None)):
                setattr(Graph, _name, _fn)
        except Exception:
            pass
