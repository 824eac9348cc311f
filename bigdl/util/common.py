"""P/util/common.py surface without the JVM: engine init, ``JTensor``/``Sample`` value classes, RNG, helpers."""
from __future__ import annotations

import os
import tempfile

import numpy as np

from bigdl_b200.dataset.base import Sample as _Sample
from bigdl_b200.utils.engine import Engine
from bigdl_b200.utils.logger_filter import redirectSparkInfoLogs, show_bigdl_info_logs as _show
from bigdl_b200.utils.random_generator import RNG as _RNG


class JavaValue:
    """There is no JVM peer: the Python object is the value."""

    def __init__(self, jvalue=None, bigdl_type="float", *args):
        self.value, self.bigdl_type = jvalue if jvalue is not None else self, bigdl_type

    def jvm_class_constructor(self):
        """Name of the factory the Python API called on the JVM side (``"create" + class name``)."""
        return "create" + type(self).__name__


class JTensor:
    """ndarray wrapper with the reference's field names (storage, shape, bigdl_type, indices for sparse)."""

    def __init__(self, storage, shape, bigdl_type="float", indices=None):
        self.storage = np.asarray(storage, dtype=np.float32).reshape(-1)
        self.shape = np.asarray(shape, dtype=np.int32)
        self.bigdl_type, self.indices = bigdl_type, indices

    @classmethod
    def from_ndarray(cls, a_ndarray, bigdl_type="float"):
        if a_ndarray is None:
            return None
        a = np.asarray(a_ndarray)
        return cls(a.reshape(-1), a.shape if a.shape else (a.size,), bigdl_type)

    @classmethod
    def sparse(cls, a_ndarray, i_ndarray, shape, bigdl_type="float"):
        return cls(np.asarray(a_ndarray), shape, bigdl_type, np.asarray(i_ndarray, dtype=np.int32))

    def to_ndarray(self):
        assert self.indices is None, "sparseTensor to ndarray is not supported"
        return self.storage.reshape(tuple(int(s) for s in self.shape))

    def __repr__(self):
        return f"JTensor: storage: {self.storage}, shape: {self.shape}, {self.bigdl_type}"


class Sample(_Sample):
    @classmethod
    def from_ndarray(cls, features, labels, bigdl_type="float"):
        if isinstance(labels, (int, float)):
            labels = np.array([labels], dtype=np.float32)
        return cls(features, labels)

    @classmethod
    def from_jtensor(cls, features, labels, bigdl_type="float"):
        f = [t.to_ndarray() for t in features] if isinstance(features, (list, tuple)) else features.to_ndarray()
        l = labels.to_ndarray() if isinstance(labels, JTensor) else labels
        return cls(f, l)


class EvaluatedResult:
    def __init__(self, result, total_num, method):
        self.result, self.total_num, self.method = result, total_num, method

    def __str__(self):
        return f"Evaluated result: {self.result}, total_num: {self.total_num}, method: {self.method}"


class RNG:
    def __init__(self, bigdl_type="float"):
        self.bigdl_type = bigdl_type

    def set_seed(self, seed):
        _RNG.setSeed(seed)

    def uniform(self, a, b, size):
        n = int(np.prod(size))
        return np.array([_RNG.uniform(a, b) for _ in range(n)], dtype=np.float32).reshape(size)


def init_engine(bigdl_type="float"):
    Engine.init()


def init_executor_gateway(sc=None, bigdl_type="float"):
    pass


def get_node_and_core_number(bigdl_type="float"):
    return Engine.nodeNumber(), Engine.coreNumber()


def redire_spark_logs(bigdl_type="float", log_path=os.getcwd() + "/bigdl.log"):
    redirectSparkInfoLogs(log_path)


def show_bigdl_info_logs(bigdl_type="float"):
    _show()


def get_bigdl_conf():
    from bigdl_b200.utils import config

    return {k: config.get(k) for k in config.keys()}


def create_spark_conf():
    """No Spark here: returns the framework configuration as a plain dict."""
    return get_bigdl_conf()


def get_spark_context(conf=None):
    return None


def get_spark_sql_context(sc=None):
    return None


def get_dtype(bigdl_type):
    """numpy dtype of a bigdl_type string ("float" / "double")."""
    import numpy as np

    return np.float64 if bigdl_type == "double" else np.float32


def to_list(a):
    return a if isinstance(a, list) else [a]


def to_sample_rdd(x, y, numSlices=None):
    return [Sample.from_ndarray(x[i], y[i]) for i in range(len(x))]


def create_tmp_path():
    return os.path.join(tempfile.gettempdir(), next(tempfile._get_candidate_names()))


def text_from_path(path):
    with open(path) as f:
        return f.read()


def get_local_file(a_path):
    return a_path


def is_distributed(path):
    return "://" in path and not path.startswith("file://")


def get_activation_by_name(activation_name, activation_id=None):
    from bigdl_b200.nn.keras.layers import activation_module

    m = activation_module(activation_name)
    if m is not None and activation_id:
        m.setName(activation_id)
    return m


def callBigDlFunc(bigdl_type, name, *args):
    raise NotImplementedError(f"there is no JVM bridge in this build; '{name}' is implemented natively in bigdl_b200")


# ---- names of the py4j plumbing, kept so code that imports them still loads; there is no gateway behind them
class SingletonMixin:
    _instance = None

    @classmethod
    def instance(cls, *args, **kw):
        if cls._instance is None:
            cls._instance = cls(*args, **kw)
        return cls._instance


class GatewayWrapper(SingletonMixin):
    def __init__(self, bigdl_type="float", port=25333):
        self.value = None


class JavaCreator(SingletonMixin):
    __creator_class = ["com.intel.analytics.bigdl.python.api.PythonBigDLKeras"]

    @classmethod
    def add_creator_class(cls, jinvoker):
        JavaCreator.__creator_class.append(jinvoker)

    @classmethod
    def get_creator_class(cls):
        return list(JavaCreator.__creator_class)

    @classmethod
    def set_creator_class(cls, cclass):
        JavaCreator.__creator_class = [cclass] if isinstance(cclass, str) else list(cclass)

    def __init__(self, bigdl_type="float", gateway=None):
        self.value = []


class JActivity:
    """An activity (tensor or table) handed across the old JVM boundary: just holds the value."""

    def __init__(self, value):
        self.value = value


def callJavaFunc(func, *args):
    """The reference converts arguments to Java objects and back; here ``func`` is a Python callable."""
    return func(*args)


def extend_spark_driver_cp(sparkConf, path):
    """No Spark driver class path to extend; returns the configuration unchanged."""
    return sparkConf
