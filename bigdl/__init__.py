"""``bigdl`` — drop-in import surface for code written against the reference's Python API (pyspark/bigdl/*).

Every module here re-exports the B200-native implementation from ``bigdl_b200`` under the reference's module paths and
class names: ``bigdl.nn.layer``, ``bigdl.nn.criterion``, ``bigdl.nn.initialization_method``, ``bigdl.nn.keras.*``,
``bigdl.optim.optimizer``, ``bigdl.util.common``, ``bigdl.util.engine``, ``bigdl.util.tf_utils``, ``bigdl.dataset.*``,
``bigdl.transform.vision.image``, ``bigdl.dlframes.*``, ``bigdl.models.*``, ``bigdl.keras.converter``,
``bigdl.contrib.onnx``. There is no py4j/JVM bridge: "JavaValue" objects are the Python objects themselves, RDDs are
Python sequences (sharded across ranks by the distributed dataset), and ``bigdl_type`` arguments are accepted and ignored.
"""
from bigdl_b200 import __version__ as __version__  # noqa: F401
