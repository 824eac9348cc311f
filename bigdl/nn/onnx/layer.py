from bigdl_b200.nn.onnx import Gemm, Reshape, Shape  # noqa: F401

from bigdl_b200.nn.ops import Gather  # noqa: F401,E402
from bigdl_b200.nn.tf import Const as Constant  # noqa: F401,E402
