from bigdl_b200.nn.onnx import Gemm, Reshape, Shape  # noqa: F401
