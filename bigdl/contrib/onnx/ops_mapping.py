from bigdl_b200.interop.onnx_ops import _convert_map  # noqa: F401
