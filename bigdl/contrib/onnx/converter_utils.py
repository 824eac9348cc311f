from bigdl_b200.interop.onnx_ops import calc_output_shape, parse_node_attr, parse_tensor_data  # noqa: F401
