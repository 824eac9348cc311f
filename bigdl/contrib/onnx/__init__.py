from bigdl_b200.interop.onnx_loader import load, load_model, load_model_proto  # noqa: F401
