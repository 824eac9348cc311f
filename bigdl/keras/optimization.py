from bigdl_b200.interop.keras_helpers import OptimConverter  # noqa: F401
