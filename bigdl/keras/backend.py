from bigdl_b200.interop.keras_backend import KerasModelWrapper, with_bigdl_backend  # noqa: F401
