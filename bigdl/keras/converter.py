from bigdl_b200.interop.keras_converter import DefinitionLoader, WeightLoader, load_keras  # noqa: F401
