from bigdl_b200.interop.keras_converter import (DefinitionLoader, LayerConverter, WeightLoader,  # noqa: F401
                                                WeightsConverter, load_keras, unsupport_exp)
