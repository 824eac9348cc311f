"""fp8 quantised inference (reference: S/nn/quantized/*, core/bigquant, S/nn/MklInt8Convertible.scala)."""
from .layers import Linear, QuantizedModule, SpatialConvolution, SpatialDilatedConvolution
from .quantization import Desc, QuantSerializer, Quantization
from .quantizer import (CellQuantizer, ContainerQuantizer, GraphQuantizer, ModuleQuantizer, Quantizable,  # noqa: F401
                        Quantizer)

__all__ = ["Linear", "QuantizedModule", "SpatialConvolution", "SpatialDilatedConvolution", "Quantizer", "Quantization", "Desc",
           "QuantSerializer"]
