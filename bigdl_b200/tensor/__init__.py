from .dnn import DnnTensor
from .quantized import QuantizedTensor
from .sparse import SparseTensor
from .tensor import Storage, Tensor, cat

__all__ = ["Tensor", "Storage", "SparseTensor", "QuantizedTensor", "DnnTensor", "cat"]
