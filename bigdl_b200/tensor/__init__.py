from .dnn import DnnTensor
from .quantized import QuantizedTensor
from .sparse import SparseTensor
from .tensor import Storage, Tensor, cat

__all__ = ["Tensor", "Storage", "SparseTensor", "QuantizedTensor", "DnnTensor", "cat"]

from .tensor import Tensor as DenseTensor  # noqa: E402,F401  (the reference's concrete dense class)
from .tensor import Storage as ArrayStorage  # noqa: E402,F401
from .tensor import Tensor as TensorMath  # noqa: E402,F401  (trait of the math methods: they are Tensor methods here)
from .tensor import Tensor as DenseTensorMath  # noqa: E402,F401
from .types import *  # noqa: E402,F401,F403
from .types import (BooleanType, ByteType, CharType, DenseType, DoubleType, FloatType, IntType, LongType,  # noqa: E402,F401
                    MklDnnType, QuantizedType, ShortType, SparseType, StringType, TensorDataType, TensorType)
