from .all_reduce_parameter import AllReduceParameter, aligned_partition
from .compressed_tensor import CompressedTensor, FP16CompressedTensor, FP16SplitsCompressedTensor, SerializerInstance
from .distri_parameter_synchronizer import DistriParameterSynchronizer
from .model_broadcast import (CachedModels, DefaultModelBroadcastFactory, ModelBroadcast, ModelBroadcastFactory,
                              ModelBroadcastImp, ModelInfo, ProtoBufferModelBroadcastFactory)
from .parameter_processor import (ConstantClippingProcessor, L2NormClippingProcessor, LarsProcessor,
                                  ParameterProcessor)
from .parameter_synchronizer import ParameterSynchronizer, enable_sync_batchnorm
from .straggler import DropPolicy, kth_largest

__all__ = ["AllReduceParameter", "aligned_partition", "CompressedTensor", "FP16CompressedTensor",
           "FP16SplitsCompressedTensor", "SerializerInstance", "DistriParameterSynchronizer", "ModelBroadcast",
           "ConstantClippingProcessor", "L2NormClippingProcessor", "LarsProcessor", "ParameterProcessor",
           "ParameterSynchronizer", "enable_sync_batchnorm", "DropPolicy", "kth_largest"]
