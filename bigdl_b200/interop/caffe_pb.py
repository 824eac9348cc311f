"""The subset of ``caffe.proto`` (package ``caffe``) the loader and persister use, built at runtime. Field numbers,
enum values and defaults are those of BVLC Caffe's caffe.proto, which the reference compiles into ``Caffe.java``."""
from .protobuild import build

_ENUMS = {
    "Phase": {"TRAIN": 0, "TEST": 1},
    "PoolMethod": {"MAX": 0, "AVE": 1, "STOCHASTIC": 2},
    "EltwiseOp": {"PROD": 0, "SUM": 1, "MAXOP": 2},
    "NormRegion": {"ACROSS_CHANNELS": 0, "WITHIN_CHANNEL": 1},
    "V1LayerType": {"NONE": 0, "ABSVAL": 35, "ACCURACY": 1, "ARGMAX": 30, "BNLL": 2, "CONCAT": 3, "CONTRASTIVE_LOSS": 37,
                    "CONVOLUTION": 4, "DATA": 5, "DECONVOLUTION": 39, "DROPOUT": 6, "DUMMY_DATA": 32,
                    "EUCLIDEAN_LOSS": 7, "ELTWISE": 25, "EXP": 38, "FLATTEN": 8, "HDF5_DATA": 9, "HDF5_OUTPUT": 10,
                    "HINGE_LOSS": 28, "IM2COL": 11, "IMAGE_DATA": 12, "INFOGAIN_LOSS": 13, "INNER_PRODUCT": 14,
                    "LRN": 15, "MEMORY_DATA": 29, "MULTINOMIAL_LOGISTIC_LOSS": 16, "MVN": 34, "POOLING": 17,
                    "POWER": 26, "RELU": 18, "SIGMOID": 19, "SIGMOID_CROSS_ENTROPY_LOSS": 27, "SILENCE": 36,
                    "SOFTMAX": 20, "SOFTMAX_LOSS": 21, "SPLIT": 22, "SLICE": 33, "TANH": 23, "WINDOW_DATA": 24,
                    "THRESHOLD": 31},
}

_COMMON_PARAMS = [
    ("concat_param", "ConcatParameter"), ("convolution_param", "ConvolutionParameter"),
    ("dropout_param", "DropoutParameter"), ("eltwise_param", "EltwiseParameter"),
    ("inner_product_param", "InnerProductParameter"), ("lrn_param", "LRNParameter"),
    ("pooling_param", "PoolingParameter"), ("power_param", "PowerParameter"), ("relu_param", "ReLUParameter"),
    ("softmax_param", "SoftmaxParameter"), ("slice_param", "SliceParameter"), ("threshold_param", "ThresholdParameter"),
    ("exp_param", "ExpParameter"),
]
_V2_NUM = {"concat_param": 104, "convolution_param": 106, "dropout_param": 108, "eltwise_param": 110,
           "exp_param": 111, "inner_product_param": 117, "lrn_param": 118, "pooling_param": 121, "power_param": 122,
           "relu_param": 123, "softmax_param": 125, "slice_param": 126, "threshold_param": 128}
_V1_NUM = {"concat_param": 9, "convolution_param": 10, "dropout_param": 12, "eltwise_param": 24, "exp_param": 41,
           "inner_product_param": 17, "lrn_param": 18, "pooling_param": 19, "power_param": 21, "relu_param": 30,
           "softmax_param": 39, "slice_param": 31, "threshold_param": 25}

_MESSAGES = {
    "BlobShape": [("dim", 1, "int64", 1)],
    "BlobProto": [("shape", 7, "BlobShape", 0), ("data", 5, "float", 1), ("diff", 6, "float", 1),
                  ("double_data", 8, "double", 1), ("double_diff", 9, "double", 1), ("num", 1, "int32", 0, 0),
                  ("channels", 2, "int32", 0, 0), ("height", 3, "int32", 0, 0), ("width", 4, "int32", 0, 0)],
    "NetStateRule": [("phase", 1, "Phase", 0)],
    "ParamSpec": [("name", 1, "string", 0), ("lr_mult", 3, "float", 0, 1.0), ("decay_mult", 4, "float", 0, 1.0)],
    "NetParameter": [("name", 1, "string", 0), ("layers", 2, "V1LayerParameter", 1), ("input", 3, "string", 1),
                     ("input_dim", 4, "int32", 1), ("input_shape", 8, "BlobShape", 1),
                     ("force_backward", 5, "bool", 0, False), ("layer", 100, "LayerParameter", 1)],
    "LayerParameter": [("name", 1, "string", 0), ("type", 2, "string", 0), ("bottom", 3, "string", 1),
                       ("top", 4, "string", 1), ("phase", 10, "Phase", 0), ("loss_weight", 5, "float", 1),
                       ("param", 6, "ParamSpec", 1), ("blobs", 7, "BlobProto", 1), ("include", 8, "NetStateRule", 1),
                       ("exclude", 9, "NetStateRule", 1),
                       ("batch_norm_param", 139, "BatchNormParameter", 0), ("bias_param", 141, "BiasParameter", 0),
                       ("elu_param", 140, "ELUParameter", 0), ("flatten_param", 135, "FlattenParameter", 0),
                       ("input_param", 143, "InputParameter", 0), ("log_param", 134, "LogParameter", 0),
                       ("prelu_param", 131, "PReLUParameter", 0), ("reshape_param", 133, "ReshapeParameter", 0),
                       ("scale_param", 142, "ScaleParameter", 0), ("tile_param", 138, "TileParameter", 0)] +
                      [(n, _V2_NUM[n], t, 0) for n, t in _COMMON_PARAMS],
    "V1LayerParameter": [("bottom", 2, "string", 1), ("top", 3, "string", 1), ("name", 4, "string", 0),
                         ("include", 32, "NetStateRule", 1), ("exclude", 33, "NetStateRule", 1),
                         ("type", 5, "V1LayerType", 0), ("blobs", 6, "BlobProto", 1), ("blobs_lr", 7, "float", 1),
                         ("weight_decay", 8, "float", 1)] + [(n, _V1_NUM[n], t, 0) for n, t in _COMMON_PARAMS],
    "FillerParameter": [("type", 1, "string", 0, "constant"), ("value", 2, "float", 0, 0.0)],
    "ConvolutionParameter": [("num_output", 1, "uint32", 0), ("bias_term", 2, "bool", 0, True), ("pad", 3, "uint32", 1),
                             ("kernel_size", 4, "uint32", 1), ("stride", 6, "uint32", 1), ("dilation", 18, "uint32", 1),
                             ("pad_h", 9, "uint32", 0, 0), ("pad_w", 10, "uint32", 0, 0), ("kernel_h", 11, "uint32", 0),
                             ("kernel_w", 12, "uint32", 0), ("stride_h", 13, "uint32", 0), ("stride_w", 14, "uint32", 0),
                             ("group", 5, "uint32", 0, 1), ("weight_filler", 7, "FillerParameter", 0),
                             ("bias_filler", 8, "FillerParameter", 0)],
    "PoolingParameter": [("pool", 1, "PoolMethod", 0, "MAX"), ("pad", 4, "uint32", 0, 0), ("pad_h", 9, "uint32", 0, 0),
                         ("pad_w", 10, "uint32", 0, 0), ("kernel_size", 2, "uint32", 0), ("kernel_h", 5, "uint32", 0),
                         ("kernel_w", 6, "uint32", 0), ("stride", 3, "uint32", 0, 1), ("stride_h", 7, "uint32", 0),
                         ("stride_w", 8, "uint32", 0), ("global_pooling", 12, "bool", 0, False)],
    "InnerProductParameter": [("num_output", 1, "uint32", 0), ("bias_term", 2, "bool", 0, True),
                              ("axis", 5, "int32", 0, 1), ("transpose", 6, "bool", 0, False)],
    "LRNParameter": [("local_size", 1, "uint32", 0, 5), ("alpha", 2, "float", 0, 1.0), ("beta", 3, "float", 0, 0.75),
                     ("norm_region", 4, "NormRegion", 0, "ACROSS_CHANNELS"), ("k", 5, "float", 0, 1.0)],
    "DropoutParameter": [("dropout_ratio", 1, "float", 0, 0.5)],
    "BatchNormParameter": [("use_global_stats", 1, "bool", 0), ("moving_average_fraction", 2, "float", 0, 0.999),
                           ("eps", 3, "float", 0, 1e-5)],
    "ScaleParameter": [("axis", 1, "int32", 0, 1), ("num_axes", 2, "int32", 0, 1), ("bias_term", 4, "bool", 0, False)],
    "BiasParameter": [("axis", 1, "int32", 0, 1), ("num_axes", 2, "int32", 0, 1)],
    "EltwiseParameter": [("operation", 1, "EltwiseOp", 0, "SUM"), ("coeff", 2, "float", 1),
                         ("stable_prod_grad", 3, "bool", 0, True)],
    "ConcatParameter": [("axis", 2, "int32", 0, 1), ("concat_dim", 1, "uint32", 0, 1)],
    "ReLUParameter": [("negative_slope", 1, "float", 0, 0.0)],
    "SoftmaxParameter": [("axis", 2, "int32", 0, 1)],
    "InputParameter": [("shape", 1, "BlobShape", 1)],
    "PowerParameter": [("power", 1, "float", 0, 1.0), ("scale", 2, "float", 0, 1.0), ("shift", 3, "float", 0, 0.0)],
    "ReshapeParameter": [("shape", 1, "BlobShape", 0), ("axis", 2, "int32", 0, 0), ("num_axes", 3, "int32", 0, -1)],
    "FlattenParameter": [("axis", 1, "int32", 0, 1), ("end_axis", 2, "int32", 0, -1)],
    "SliceParameter": [("axis", 3, "int32", 0, 1), ("slice_point", 2, "uint32", 1), ("slice_dim", 1, "uint32", 0, 1)],
    "ELUParameter": [("alpha", 1, "float", 0, 1.0)],
    "PReLUParameter": [("filler", 1, "FillerParameter", 0), ("channel_shared", 2, "bool", 0, False)],
    "ThresholdParameter": [("threshold", 1, "float", 0, 0.0)],
    "TileParameter": [("axis", 1, "int32", 0, 1), ("tiles", 2, "int32", 0)],
    "ExpParameter": [("base", 1, "float", 0, -1.0), ("scale", 2, "float", 0, 1.0), ("shift", 3, "float", 0, 0.0)],
    "LogParameter": [("base", 1, "float", 0, -1.0), ("scale", 2, "float", 0, 1.0), ("shift", 3, "float", 0, 0.0)],
}

_PACKED = {("BlobShape", "dim"), ("BlobProto", "data"), ("BlobProto", "diff"), ("BlobProto", "double_data"),
           ("BlobProto", "double_diff")}

_cls = build("caffe", "caffe_subset.proto", _ENUMS, _MESSAGES, "proto2", _PACKED)
NetParameter, LayerParameter, V1LayerParameter, BlobProto, BlobShape = (_cls["NetParameter"], _cls["LayerParameter"],
                                                                         _cls["V1LayerParameter"], _cls["BlobProto"],
                                                                         _cls["BlobShape"])
V1_TYPE_NAMES = {v: k for k, v in _ENUMS["V1LayerType"].items()}
