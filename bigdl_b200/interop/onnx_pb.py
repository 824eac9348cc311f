"""Runtime-built subset of ``onnx.proto`` (package ``onnx``)."""
import numpy as np

from .protobuild import build

_MESSAGES = {
    "TensorShapeDim": [("dim_value", 1, "int64", 0), ("dim_param", 2, "string", 0)],
    "TensorShapeProto": [("dim", 1, "TensorShapeDim", 1)],
    "TypeTensor": [("elem_type", 1, "int32", 0), ("shape", 2, "TensorShapeProto", 0)],
    "TypeProto": [("tensor_type", 1, "TypeTensor", 0)],
    "ValueInfoProto": [("name", 1, "string", 0), ("type", 2, "TypeProto", 0), ("doc_string", 3, "string", 0)],
    "TensorProto": [("dims", 1, "int64", 1), ("data_type", 2, "int32", 0), ("float_data", 4, "float", 1),
                    ("int32_data", 5, "int32", 1), ("string_data", 6, "bytes", 1), ("int64_data", 7, "int64", 1),
                    ("name", 8, "string", 0), ("raw_data", 9, "bytes", 0), ("double_data", 10, "double", 1)],
    "AttributeProto": [("name", 1, "string", 0), ("f", 2, "float", 0), ("i", 3, "int64", 0), ("s", 4, "bytes", 0),
                       ("t", 5, "TensorProto", 0), ("floats", 7, "float", 1), ("ints", 8, "int64", 1),
                       ("strings", 9, "bytes", 1), ("tensors", 10, "TensorProto", 1), ("type", 20, "int32", 0)],
    "NodeProto": [("input", 1, "string", 1), ("output", 2, "string", 1), ("name", 3, "string", 0),
                  ("op_type", 4, "string", 0), ("attribute", 5, "AttributeProto", 1), ("domain", 7, "string", 0)],
    "GraphProto": [("node", 1, "NodeProto", 1), ("name", 2, "string", 0), ("initializer", 5, "TensorProto", 1),
                   ("input", 11, "ValueInfoProto", 1), ("output", 12, "ValueInfoProto", 1),
                   ("value_info", 13, "ValueInfoProto", 1)],
    "OperatorSetIdProto": [("domain", 1, "string", 0), ("version", 2, "int64", 0)],
    "ModelProto": [("ir_version", 1, "int64", 0), ("producer_name", 2, "string", 0), ("producer_version", 3, "string", 0),
                   ("domain", 4, "string", 0), ("model_version", 5, "int64", 0), ("graph", 7, "GraphProto", 0),
                   ("opset_import", 8, "OperatorSetIdProto", 1)],
}
_PACKED = {(m, f[0]) for m, fs in _MESSAGES.items() for f in fs
           if f[3] and f[2] in ("int32", "int64", "float", "double")}
_cls = build("onnx", "onnx_subset.proto", {}, _MESSAGES, "proto2", _PACKED)
ModelProto, GraphProto, NodeProto, TensorProto, AttributeProto, ValueInfoProto = (
    _cls["ModelProto"], _cls["GraphProto"], _cls["NodeProto"], _cls["TensorProto"], _cls["AttributeProto"],
    _cls["ValueInfoProto"])

_NP = {1: np.float32, 2: np.uint8, 3: np.int8, 5: np.int16, 6: np.int32, 7: np.int64, 9: np.bool_, 10: np.float16,
       11: np.float64}
_DT = {np.dtype(v): k for k, v in _NP.items()}


def to_numpy(t) -> np.ndarray:
    dt = np.dtype(_NP[t.data_type])
    dims = [int(d) for d in t.dims]
    if t.raw_data:
        arr = np.frombuffer(t.raw_data, dtype=dt.newbyteorder("<")).astype(dt)
    elif t.data_type == 1:
        arr = np.array(t.float_data, dtype=dt)
    elif t.data_type == 11:
        arr = np.array(t.double_data, dtype=dt)
    elif t.data_type == 7:
        arr = np.array(t.int64_data, dtype=dt)
    else:
        arr = np.array(t.int32_data, dtype=dt)
    return arr.reshape(dims).copy()


def from_numpy(arr, name="") -> "TensorProto":
    arr = np.ascontiguousarray(arr)
    t = TensorProto()
    t.name, t.data_type = name, _DT[arr.dtype]
    t.dims.extend(arr.shape)
    t.raw_data = arr.astype(arr.dtype.newbyteorder("<")).tobytes()
    return t
