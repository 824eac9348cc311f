"""Runtime-built subset of TensorFlow's graph protos (graph.proto, node_def.proto, attr_value.proto, tensor.proto,
tensor_shape.proto, types.proto; package ``tensorflow``) — what the reference compiles into ``org.tensorflow.framework``
Java classes for its TensorflowLoader / TensorflowSaver."""
import numpy as np

from .protobuild import build

DT = {"DT_INVALID": 0, "DT_FLOAT": 1, "DT_DOUBLE": 2, "DT_INT32": 3, "DT_UINT8": 4, "DT_INT16": 5, "DT_INT8": 6,
      "DT_STRING": 7, "DT_COMPLEX64": 8, "DT_INT64": 9, "DT_BOOL": 10, "DT_QINT8": 11, "DT_QUINT8": 12,
      "DT_QINT32": 13, "DT_BFLOAT16": 14, "DT_UINT16": 17, "DT_HALF": 19, "DT_RESOURCE": 20}

_MESSAGES = {
    "TensorShapeDim": [("size", 1, "int64", 0), ("name", 2, "string", 0)],
    "TensorShapeProto": [("dim", 2, "TensorShapeDim", 1), ("unknown_rank", 3, "bool", 0)],
    "TensorProto": [("dtype", 1, "DataType", 0), ("tensor_shape", 2, "TensorShapeProto", 0),
                    ("version_number", 3, "int32", 0), ("tensor_content", 4, "bytes", 0), ("half_val", 13, "int32", 1),
                    ("float_val", 5, "float", 1), ("double_val", 6, "double", 1), ("int_val", 7, "int32", 1),
                    ("string_val", 8, "bytes", 1), ("int64_val", 10, "int64", 1), ("bool_val", 11, "bool", 1)],
    "NameAttrList": [("name", 1, "string", 0), ("attr", 2, "map:AttrValue", 0)],
    "ListValue": [("s", 2, "bytes", 1), ("i", 3, "int64", 1), ("f", 4, "float", 1), ("b", 5, "bool", 1),
                  ("type", 6, "DataType", 1), ("shape", 7, "TensorShapeProto", 1), ("tensor", 8, "TensorProto", 1)],
    "AttrValue": [("list", 1, "ListValue", 0), ("s", 2, "bytes", 0), ("i", 3, "int64", 0), ("f", 4, "float", 0),
                  ("b", 5, "bool", 0), ("type", 6, "DataType", 0), ("shape", 7, "TensorShapeProto", 0),
                  ("tensor", 8, "TensorProto", 0), ("placeholder", 9, "string", 0), ("func", 10, "NameAttrList", 0)],
    "NodeDef": [("name", 1, "string", 0), ("op", 2, "string", 0), ("input", 3, "string", 1), ("device", 4, "string", 0),
                ("attr", 5, "map:AttrValue", 0)],
    "VersionDef": [("producer", 1, "int32", 0), ("min_consumer", 2, "int32", 0)],
    "GraphDef": [("node", 1, "NodeDef", 1), ("versions", 4, "VersionDef", 0), ("version", 3, "int32", 0)],
}
# proto2 syntax on purpose: identical wire format, but scalar attr fields keep presence (``HasField``), which the
# loader needs to tell "i: 0" from "not set". Repeated numeric fields are marked packed like TensorFlow writes them.
_PACKED = {(m, f[0]) for m, fs in _MESSAGES.items() for f in fs
           if f[3] and f[2] in ("int32", "int64", "float", "double", "bool", "DataType")}
_cls = build("tensorflow", "tf_subset.proto", {"DataType": DT}, _MESSAGES, "proto2", _PACKED)
GraphDef, NodeDef, AttrValue, TensorProto, TensorShapeProto = (_cls["GraphDef"], _cls["NodeDef"], _cls["AttrValue"],
                                                               _cls["TensorProto"], _cls["TensorShapeProto"])

NP_OF = {1: np.float32, 2: np.float64, 3: np.int32, 4: np.uint8, 5: np.int16, 6: np.int8, 9: np.int64, 10: np.bool_,
         17: np.uint16, 19: np.float16}
DT_OF_NP = {np.dtype(v): k for k, v in NP_OF.items()}


def tensor_to_numpy(t, byte_order="little_endian"):
    shape = [int(d.size) for d in t.tensor_shape.dim]
    if t.dtype == 7:
        vals = list(t.string_val)
        return np.array(vals, dtype=object).reshape(shape) if shape else (vals[0] if vals else b"")
    dt = np.dtype(NP_OF[t.dtype])
    n = int(np.prod(shape)) if shape else 1
    if t.tensor_content:
        arr = np.frombuffer(t.tensor_content, dtype=dt.newbyteorder("<" if byte_order == "little_endian" else ">"))
        arr = arr.astype(dt)
    else:
        src = {1: t.float_val, 2: t.double_val, 3: t.int_val, 4: t.int_val, 5: t.int_val, 6: t.int_val,
               9: t.int64_val, 10: t.bool_val, 17: t.int_val, 19: t.half_val}[t.dtype]
        vals = list(src)
        if t.dtype == 19:
            arr = np.array(vals, dtype=np.uint16).view(np.float16)
        else:
            arr = np.array(vals, dtype=dt)
        if arr.size == 1 and n > 1:
            arr = np.full(n, arr[0], dtype=dt)           # scalar splat
        elif arr.size == 0:
            arr = np.zeros(n, dtype=dt)
    return arr.reshape(shape).copy()


def numpy_to_tensor(arr) -> "TensorProto":
    arr = np.asarray(arr)
    t = TensorProto()
    if arr.dtype == object or arr.dtype.kind in "US":
        t.dtype = 7
        for v in arr.reshape(-1):
            t.string_val.append(v if isinstance(v, bytes) else str(v).encode())
    else:
        t.dtype = DT_OF_NP[arr.dtype]
        t.tensor_content = np.ascontiguousarray(arr).astype(arr.dtype.newbyteorder("<")).tobytes()
    for s in arr.shape:
        t.tensor_shape.dim.add().size = int(s)
    return t
