"""Runtime-built protobuf classes for the ``.bigdl`` model format.

The wire format is the reference's ``com.intel.analytics.bigdl.serialization`` package
(spark/dl/src/main/resources/serialization/bigdl.proto:1-190). There is no ``protoc`` in this environment, so the
FileDescriptorProto is assembled here from a compact table (message → [(name, number, type, label)]) and handed to the
protobuf runtime; the resulting classes read and write bytes that the reference's generated ``Bigdl.java`` accepts.
"""
from __future__ import annotations

from google.protobuf import any_pb2, descriptor_pb2, descriptor_pool, message_factory

_PKG = "com.intel.analytics.bigdl.serialization"
F = descriptor_pb2.FieldDescriptorProto

_SCALAR = {
    "string": F.TYPE_STRING, "bool": F.TYPE_BOOL, "int32": F.TYPE_INT32, "int64": F.TYPE_INT64,
    "float": F.TYPE_FLOAT, "double": F.TYPE_DOUBLE, "bytes": F.TYPE_BYTES,
}

_ENUMS = {
    "VarFormat": ["EMPTY_FORMAT", "DEFAULT", "ONE_D", "IN_OUT", "OUT_IN", "IN_OUT_KW_KH", "OUT_IN_KW_KH",
                  "GP_OUT_IN_KW_KH", "GP_IN_OUT_KW_KH", "OUT_IN_KT_KH_KW"],
    "InitMethodType": ["EMPTY_INITIALIZATION", "RANDOM_UNIFORM", "RANDOM_UNIFORM_PARAM", "RANDOM_NORMAL", "ZEROS",
                       "ONES", "CONST", "XAVIER", "BILINEARFILLER"],
    "RegularizerType": ["L1L2Regularizer", "L1Regularizer", "L2Regularizer"],
    "InputDataFormat": ["NCHW", "NHWC"],
    "TensorType": ["DENSE", "QUANT"],
    "DataType": ["INT32", "INT64", "FLOAT", "DOUBLE", "STRING", "BOOL", "CHAR", "SHORT", "BYTES", "REGULARIZER",
                 "TENSOR", "VARIABLE_FORMAT", "INITMETHOD", "MODULE", "NAME_ATTR_LIST", "ARRAY_VALUE", "DATA_FORMAT",
                 "CUSTOM", "SHAPE"],
}

# (field name, number, type, repeated?)  — type is a scalar name, an enum name, a message name, or "map:<value type>"
_MESSAGES = {
    "BigDLModule": [
        ("name", 1, "string", 0), ("subModules", 2, "BigDLModule", 1), ("weight", 3, "BigDLTensor", 0),
        ("bias", 4, "BigDLTensor", 0), ("preModules", 5, "string", 1), ("nextModules", 6, "string", 1),
        ("moduleType", 7, "string", 0), ("attr", 8, "map:AttrValue", 0), ("version", 9, "string", 0),
        ("train", 10, "bool", 0), ("namePostfix", 11, "string", 0), ("id", 12, "int32", 0),
        ("inputShape", 13, "Shape", 0), ("outputShape", 14, "Shape", 0), ("hasParameters", 15, "bool", 0),
        ("parameters", 16, "BigDLTensor", 1), ("isMklInt8Enabled", 17, "bool", 0), ("inputDimMasks", 18, "int32", 0),
        ("inputScales", 19, "AttrValue", 1), ("outputDimMasks", 20, "int32", 0), ("outputScales", 21, "AttrValue", 1),
        ("weightDimMasks", 22, "int32", 0), ("weightScales", 23, "AttrValue", 1),
    ],
    "InitMethod": [("methodType", 1, "InitMethodType", 0), ("data", 2, "double", 1)],
    "BigDLTensor": [
        ("datatype", 1, "DataType", 0), ("size", 2, "int32", 1), ("stride", 3, "int32", 1), ("offset", 4, "int32", 0),
        ("dimension", 5, "int32", 0), ("nElements", 6, "int32", 0), ("isScalar", 7, "bool", 0),
        ("storage", 8, "TensorStorage", 0), ("id", 9, "int32", 0), ("tensorType", 10, "TensorType", 0),
    ],
    "TensorStorage": [
        ("datatype", 1, "DataType", 0), ("float_data", 2, "float", 1), ("double_data", 3, "double", 1),
        ("bool_data", 4, "bool", 1), ("string_data", 5, "string", 1), ("int_data", 6, "int32", 1),
        ("long_data", 7, "int64", 1), ("bytes_data", 8, "bytes", 1), ("id", 9, "int32", 0),
    ],
    "Regularizer": [("regularizerType", 1, "RegularizerType", 0), ("regularData", 2, "double", 1)],
    "NameAttrList": [("name", 1, "string", 0), ("attr", 2, "map:AttrValue", 0)],
}

_ATTR_ARRAY = [
    ("size", 1, "int32", 0), ("datatype", 2, "DataType", 0), ("i32", 3, "int32", 1), ("i64", 4, "int64", 1),
    ("flt", 5, "float", 1), ("dbl", 6, "double", 1), ("str", 7, "string", 1), ("boolean", 8, "bool", 1),
    ("Regularizer", 9, "Regularizer", 1), ("tensor", 10, "BigDLTensor", 1), ("variableFormat", 11, "VarFormat", 1),
    ("initMethod", 12, "InitMethod", 1), ("bigDLModule", 13, "BigDLModule", 1), ("nameAttrList", 14, "NameAttrList", 1),
    ("dataFormat", 15, "InputDataFormat", 1), ("custom", 16, "google.protobuf.Any", 1), ("shape", 17, "Shape", 1),
]
_ATTR_PLAIN = [("dataType", 1, "DataType", 0), ("subType", 2, "string", 0)]
_ATTR_ONEOF = [
    ("int32Value", 3, "int32"), ("int64Value", 4, "int64"), ("floatValue", 5, "float"), ("doubleValue", 6, "double"),
    ("stringValue", 7, "string"), ("boolValue", 8, "bool"), ("regularizerValue", 9, "Regularizer"),
    ("tensorValue", 10, "BigDLTensor"), ("variableFormatValue", 11, "VarFormat"), ("initMethodValue", 12, "InitMethod"),
    ("bigDLModuleValue", 13, "BigDLModule"), ("nameAttrListValue", 14, "NameAttrList"),
    ("arrayValue", 15, "AttrValue.ArrayValue"), ("dataFormatValue", 16, "InputDataFormat"),
    ("customValue", 17, "google.protobuf.Any"), ("shape", 18, "Shape"),
]


def _set_type(fd, typ: str):
    if typ in _SCALAR:
        fd.type = _SCALAR[typ]
    elif typ in _ENUMS or typ == "Shape.ShapeType":
        fd.type = F.TYPE_ENUM
        fd.type_name = f".{_PKG}.{typ}"
    elif typ == "google.protobuf.Any":
        fd.type = F.TYPE_MESSAGE
        fd.type_name = ".google.protobuf.Any"
    else:
        fd.type = F.TYPE_MESSAGE
        fd.type_name = f".{_PKG}.{typ}"


def _add_field(msg, owner: str, name, number, typ, repeated, oneof_index=None):
    fd = msg.field.add()
    fd.name, fd.number = name, number
    fd.json_name = name
    if typ.startswith("map:"):
        entry = msg.nested_type.add()
        entry.name = name[0].upper() + name[1:] + "Entry"
        entry.options.map_entry = True
        k = entry.field.add()
        k.name, k.number, k.type, k.label, k.json_name = "key", 1, F.TYPE_STRING, F.LABEL_OPTIONAL, "key"
        v = entry.field.add()
        v.name, v.number, v.label, v.json_name = "value", 2, F.LABEL_OPTIONAL, "value"
        _set_type(v, typ[4:])
        fd.type, fd.label = F.TYPE_MESSAGE, F.LABEL_REPEATED
        fd.type_name = f".{_PKG}.{owner}.{entry.name}"
        return
    _set_type(fd, typ)
    fd.label = F.LABEL_REPEATED if repeated else F.LABEL_OPTIONAL
    if oneof_index is not None:
        fd.oneof_index = oneof_index


def _build_file() -> descriptor_pb2.FileDescriptorProto:
    f = descriptor_pb2.FileDescriptorProto()
    f.name = "bigdl_b200/bigdl.proto"
    f.package = _PKG
    f.syntax = "proto3"
    f.dependency.append("google/protobuf/any.proto")
    for ename, values in _ENUMS.items():
        e = f.enum_type.add()
        e.name = ename
        for i, v in enumerate(values):
            ev = e.value.add()
            ev.name, ev.number = v, i
    for mname, fields in _MESSAGES.items():
        m = f.message_type.add()
        m.name = mname
        for name, number, typ, rep in fields:
            _add_field(m, mname, name, number, typ, rep)
    # AttrValue with nested ArrayValue and the ``value`` oneof
    av = f.message_type.add()
    av.name = "AttrValue"
    arr = av.nested_type.add()
    arr.name = "ArrayValue"
    for name, number, typ, rep in _ATTR_ARRAY:
        _add_field(arr, "AttrValue.ArrayValue", name, number, typ, rep)
    for name, number, typ, rep in _ATTR_PLAIN:
        _add_field(av, "AttrValue", name, number, typ, rep)
    av.oneof_decl.add().name = "value"
    for name, number, typ in _ATTR_ONEOF:
        _add_field(av, "AttrValue", name, number, typ, 0, oneof_index=0)
    # Shape with nested enum
    sh = f.message_type.add()
    sh.name = "Shape"
    se = sh.enum_type.add()
    se.name = "ShapeType"
    for i, v in enumerate(["SINGLE", "MULTI"]):
        ev = se.value.add()
        ev.name, ev.number = v, i
    for name, number, typ, rep in [("shapeType", 1, "Shape.ShapeType", 0), ("ssize", 2, "int32", 0),
                                   ("shapeValue", 3, "int32", 1), ("shape", 4, "Shape", 1)]:
        _add_field(sh, "Shape", name, number, typ, rep)
    return f


_pool = descriptor_pool.Default()
try:
    _file = _pool.FindFileByName("bigdl_b200/bigdl.proto")
except KeyError:
    _ = any_pb2.DESCRIPTOR          # make sure the dependency is registered
    _file = _pool.Add(_build_file()) or _pool.FindFileByName("bigdl_b200/bigdl.proto")


def _cls(name: str):
    return message_factory.GetMessageClass(_pool.FindMessageTypeByName(f"{_PKG}.{name}"))


BigDLModule = _cls("BigDLModule")
BigDLTensor = _cls("BigDLTensor")
TensorStorage = _cls("TensorStorage")
AttrValue = _cls("AttrValue")
ArrayValue = _cls("AttrValue.ArrayValue")
NameAttrList = _cls("NameAttrList")
InitMethod = _cls("InitMethod")
Regularizer = _cls("Regularizer")
Shape = _cls("Shape")


class _Enum:
    def __init__(self, name):
        self._d = _pool.FindEnumTypeByName(f"{_PKG}.{name}")
        for v in self._d.values:
            setattr(self, v.name, v.number)

    def Name(self, number: int) -> str:
        return self._d.values_by_number[number].name

    def Value(self, name: str) -> int:
        return self._d.values_by_name[name].number


DataType = _Enum("DataType")
TensorType = _Enum("TensorType")
VarFormat = _Enum("VarFormat")
InitMethodType = _Enum("InitMethodType")
RegularizerType = _Enum("RegularizerType")
InputDataFormat = _Enum("InputDataFormat")
EOF_MARK = None
