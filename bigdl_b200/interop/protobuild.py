"""Build protobuf message classes at runtime from a compact table (no ``protoc`` in this environment).

``build(package, file_name, enums, messages, syntax)`` → {message name: class}. A field row is
``(name, number, type, repeated[, default])`` where type is a scalar name, an enum name, another message name of the same
table, or ``map:<value type>`` (string keys). proto2 files honour ``default``; repeated scalar fields are packed for
proto3 and left unpacked-but-accepting-packed for proto2, which is what the protobuf runtime does on parse anyway."""
from __future__ import annotations

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

F = descriptor_pb2.FieldDescriptorProto
_SCALAR = {"string": F.TYPE_STRING, "bool": F.TYPE_BOOL, "int32": F.TYPE_INT32, "int64": F.TYPE_INT64,
           "uint32": F.TYPE_UINT32, "uint64": F.TYPE_UINT64, "float": F.TYPE_FLOAT, "double": F.TYPE_DOUBLE,
           "bytes": F.TYPE_BYTES}


def build(package: str, file_name: str, enums: dict, messages: dict, syntax: str = "proto2", packed=()):
    fdp = descriptor_pb2.FileDescriptorProto()
    fdp.name, fdp.package, fdp.syntax = file_name, package, syntax
    for ename, values in enums.items():
        e = fdp.enum_type.add()
        e.name = ename
        items = values.items() if isinstance(values, dict) else [(v, i) for i, v in enumerate(values)]
        for vname, num in items:
            v = e.value.add()
            v.name, v.number = vname, num
    for mname, fields in messages.items():
        m = fdp.message_type.add()
        m.name = mname
        for row in fields:
            name, num, typ, rep = row[:4]
            default = row[4] if len(row) > 4 else None
            fd = m.field.add()
            fd.name, fd.number = name, num
            fd.label = F.LABEL_REPEATED if rep else F.LABEL_OPTIONAL
            if typ.startswith("map:"):
                vt = typ[4:]
                entry = m.nested_type.add()
                entry.name = "".join(p.capitalize() for p in name.split("_")) + "Entry"
                entry.options.map_entry = True
                k = entry.field.add()
                k.name, k.number, k.label, k.type = "key", 1, F.LABEL_OPTIONAL, F.TYPE_STRING
                v = entry.field.add()
                v.name, v.number, v.label = "value", 2, F.LABEL_OPTIONAL
                _set_type(v, vt, package, enums)
                fd.label, fd.type, fd.type_name = F.LABEL_REPEATED, F.TYPE_MESSAGE, f".{package}.{mname}.{entry.name}"
                continue
            _set_type(fd, typ, package, enums)
            if default is not None and syntax == "proto2":
                fd.default_value = str(default).lower() if isinstance(default, bool) else str(default)
            if (mname, name) in packed:
                fd.options.packed = True
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    out = {}
    for mname in messages:
        out[mname] = message_factory.GetMessageClass(pool.FindMessageTypeByName(f"{package}.{mname}"))
    return out


def _set_type(fd, typ, package, enums):
    if typ in _SCALAR:
        fd.type = _SCALAR[typ]
    elif typ in enums:
        fd.type, fd.type_name = F.TYPE_ENUM, f".{package}.{typ}"
    else:
        fd.type, fd.type_name = F.TYPE_MESSAGE, f".{package}.{typ}"
