"""``saveModule`` / ``loadModule``: the portable ``.bigdl`` protobuf model format.

Reference: S/utils/serializer/ModuleSerializable.scala:150-310 (ctor-parameter reflection → ``attr[paramName]``,
``parameters`` + root attr ``global_storage`` so shared storages are written once), ModuleSerializer.scala (class
registry), ModuleLoader.scala:40-117,257-330 (protobuf-only, or protobuf + separate weight file
``[int magic 3721][int count]{[int id][int dtype][int size][data]}…[int md5len][md5]``), converters/*.

Mapping rules used here:
  * ``moduleType`` = ``com.intel.analytics.bigdl.nn.<ClassName>`` (JVM class name, so files interoperate by name);
  * ctor arguments are recorded at construction (``Module.__init_subclass__`` hook below) and written under the
    camelCase form of the Python keyword (``input_size`` → ``inputSize``), which is how the reference's Python API maps
    onto the Scala constructors; loading converts back;
  * tensors: 1-based ``offset``, ``size``/``stride`` lists, DENSE/QUANT tensor types, FLOAT storage by default;
  * weights are the module's ``parameters()`` in order (conv weights are written in the reference's
    (nGroup, out/g, in/g, kH, kW) logical order, not our OHWI physical order).
"""
from __future__ import annotations

import functools
import hashlib
import inspect
import os
import re
import struct
import sys

import numpy as np
import torch

from ..utils.table import Table
from . import bigdl_pb as pb

_JVM_PREFIX = "com.intel.analytics.bigdl.nn."
MAGIC_NO = 3721


def _camel(s: str) -> str:
    parts = s.split("_")
    return parts[0] + "".join(p[:1].upper() + p[1:] for p in parts[1:])


def _snake(s: str) -> str:
    return re.sub(r"(?<!^)(?=[A-Z])", "_", s).lower()


# ---------------------------------------------------------------- tensors
class _Ctx:
    def __init__(self):
        self.storages: dict[int, pb.TensorStorage] = {}     # storage id → TensorStorage (with data)
        self.tensors: dict[int, pb.BigDLTensor] = {}        # tensor id → BigDLTensor (no data)
        self.by_ptr: dict[tuple, int] = {}
        self.next_id = 1

    def new_id(self):
        self.next_id += 1
        return self.next_id


# Fixed-width packed fields of TensorStorage go in and out as raw little-endian bytes (one wire-format record: tag,
# byte length, payload) instead of one Python object per element — a VGG-16 snapshot is 134 M floats.
_RAW_FIELDS = {"float_data": (2, np.float32), "double_data": (3, np.float64), "bool_data": (4, np.uint8)}
_RAW_MIN = 256


def _varint(n: int) -> bytes:
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        out.append(b | (0x80 if n else 0))
        if not n:
            return bytes(out)


def packed_merge(msg, field_number: int, arr: np.ndarray):
    """Append ``arr`` to the packed fixed-width repeated field ``field_number`` of ``msg`` as one wire record."""
    raw = np.ascontiguousarray(arr).tobytes()
    msg.MergeFromString(_varint(field_number << 3 | 2) + _varint(len(raw)) + raw)


def _fill_storage(st, field: str, flat: torch.Tensor):
    if field in _RAW_FIELDS and flat.numel() >= _RAW_MIN and sys.byteorder == "little":
        num, npdt = _RAW_FIELDS[field]
        packed_merge(st, num, flat.numpy().astype(npdt, copy=False))
    else:
        getattr(st, field).extend(flat.tolist())


def _read_varint(buf, pos):
    v = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, pos
        shift += 7


def packed_read(msg, field_number: int, npdt) -> np.ndarray:
    """All values of the packed fixed-width repeated field ``field_number`` of ``msg``, read from its wire bytes."""
    buf = memoryview(msg.SerializeToString())
    pos, parts = 0, []
    while pos < len(buf):
        tag, pos = _read_varint(buf, pos)
        num, wt = tag >> 3, tag & 7
        if wt == 0:
            _, pos = _read_varint(buf, pos)
        elif wt == 1:
            pos += 8
        elif wt == 5:
            pos += 4
        elif wt == 2:
            n, pos = _read_varint(buf, pos)
            if num == field_number:
                parts.append(np.frombuffer(buf[pos:pos + n], dtype=npdt))
            pos += n
        else:
            raise ValueError(f"unexpected wire type {wt} in {type(msg).__name__}")
    if not parts:
        return np.zeros(0, dtype=npdt)
    return parts[0].copy() if len(parts) == 1 else np.concatenate(parts)


def _storage_flat(st, field: str, dtype) -> torch.Tensor:
    """The ``field`` values of a TensorStorage as a flat tensor."""
    vals = getattr(st, field)
    if field not in _RAW_FIELDS or len(vals) < _RAW_MIN or sys.byteorder != "little":
        return torch.tensor(list(vals), dtype=dtype)
    want, npdt = _RAW_FIELDS[field]
    return torch.from_numpy(packed_read(st, want, npdt)).to(dtype)


def tensor_to_proto(t: torch.Tensor, ctx: _Ctx | None, inline: bool = False) -> pb.BigDLTensor:
    t = t.detach().cpu()
    out = pb.BigDLTensor()
    if t.dtype in (torch.float32, torch.bfloat16, torch.float16):
        dt, field, data = pb.DataType.FLOAT, "float_data", t.float()
    elif t.dtype == torch.float64:
        dt, field, data = pb.DataType.DOUBLE, "double_data", t
    elif t.dtype in (torch.int32, torch.int16, torch.int8, torch.uint8):
        dt, field, data = pb.DataType.INT32, "int_data", t.int()
    elif t.dtype == torch.int64:
        dt, field, data = pb.DataType.INT64, "long_data", t
    elif t.dtype == torch.bool:
        dt, field, data = pb.DataType.BOOL, "bool_data", t
    else:
        raise TypeError(f"unsupported tensor dtype {t.dtype}")
    data = data.contiguous()
    out.datatype = dt
    out.size.extend(list(data.shape))
    out.stride.extend(list(data.stride()))
    out.offset = 1                      # 1-based storage offset
    out.dimension = data.dim()
    out.nElements = data.numel()
    out.isScalar = data.dim() == 0
    out.tensorType = pb.TensorType.DENSE
    tid = ctx.new_id() if ctx else 0
    sid = ctx.new_id() if ctx else 0
    out.id = tid
    if ctx is None or inline:
        out.storage.datatype, out.storage.id = dt, sid
        _fill_storage(out.storage, field, data.reshape(-1))
    else:
        st = pb.TensorStorage(datatype=dt, id=sid)
        _fill_storage(st, field, data.reshape(-1))
        out.storage.datatype, out.storage.id = dt, sid
        ctx.storages[sid] = st
        ctx.tensors[tid] = out
    return out


def proto_to_tensor(p: pb.BigDLTensor, storages: dict[int, torch.Tensor]) -> torch.Tensor | None:
    sid = p.storage.id
    st = p.storage
    flat = None
    for field, dtype in (("float_data", torch.float32), ("double_data", torch.float64), ("int_data", torch.int32),
                         ("long_data", torch.int64), ("bool_data", torch.bool)):
        vals = getattr(st, field)
        if len(vals):
            flat = _storage_flat(st, field, dtype)
            break
    if flat is None:
        flat = storages.get(sid)
        if flat is None:
            if p.nElements == 0:
                return torch.zeros(list(p.size))
            return None
    else:
        storages.setdefault(sid, flat)
    if p.isScalar or p.dimension == 0:
        return flat[max(p.offset - 1, 0)].clone()
    size, stride = list(p.size), list(p.stride)
    if not stride:
        stride = list(torch.empty(size).stride())
    return torch.as_strided(flat, size, stride, max(p.offset - 1, 0)).clone()


# ---------------------------------------------------------------- attribute values
def _set_attr(av: pb.AttrValue, v, ctx: _Ctx):
    from ..nn.initialization_method import InitializationMethod
    from ..nn.module import Module
    from ..optim.regularizer import L1L2Regularizer

    if isinstance(v, bool):
        av.dataType, av.boolValue = pb.DataType.BOOL, v
    elif isinstance(v, int):
        av.dataType, av.int32Value = pb.DataType.INT32, v
    elif isinstance(v, float):
        av.dataType, av.doubleValue = pb.DataType.DOUBLE, v
    elif isinstance(v, str):
        av.dataType, av.stringValue = pb.DataType.STRING, v
    elif isinstance(v, (torch.Tensor, np.ndarray)):
        av.dataType = pb.DataType.TENSOR
        av.tensorValue.CopyFrom(tensor_to_proto(torch.as_tensor(v), ctx, inline=True))
    elif isinstance(v, Module):
        av.dataType = pb.DataType.MODULE
        av.bigDLModuleValue.CopyFrom(_module_to_proto(v, ctx))
    elif isinstance(v, L1L2Regularizer):
        av.dataType = pb.DataType.REGULARIZER
        av.regularizerValue.regularizerType = pb.RegularizerType.L1L2Regularizer
        av.regularizerValue.regularData.extend([v.l1, v.l2])
    elif isinstance(v, InitializationMethod):
        av.dataType = pb.DataType.INITMETHOD
        name = type(v).__name__
        kinds = {"RandomUniform": pb.InitMethodType.RANDOM_UNIFORM, "RandomNormal": pb.InitMethodType.RANDOM_NORMAL,
                 "Zeros": pb.InitMethodType.ZEROS, "Ones": pb.InitMethodType.ONES, "ConstInitMethod": pb.InitMethodType.CONST,
                 "Xavier": pb.InitMethodType.XAVIER, "BilinearFiller": pb.InitMethodType.BILINEARFILLER}
        kind = kinds.get(name, pb.InitMethodType.EMPTY_INITIALIZATION)
        if name == "RandomUniform" and v.lower is not None:
            kind = pb.InitMethodType.RANDOM_UNIFORM_PARAM
            av.initMethodValue.data.extend([v.lower, v.upper])
        elif name == "RandomNormal":
            av.initMethodValue.data.extend([v.mean, v.stdv])
        elif name == "ConstInitMethod":
            av.initMethodValue.data.append(v.value)
        av.initMethodValue.methodType = kind
    elif isinstance(v, (list, tuple)):
        av.dataType = pb.DataType.ARRAY_VALUE
        arr = av.arrayValue
        arr.size = len(v)
        if all(isinstance(x, bool) for x in v) and v:
            arr.datatype = pb.DataType.BOOL
            arr.boolean.extend(v)
        elif all(isinstance(x, int) for x in v):
            arr.datatype = pb.DataType.INT32
            arr.i32.extend(v)
        elif all(isinstance(x, (int, float)) for x in v):
            arr.datatype = pb.DataType.DOUBLE
            arr.dbl.extend(float(x) for x in v)
        elif all(isinstance(x, str) for x in v):
            arr.datatype = pb.DataType.STRING
            arr.str.extend(v)
        elif all(isinstance(x, Module) for x in v):
            arr.datatype = pb.DataType.MODULE
            for x in v:
                arr.bigDLModule.add().CopyFrom(_module_to_proto(x, ctx))
        elif all(isinstance(x, (torch.Tensor, np.ndarray)) for x in v):
            arr.datatype = pb.DataType.TENSOR
            for x in v:
                arr.tensor.add().CopyFrom(tensor_to_proto(torch.as_tensor(x), ctx, inline=True))
        else:      # nested / mixed python structure: keep it as a custom JSON string
            import json

            av.dataType, av.subType = pb.DataType.CUSTOM, "json"
            av.ClearField("arrayValue")
            av.stringValue = json.dumps(_jsonable(v))
    else:
        raise TypeError(f"cannot serialise ctor argument of type {type(v)}")


def _jsonable(v):
    if isinstance(v, (list, tuple)):
        return [_jsonable(x) for x in v]
    if isinstance(v, (torch.Tensor, np.ndarray)):
        return np.asarray(v).tolist()
    return v


def _get_attr(av: pb.AttrValue, storages):
    from ..nn import initialization_method as im
    from ..optim.regularizer import L1L2Regularizer

    dt = av.dataType
    which = av.WhichOneof("value")
    if which is None:
        return None
    if dt == pb.DataType.CUSTOM and av.subType == "json":
        import json

        return json.loads(av.stringValue)
    if which in ("int32Value", "int64Value", "floatValue", "doubleValue", "stringValue", "boolValue"):
        return getattr(av, which)
    if which == "tensorValue":
        return proto_to_tensor(av.tensorValue, storages)
    if which == "bigDLModuleValue":
        return _proto_to_module(av.bigDLModuleValue, storages)
    if which == "regularizerValue":
        d = list(av.regularizerValue.regularData)
        return L1L2Regularizer(d[0], d[1]) if len(d) >= 2 else None
    if which == "initMethodValue":
        k, d = av.initMethodValue.methodType, list(av.initMethodValue.data)
        T = pb.InitMethodType
        return {T.RANDOM_UNIFORM: lambda: im.RandomUniform(), T.RANDOM_UNIFORM_PARAM: lambda: im.RandomUniform(d[0], d[1]),
                T.RANDOM_NORMAL: lambda: im.RandomNormal(d[0], d[1]), T.ZEROS: lambda: im.Zeros(), T.ONES: lambda: im.Ones(),
                T.CONST: lambda: im.ConstInitMethod(d[0]), T.XAVIER: lambda: im.Xavier(),
                T.BILINEARFILLER: lambda: im.BilinearFiller()}.get(k, lambda: None)()
    if which == "dataFormatValue":
        return "NCHW" if av.dataFormatValue == pb.InputDataFormat.NCHW else "NHWC"
    if which == "arrayValue":
        a = av.arrayValue
        for field in ("i32", "i64", "flt", "dbl", "str", "boolean"):
            vals = list(getattr(a, field))
            if vals:
                return vals
        if len(a.bigDLModule):
            return [_proto_to_module(m, storages) for m in a.bigDLModule]
        if len(a.tensor):
            return [proto_to_tensor(t, storages) for t in a.tensor]
        return []
    return None


# ---------------------------------------------------------------- modules
def _logical_param(m, name: str, p: torch.Tensor) -> torch.Tensor:
    """Parameter in the reference's logical layout (conv weights get the leading group axis)."""
    from ..nn.layers_core import SpatialConvolution

    t = p.detach().float().cpu().contiguous()
    if isinstance(m, SpatialConvolution) and name == "weight":
        g = m.n_group
        return t.reshape(g, m.n_output_plane // g, m.n_input_plane // g, m.kernel_h, m.kernel_w)
    return t


def _module_to_proto(m, ctx: _Ctx) -> pb.BigDLModule:
    from ..nn.module import Graph, Node

    out = pb.BigDLModule()
    if m.hasName():
        out.name = m.getName()
    cls = type(m)
    out.moduleType = _JVM_PREFIX + cls.__name__ if cls.__module__.startswith("bigdl_b200.nn") else \
        f"{cls.__module__}.{cls.__name__}"
    if not cls.__module__.startswith(("bigdl_b200.", "bigdl.")):
        _USER_CLASSES.setdefault(out.moduleType, cls)     # a class this process saved is one it may load back
    out.version = "0.10.0"
    out.train = m.isTraining()
    out.id = id(m) & 0x7FFFFFFF
    out.namePostfix = str(m._uid)
    if isinstance(m, Graph):
        names = {}
        for n in m.exec_order:
            names[id(n)] = n.element.getName()
        for n in m.exec_order:
            sub = out.subModules.add()
            sub.CopyFrom(_module_to_proto(n.element, ctx))
            sub.name = n.element.getName()
            sub.preModules.extend(names[id(p)] for p in n.prev)
            sub.nextModules.extend(names[id(x)] for x in n.next if id(x) in names)
            edges = pb.NameAttrList(name=f"{sub.name}_edges")
            for p, idx in zip(n.prev, n.edge_index):
                av = edges.attr[names[id(p)]]
                av.dataType, av.int32Value = pb.DataType.INT32, -1 if idx is None else idx
            ev = out.attr[f"{sub.name}_edges"]
            ev.dataType = pb.DataType.NAME_ATTR_LIST
            ev.nameAttrListValue.CopyFrom(edges)
        _set_attr(out.attr["inputNames"], [names[id(n)] for n in m.input_nodes], ctx)
        _set_attr(out.attr["outputNames"], [names[id(n)] for n in m.output_nodes], ctx)
        _set_attr(out.attr["stopGradientLayers"], sorted(m._stop_gradient), ctx)
        return out
    for k, v in (m.__dict__.get("_ctor_args") or {}).items():
        if v is None or k in ("init_weight", "init_bias", "init_grad_weight", "init_grad_bias"):
            continue
        if isinstance(v, Node):
            continue
        try:
            _set_attr(out.attr[_camel(k)], v, ctx)
        except TypeError:
            del out.attr[_camel(k)]
    if hasattr(m, "_ser_children"):
        for c in m._ser_children():
            out.subModules.add().CopyFrom(_module_to_proto(c, ctx))
    elif m.modules and not _rebuilds_children(m):
        for c in m.modules:
            out.subModules.add().CopyFrom(_module_to_proto(c, ctx))
    own = list(m._parameters.items())
    if _rebuilds_children(m):
        # composite layers (cells, attention blocks, …) rebuild their children from ctor args: store every leaf tensor
        own = [(f"{i}", p) for i, (_, _, p) in enumerate(m.leaf_parameters())]
        own_mods = [mm for mm, _, _ in m.leaf_parameters()]
        own_names = [nn_ for _, nn_, _ in m.leaf_parameters()]
    out.hasParameters = bool(own)
    for i, (n, p) in enumerate(own):
        if _rebuilds_children(m):
            t = _logical_param(own_mods[i], own_names[i], p)
        else:
            t = _logical_param(m, n, p)
        out.parameters.add().CopyFrom(tensor_to_proto(t, ctx))
    for n, b in m._buffers.items():
        if isinstance(b, torch.Tensor):
            _set_attr(out.attr[n], b, ctx)
    for flag in ("ceil_mode",):
        if flag in m.__dict__ and flag not in (m.__dict__.get("_ctor_args") or {}):
            _set_attr(out.attr[_camel(flag)], bool(m.__dict__[flag]), ctx)
    if getattr(m, "_mkl_int8", None):
        _save_int8(m, out)
    return out


def _rebuilds_children(m) -> bool:
    """True for composite modules whose sub-modules are created by their own constructor (Scale, LSTM, Attention,
    Transformer, FPN, …): they are rebuilt from ctor arguments and only their leaf tensors are stored. Containers the
    user fills with ``add`` (Sequential, Concat, Recurrent, …) store ``subModules`` instead."""
    from ..nn.module import Graph

    if isinstance(m, Graph) or hasattr(m, "_ser_children"):
        return False
    return bool(m.modules) and bool(m.__dict__.get("_children_from_ctor", False))


def _save_int8(m, out):
    s = m._mkl_int8
    out.isMklInt8Enabled = True
    out.inputDimMasks, out.outputDimMasks, out.weightDimMasks = s.get("inputDimMask", 0), s.get("outputDimMask", 0), \
        s.get("weightDimMask", 0)
    for key, field in (("inputScales", out.inputScales), ("outputScales", out.outputScales),
                       ("weightScales", out.weightScales)):
        for arr in s.get(key, []):
            av = field.add()
            av.dataType = pb.DataType.ARRAY_VALUE
            av.arrayValue.size, av.arrayValue.datatype = len(arr), pb.DataType.FLOAT
            av.arrayValue.flt.extend(float(x) for x in arr)


def _resolve_class(module_type: str):
    import importlib

    from ..nn import layer as L

    if module_type.startswith(_JVM_PREFIX):
        name = module_type[len(_JVM_PREFIX):].split(".")[-1]
        alias = {"RnnCell": "RnnCell", "StaticGraph": "Graph", "DynamicGraph": "Graph"}
        name = alias.get(name, name)
        if hasattr(L, name):
            return getattr(L, name)
        for sub in ("keras", "quantized"):
            try:
                mod = importlib.import_module(f"bigdl_b200.nn.{sub}" if sub == "keras" else f"bigdl_b200.{sub}.layers")
                if hasattr(mod, name):
                    return getattr(mod, name)
            except ImportError:
                pass
        raise KeyError(f"no layer registered for module type {module_type}")
    # Anything outside the BigDL namespace must have been registered explicitly (register_module_class): a model file
    # is data, and resolving arbitrary dotted names from it would let a crafted file instantiate any importable class.
    cls = _USER_CLASSES.get(module_type)
    if cls is None:
        modname, _, cname = module_type.rpartition(".")
        if modname.split(".")[0] in ("bigdl_b200", "bigdl"):
            cls = getattr(importlib.import_module(modname), cname, None)
    from ..nn.module import Module as _Module

    if cls is None or not (inspect.isclass(cls) and issubclass(cls, _Module)) and not _is_node_factory(cls):
        raise KeyError(f"module type {module_type!r} is not a registered layer class "
                       "(use serialization.module_io.register_module_class for custom layers)")
    return cls


_USER_CLASSES: dict = {}


def register_module_class(cls, name: str | None = None):
    """Allow ``loadModule`` to rebuild a user-defined layer (the reference's ModuleSerializer.registerModule,
    S/utils/serializer/ModuleSerializer.scala:163-170). ``name`` defaults to the class's dotted path."""
    from ..nn.module import Module as _Module

    if not (inspect.isclass(cls) and issubclass(cls, _Module)):
        raise TypeError("only Module subclasses can be registered for deserialisation")
    _USER_CLASSES[name or f"{cls.__module__}.{cls.__qualname__}"] = cls
    return cls


def _is_node_factory(obj) -> bool:
    """``Input`` is a function that wraps a placeholder module in a graph Node; it lives in bigdl_b200.nn."""
    return callable(obj) and getattr(obj, "__module__", "").startswith("bigdl_b200.nn") and getattr(obj, "__name__", "") == "Input"


def _proto_to_module(p: pb.BigDLModule, storages):
    from ..nn.module import Graph, Node

    cls = _resolve_class(p.moduleType)
    attrs = {k: _get_attr(v, storages) for k, v in p.attr.items() if k != "global_storage"}
    if cls is Graph:
        nodes: dict[str, Node] = {}
        subs = list(p.subModules)
        for s in subs:
            nodes[s.name] = Node(_proto_to_module(s, storages))
        for s in subs:
            edges = p.attr.get(f"{s.name}_edges")
            for pre in s.preModules:
                idx = None
                if edges is not None and pre in edges.nameAttrListValue.attr:
                    v = edges.nameAttrListValue.attr[pre].int32Value
                    idx = None if v < 0 else v
                nodes[pre].add(nodes[s.name], idx)
        g = Graph([nodes[n] for n in attrs.get("inputNames", [])], [nodes[n] for n in attrs.get("outputNames", [])])
        if attrs.get("stopGradientLayers"):
            g.stopGradient(attrs["stopGradientLayers"])
        m = g
    else:
        sig = inspect.signature(cls.__init__)
        kwargs = {}
        for k, v in attrs.items():
            sk = _snake(k)
            if sk in sig.parameters:
                kwargs[sk] = v
            elif k in sig.parameters:
                kwargs[k] = v
        var_pos = [n for n, prm in sig.parameters.items() if prm.kind == inspect.Parameter.VAR_POSITIONAL]
        if var_pos and var_pos[0] in kwargs:
            m = cls(*kwargs.pop(var_pos[0]), **kwargs)
        else:
            m = cls(**kwargs)
        if type(m).__name__ == "Node":          # ``Input()`` builds a graph node around the placeholder module
            m = m.element
        if p.subModules and hasattr(m, "_deser_children"):
            m._deser_children([_proto_to_module(s, storages) for s in p.subModules])
        elif p.subModules and not _rebuilds_children(m):
            m.modules = []
            for s in p.subModules:
                m.add(_proto_to_module(s, storages))
        for k, v in attrs.items():
            if k in m._buffers and isinstance(v, torch.Tensor):
                m._buffers[k] = v.reshape(m._buffers[k].shape).to(m._buffers[k].dtype)
        if "ceilMode" in attrs and hasattr(m, "ceil_mode"):
            m.ceil_mode = bool(attrs["ceilMode"])
        targets = [q for _, _, q in m.leaf_parameters()] if _rebuilds_children(m) else list(m._parameters.values())
        if len(p.parameters) and len(p.parameters) != len(targets):
            raise ValueError(f"module {p.name or p.moduleType}: the file holds {len(p.parameters)} parameter tensors "
                             f"but the rebuilt {type(m).__name__} has {len(targets)}")
        if len(p.parameters):
            with torch.no_grad():
                for q, tp in zip(targets, p.parameters):
                    t = proto_to_tensor(tp, storages)
                    if t is not None:
                        if t.numel() != q.numel():
                            raise ValueError(f"module {p.name or p.moduleType}: parameter of {t.numel()} elements "
                                             f"does not fit {tuple(q.shape)}")
                        q.copy_(t.reshape(q.shape).to(q.dtype))
        elif len(p.parameters) == 0 and (p.HasField("weight") or p.HasField("bias")):   # deprecated fields
            with torch.no_grad():
                for nm in ("weight", "bias"):
                    if p.HasField(nm) and nm in m._parameters:
                        t = proto_to_tensor(getattr(p, nm), storages)
                        if t is not None:
                            m._parameters[nm].copy_(t.reshape(m._parameters[nm].shape))
    if p.name:
        m.setName(p.name)
    (m.training if p.train else m.evaluate)()
    if p.isMklInt8Enabled:
        m._mkl_int8 = {
            "inputDimMask": p.inputDimMasks, "outputDimMask": p.outputDimMasks, "weightDimMask": p.weightDimMasks,
            "inputScales": [list(a.arrayValue.flt) for a in p.inputScales],
            "outputScales": [list(a.arrayValue.flt) for a in p.outputScales],
            "weightScales": [list(a.arrayValue.flt) for a in p.weightScales]}
    m._mark_weights_dirty()
    return m


# ---------------------------------------------------------------- file level
_DT_CODE = {"float_data": 2, "double_data": 3, "int_data": 0, "long_data": 1, "bool_data": 5}
_DT_TORCH = {"float_data": torch.float32, "double_data": torch.float64, "int_data": torch.int32,
             "long_data": torch.int64, "bool_data": torch.bool}


def save_module(module, path: str, weight_path: str | None = None, overwrite: bool = False):
    for f in (path, weight_path):
        if f and os.path.exists(f) and not overwrite:
            raise FileExistsError(f"{f} already exists (pass overWrite=True)")
    from ..serialization.snapshot import _portable_copy

    m = _portable_copy(module)
    ctx = _Ctx()
    root = _module_to_proto(m, ctx)
    root.attr["global_storage"].dataType = pb.DataType.NAME_ATTR_LIST
    gs = root.attr["global_storage"].nameAttrListValue
    gs.name = "global_storage"
    if weight_path is None:
        for tid, t in ctx.tensors.items():
            av = gs.attr[str(tid)]
            av.dataType = pb.DataType.TENSOR
            av.tensorValue.CopyFrom(t)
            av.tensorValue.storage.CopyFrom(ctx.storages[t.storage.id])
    else:
        for tid, t in ctx.tensors.items():
            av = gs.attr[str(tid)]
            av.dataType = pb.DataType.TENSOR
            av.tensorValue.CopyFrom(t)
        _write_weight_file(weight_path, ctx)
    _write_bytes(path, [root.SerializeToString()])


def _write_bytes(path: str, chunks):
    """Local paths are written atomically (tmp + rename); hdfs:// s3a:// ... go through utils.File (ModulePersister
    .saveToFile takes the same path kinds, S/utils/serializer/ModuleLoader.scala:199-243)."""
    if "://" in path and not path.startswith("file://"):
        from ..utils.file import File
        File.saveBytes(b"".join(chunks), path, isOverwrite=True)
        return
    path = path[7:] if path.startswith("file://") else path
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path + ".tmp", "wb") as f:
        for c in chunks:
            f.write(c)
    os.replace(path + ".tmp", path)


def _read_bytes(path: str) -> bytes:
    if "://" in path:
        from ..utils.file import File
        return File.readBytes(path)
    with open(path, "rb") as f:
        return f.read()


def _write_weight_file(path: str, ctx: _Ctx):
    """[int magic][int count]{[int id][int dtype][int size][data]}…[int md5len][md5], big-endian like Java's
    DataOutputStream (ModuleLoader.scala:257-271)."""
    md5 = hashlib.md5()
    chunks = []

    def put(b: bytes):
        md5.update(b)
        chunks.append(b)

    put(struct.pack(">i", MAGIC_NO))
    put(struct.pack(">i", len(ctx.storages)))
    for sid, st in ctx.storages.items():
        for field, code in _DT_CODE.items():
            vals = getattr(st, field)
            if len(vals):
                arr = _storage_flat(st, field, _DT_TORCH[field]).numpy()
                fmt = {"float_data": ">f4", "double_data": ">f8", "int_data": ">i4", "long_data": ">i8",
                       "bool_data": ">i1"}[field]
                put(struct.pack(">iii", sid, code, arr.size))
                put(arr.astype(fmt).tobytes())
                break
        else:
            put(struct.pack(">iii", sid, 2, 0))
    digest = md5.digest()
    _write_bytes(path, chunks + [struct.pack(">i", len(digest)), digest])


def _read_weight_file(path: str) -> dict[int, torch.Tensor]:
    data = _read_bytes(path)
    pos = 0

    def take(n):
        nonlocal pos
        b = data[pos:pos + n]
        pos += n
        return b

    magic, count = struct.unpack(">ii", take(8))
    if magic != MAGIC_NO:
        raise ValueError("Magic number mismatch, not a BigDL weight file")
    out = {}
    for _ in range(count):
        sid, code, size = struct.unpack(">iii", take(12))
        fmt, width, dt = {2: (">f4", 4, torch.float32), 3: (">f8", 8, torch.float64), 0: (">i4", 4, torch.int32),
                          1: (">i8", 8, torch.int64), 5: (">i1", 1, torch.bool)}[code]
        arr = np.frombuffer(take(size * width), dtype=fmt).astype(fmt[1:])
        out[sid] = torch.from_numpy(arr.copy()).to(dt)
    body_end = pos
    (dlen,) = struct.unpack(">i", take(4))
    stored = take(dlen)
    if hashlib.md5(data[:body_end]).digest() != stored:
        raise ValueError("check sum error, please check weight file")
    return out


def load_module(path: str, weight_path: str | None = None):
    root = pb.BigDLModule.FromString(_read_bytes(path))
    storages: dict[int, torch.Tensor] = {}
    if weight_path:
        storages.update(_read_weight_file(weight_path))
    if "global_storage" in root.attr:
        for _, av in root.attr["global_storage"].nameAttrListValue.attr.items():
            proto_to_tensor(av.tensorValue, storages)        # registers storages by id
    return _proto_to_module(root, storages)

