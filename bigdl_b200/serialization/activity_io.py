"""Activity ⇄ bytes for the serving API (S/optim/PredictionService.scala:184-280): an Activity is encoded as one
``AttrValue`` of bigdl.proto — a single tensor as ``tensorValue``; a Table as an ``arrayValue`` of tensors laid out
``[isKeyPrimitive, keys…, values…]`` where primitive keys are scalar tensors. Errors travel back to the caller as a scalar
string tensor, which the reference does instead of throwing across the service boundary."""
from __future__ import annotations

import torch

from ..utils.table import Table
from . import bigdl_pb as pb
from .module_io import proto_to_tensor, tensor_to_proto


def _string_tensor(msg: str) -> pb.BigDLTensor:
    t = pb.BigDLTensor()
    t.datatype, t.dimension, t.nElements, t.isScalar, t.offset = pb.DataType.STRING, 0, 1, True, 1
    t.storage.datatype = pb.DataType.STRING
    t.storage.string_data.append(msg)
    return t


def serialize_activity(activity, error: str | None = None) -> bytes:
    av = pb.AttrValue()
    if error is not None:
        av.dataType = pb.DataType.TENSOR
        av.tensorValue.CopyFrom(_string_tensor(error))
        return av.SerializeToString()
    if isinstance(activity, torch.Tensor):
        av.dataType = pb.DataType.TENSOR
        av.tensorValue.CopyFrom(tensor_to_proto(activity, None, inline=True))
        return av.SerializeToString()
    if isinstance(activity, Table):
        items = list(activity.items())
        for k, v in items:
            if not isinstance(v, torch.Tensor):
                raise TypeError(f"Unsupported Table value for key {k!r}: only Table[primitive | Tensor, Tensor] travels "
                                f"over the service boundary, got {type(v).__name__}")
        primitive = not (items and isinstance(items[0][0], torch.Tensor))
        keys = [k if isinstance(k, torch.Tensor) else torch.tensor(k) for k, _ in items]
        tensors = [torch.tensor(primitive)] + keys + [v for _, v in items]
        av.dataType = pb.DataType.ARRAY_VALUE
        av.arrayValue.size, av.arrayValue.datatype = len(tensors), pb.DataType.TENSOR
        for t in tensors:
            av.arrayValue.tensor.append(tensor_to_proto(t, None, inline=True))
        return av.SerializeToString()
    raise TypeError(f"Unsupported Activity type {type(activity).__name__}")


def _tensor_or_string(p: pb.BigDLTensor):
    if p.datatype == pb.DataType.STRING:
        vals = list(p.storage.string_data)
        return vals[0] if p.isScalar or len(vals) == 1 else vals
    return proto_to_tensor(p, {})


def deserialize_activity(data: bytes):
    av = pb.AttrValue()
    av.ParseFromString(data)
    if av.dataType == pb.DataType.TENSOR:
        return _tensor_or_string(av.tensorValue)
    if av.dataType == pb.DataType.ARRAY_VALUE:
        tensors = [_tensor_or_string(t) for t in av.arrayValue.tensor]
        n = (len(tensors) - 1) // 2
        primitive = bool(tensors[0])
        keys, values = tensors[1:n + 1], tensors[n + 1:]
        out = Table()
        for k, v in zip(keys, values):
            if primitive:
                k = k.item()
                k = int(k) if float(k).is_integer() else k
            out[k] = v
        return out
    raise ValueError(f"Unsupported DataType({av.dataType})!")
