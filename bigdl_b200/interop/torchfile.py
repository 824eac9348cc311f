"""Torch7 ``.t7`` binary serialization: tensors, storages, tables and nn modules
(S/utils/TorchFile.scala:79-1102 — ``Module.loadTorch`` / ``saveTorch``, ``File.loadTorch`` / ``saveTorch``).

Wire format (little-endian): every object starts with an int32 type tag — 0 nil, 1 number (double), 2 string
(int32 length + bytes), 3 table, 4 torch object, 5 boolean (int32). Tables and torch objects carry an int32 object index
so shared references are written once. A torch object then has a version string ("V 1"), its class name and a
class-specific payload: tensors write ndim, sizes[], strides[] (int64), a 1-based storage offset and the storage object;
storages write an int64 length followed by the raw elements; nn modules write their field table.
"""
from __future__ import annotations

import os
import struct

import numpy as np
import torch

from ..utils.table import Table

TYPE_NIL, TYPE_NUMBER, TYPE_STRING, TYPE_TABLE, TYPE_TORCH, TYPE_BOOLEAN = 0, 1, 2, 3, 4, 5

_STORAGE_DTYPES = {"torch.FloatStorage": np.float32, "torch.DoubleStorage": np.float64, "torch.LongStorage": np.int64,
                   "torch.IntStorage": np.int32, "torch.ShortStorage": np.int16, "torch.ByteStorage": np.uint8,
                   "torch.CharStorage": np.int8, "torch.CudaStorage": np.float32, "torch.CudaDoubleStorage": np.float64,
                   "torch.CudaLongStorage": np.int64}
_TENSOR_TO_STORAGE = {"torch.FloatTensor": "torch.FloatStorage", "torch.DoubleTensor": "torch.DoubleStorage",
                      "torch.LongTensor": "torch.LongStorage", "torch.IntTensor": "torch.IntStorage",
                      "torch.ShortTensor": "torch.ShortStorage", "torch.ByteTensor": "torch.ByteStorage",
                      "torch.CharTensor": "torch.CharStorage", "torch.CudaTensor": "torch.CudaStorage",
                      "torch.CudaDoubleTensor": "torch.CudaDoubleStorage", "torch.CudaLongTensor": "torch.CudaLongStorage"}


class TorchObject:
    """A torch class this reader has no conversion for: class name + raw field table."""

    def __init__(self, cls: str, fields):
        self.torch_class, self.fields = cls, fields

    def __repr__(self):
        return f"TorchObject({self.torch_class})"


# ------------------------------------------------------------------------------------------------ reader
class _Reader:
    def __init__(self, buf: bytes):
        self.buf, self.pos, self.objects = buf, 0, {}

    def _i32(self):
        v = struct.unpack_from("<i", self.buf, self.pos)[0]
        self.pos += 4
        return v

    def _i64(self):
        v = struct.unpack_from("<q", self.buf, self.pos)[0]
        self.pos += 8
        return v

    def _str(self):
        n = self._i32()
        s = self.buf[self.pos:self.pos + n].decode("latin-1")
        self.pos += n
        return s

    def read(self):
        tag = self._i32()
        if tag == TYPE_NIL:
            return None
        if tag == TYPE_NUMBER:
            v = struct.unpack_from("<d", self.buf, self.pos)[0]
            self.pos += 8
            return v
        if tag == TYPE_STRING:
            return self._str()
        if tag == TYPE_BOOLEAN:
            return self._i32() == 1
        if tag == TYPE_TABLE:
            idx = self._i32()
            if idx in self.objects:
                return self.objects[idx]
            t = Table()
            self.objects[idx] = t
            for _ in range(self._i32()):
                k = self.read()
                v = self.read()
                if isinstance(k, float) and k == int(k):
                    k = int(k)
                t[k] = v
            return t
        if tag == TYPE_TORCH:
            idx = self._i32()
            if idx in self.objects:
                return self.objects[idx]
            version = self._str()
            cls = self._str() if version.startswith("V ") else version
            obj = self._read_torch(cls, idx)
            self.objects[idx] = obj
            return obj
        raise ValueError(f"unsupported t7 type tag {tag} at byte {self.pos - 4}")

    def _read_torch(self, cls: str, idx: int):
        if cls in _STORAGE_DTYPES:
            n = self._i64()
            dt = np.dtype(_STORAGE_DTYPES[cls])
            arr = np.frombuffer(self.buf, dtype=dt, count=n, offset=self.pos).copy()
            self.pos += n * dt.itemsize
            return torch.from_numpy(arr)
        if cls in _TENSOR_TO_STORAGE:
            nd = self._i32()
            sizes = [self._i64() for _ in range(nd)]
            strides = [self._i64() for _ in range(nd)]
            off = self._i64() - 1
            storage = self.read()
            if storage is None or nd == 0:
                return torch.empty(0, dtype=torch.from_numpy(np.zeros(0, _STORAGE_DTYPES[_TENSOR_TO_STORAGE[cls]])).dtype)
            return torch.as_strided(storage, sizes, strides, off)
        fields = self.read()
        return convert_module(cls, fields)


def load(file_name: str):
    with open(file_name, "rb") as f:
        return _Reader(f.read()).read()


# ------------------------------------------------------------------------------------------------ module conversion
def _get(t, k, default=None):
    try:
        v = t[k]
        return default if v is None else v
    except (KeyError, IndexError):
        return default


def _children(fields):
    mods = _get(fields, "modules")
    if mods is None:
        return []
    return [mods[i] for i in sorted(k for k in mods.keys() if isinstance(k, int))]


def convert_module(cls: str, f):
    """torch class name + field table → framework module (TorchFile.scala:137-260). Unknown classes come back as
    ``TorchObject`` so tables of mixed content still load."""
    from ..nn import layer as L

    name = cls.replace("cudnn.", "nn.")
    if not name.startswith("nn."):
        return TorchObject(cls, f)
    short = name[3:]
    i = lambda k, d=0: int(_get(f, k, d))

    def set_wb(m, transpose4d=False):
        w, b = _get(f, "weight"), _get(f, "bias")
        with torch.no_grad():
            if w is not None and "weight" in m._parameters:
                m.weight.copy_(w.float().reshape(m.weight.shape))
            if b is not None and "bias" in m._parameters:
                m.bias.copy_(b.float().reshape(m.bias.shape))
        m._mark_weights_dirty()
        return m

    if short == "Sequential":
        s = L.Sequential()
        for c in _children(f):
            s.add(c)
        return s
    if short == "Concat":
        s = L.Concat(i("dimension", 2))
        for c in _children(f):
            s.add(c)
        return s
    if short == "ConcatTable":
        return L.ConcatTable(*_children(f))
    if short == "Linear":
        w = f["weight"]
        return set_wb(L.Linear(w.shape[1], w.shape[0], with_bias=_get(f, "bias") is not None))
    if short in ("SpatialConvolution", "SpatialConvolutionMM"):
        m = L.SpatialConvolution(i("nInputPlane"), i("nOutputPlane"), i("kW"), i("kH"), i("dW", 1), i("dH", 1),
                                 i("padW"), i("padH"), i("groups", 1), with_bias=_get(f, "bias") is not None)
        return set_wb(m)
    if short == "SpatialMaxPooling":
        m = L.SpatialMaxPooling(i("kW"), i("kH"), i("dW"), i("dH"), i("padW"), i("padH"), bool(_get(f, "ceil_mode", False)))
        return m
    if short == "SpatialAveragePooling":
        return L.SpatialAveragePooling(i("kW"), i("kH"), i("dW", 1), i("dH", 1), i("padW"), i("padH"),
                                       ceil_mode=bool(_get(f, "ceil_mode", False)),
                                       count_include_pad=bool(_get(f, "count_include_pad", True)),
                                       divide=bool(_get(f, "divide", True)))
    if short in ("BatchNormalization", "SpatialBatchNormalization"):
        rm = f["running_mean"]
        C = getattr(L, short)
        m = C(rm.numel(), float(_get(f, "eps", 1e-5)), float(_get(f, "momentum", 0.1)), _get(f, "affine", True) is not False
              and _get(f, "weight") is not None)
        m._buffers["runningMean"].copy_(rm.float())
        rv = _get(f, "running_var")
        if rv is not None:
            m._buffers["runningVar"].copy_(rv.float())
        return set_wb(m)
    if short == "SpatialCrossMapLRN":
        return L.SpatialCrossMapLRN(i("size", 5), float(_get(f, "alpha", 1e-4)), float(_get(f, "beta", 0.75)),
                                    float(_get(f, "k", 1.0)))
    if short == "SpatialZeroPadding":
        return L.SpatialZeroPadding(i("pad_l"), i("pad_r"), i("pad_t"), i("pad_b"))
    if short == "ReLU":
        return L.ReLU(bool(_get(f, "inplace", False)))
    if short == "Threshold":
        return L.Threshold(float(_get(f, "threshold", 1e-6)), float(_get(f, "val", 0.0)), bool(_get(f, "inplace", False)))
    if short == "LeakyReLU":
        return L.LeakyReLU(float(_get(f, "negval", 0.01)), bool(_get(f, "inplace", False)))
    if short == "Dropout":
        return L.Dropout(float(_get(f, "p", 0.5)), bool(_get(f, "inplace", False)), not bool(_get(f, "v2", True)) or True)
    if short == "CAddTable":
        return L.CAddTable(bool(_get(f, "inplace", False)))
    if short == "Reshape":
        return L.Reshape([int(v) for v in f["size"].tolist()])
    if short == "View":
        m = L.View(*[int(v) for v in f["size"].tolist()])
        n = _get(f, "numInputDims")
        if n is not None:
            m.setNumInputDims(int(n))
        return m
    if short == "SpatialConvolutionMap":
        conn = f["connTable"].float()
        m = L.SpatialConvolutionMap(conn, i("kW"), i("kH"), i("dW", 1), i("dH", 1), i("padW"), i("padH"))
        return set_wb(m)
    cls_obj = getattr(L, short, None)
    if cls_obj is not None:
        try:
            return cls_obj()          # parameter-free modules are built reflectively (TorchFile.scala:170-188)
        except TypeError:
            pass
    raise ValueError(f"unsupported module {cls}")


# ------------------------------------------------------------------------------------------------ writer
class _Writer:
    def __init__(self):
        self.out = bytearray()
        self.index = 0
        self.seen: dict[int, int] = {}

    def _i32(self, v): self.out += struct.pack("<i", int(v))
    def _i64(self, v): self.out += struct.pack("<q", int(v))

    def _str(self, s: str):
        b = s.encode("latin-1")
        self._i32(len(b))
        self.out += b

    def _next(self):
        self.index += 1
        return self.index

    def write(self, obj):
        from ..nn.module import Module

        if obj is None:
            self._i32(TYPE_NIL)
        elif isinstance(obj, bool):
            self._i32(TYPE_BOOLEAN)
            self._i32(1 if obj else 0)
        elif isinstance(obj, (int, float)):
            self._i32(TYPE_NUMBER)
            self.out += struct.pack("<d", float(obj))
        elif isinstance(obj, str):
            self._i32(TYPE_STRING)
            self._str(obj)
        elif isinstance(obj, torch.Tensor):
            self._tensor(obj)
        elif isinstance(obj, np.ndarray):
            self._tensor(torch.from_numpy(np.ascontiguousarray(obj)))
        elif isinstance(obj, Module):
            self._module(obj)
        elif isinstance(obj, (Table, dict, list, tuple)):
            self._table(obj)
        else:
            raise TypeError(f"cannot serialise {type(obj).__name__} to t7")

    def _table(self, t):
        self._i32(TYPE_TABLE)
        self._i32(self._next())
        items = list(t.items()) if isinstance(t, (Table, dict)) else [(i + 1, v) for i, v in enumerate(t)]
        self._i32(len(items))
        for k, v in items:
            self.write(k)
            self.write(v)

    def _tensor(self, t: torch.Tensor):
        t = t.detach().cpu()
        kinds = {torch.float32: "Float", torch.float64: "Double", torch.int64: "Long", torch.int32: "Int",
                 torch.int16: "Short", torch.uint8: "Byte", torch.int8: "Char"}
        if t.dtype not in kinds:
            t = t.float()
        kind = kinds[t.dtype]
        t = t.contiguous()
        self._i32(TYPE_TORCH)
        self._i32(self._next())
        self._str("V 1")
        self._str(f"torch.{kind}Tensor")
        self._i32(t.dim())
        for s in t.shape:
            self._i64(s)
        for s in t.stride():
            self._i64(s)
        self._i64(1)
        if t.numel() == 0:
            self._i32(TYPE_NIL)
            return
        self._i32(TYPE_TORCH)
        self._i32(self._next())
        self._str("V 1")
        self._str(f"torch.{kind}Storage")
        self._i64(t.numel())
        self.out += t.numpy().tobytes()

    def _module(self, m):
        from ..nn import layer as L

        f = Table()
        f["_type"] = "torch.FloatTensor"
        f["train"] = m.isTraining()
        name = type(m).__name__
        if isinstance(m, L.Linear):
            f["weight"], f["gradWeight"] = m.weight.detach(), torch.zeros_like(m.weight)
            if m.with_bias:
                f["bias"], f["gradBias"] = m.bias.detach(), torch.zeros_like(m.bias)
        elif type(m) is L.SpatialConvolution:
            if m.n_group != 1:
                raise ValueError("nGroup is not supported in torch")
            for k, v in (("nInputPlane", m.n_input_plane), ("nOutputPlane", m.n_output_plane), ("kW", m.kernel_w),
                         ("kH", m.kernel_h), ("dW", m.stride_w), ("dH", m.stride_h), ("padW", m.pad_w), ("padH", m.pad_h)):
                f[k] = v
            f["weight"] = m.weight.detach().contiguous()
            f["gradWeight"] = torch.zeros_like(f["weight"])
            if m.with_bias:
                f["bias"], f["gradBias"] = m.bias.detach(), torch.zeros_like(m.bias)
        elif isinstance(m, L.SpatialMaxPooling):
            for k, v in (("kW", m.kw), ("kH", m.kh), ("dW", m.dw), ("dH", m.dh), ("padW", m.pad_w), ("padH", m.pad_h),
                         ("ceil_mode", bool(m.ceil_mode))):
                f[k] = v
        elif type(m) is L.ReLU:
            f["inplace"], f["threshold"], f["val"] = bool(getattr(m, "inplace", False)), 0.0, 0.0
        elif isinstance(m, L.Threshold):
            f["threshold"], f["val"], f["inplace"] = float(m.th), float(m.v), False
        elif isinstance(m, L.Dropout):
            f["p"], f["inplace"], f["v2"] = float(m.p), bool(m.inplace), True
        elif isinstance(m, L.View):
            f["size"] = torch.tensor(list(m.sizes), dtype=torch.int64)
            f["numElements"] = float(np.prod(list(m.sizes)))
            if m.num_input_dims:
                f["numInputDims"] = m.num_input_dims
        elif isinstance(m, L.SpatialCrossMapLRN):
            f["size"], f["alpha"], f["beta"], f["k"] = m.size, float(m.alpha), float(m.beta), float(m.k)
        elif isinstance(m, L.LogSoftMax):
            pass
        elif isinstance(m, (L.Sequential, L.Concat)):
            f["modules"] = Table(*m.modules)
            if isinstance(m, L.Concat):
                f["dimension"] = m.dimension
        else:
            raise ValueError(f"saveTorch: unsupported module {name}")
        self._i32(TYPE_TORCH)
        self._i32(self._next())
        self._str("V 1")
        self._str(f"nn.{name}")
        self._table(f)


def save(obj, file_name: str, overWrite: bool = False):
    if os.path.exists(file_name) and not overWrite:
        raise FileExistsError(file_name)
    w = _Writer()
    w.write(obj)
    with open(file_name, "wb") as f:
        f.write(bytes(w.out))


def save_torch_module(module, path: str, overWrite: bool = False):
    save(module, path, overWrite)


def load_torch_module(path: str):
    m = load(path)
    from ..nn.module import Module

    if not isinstance(m, Module):
        raise ValueError(f"{path} does not contain an nn module (got {type(m).__name__})")
    return m


class TorchFile:
    """``TorchFile.load(path)`` / ``TorchFile.save(obj, path, type, overWrite)`` (S/utils/TorchFile.scala:70-130)."""
    load = staticmethod(load)

    @staticmethod
    def save(source, fileName: str, objectType=None, overWrite: bool = False):
        return save(source, fileName, overWrite)

    loadModule = staticmethod(load_torch_module)
    saveModule = staticmethod(save_torch_module)
