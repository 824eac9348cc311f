"""Reference quantisation arithmetic and descriptors (S/nn/quantized/Quantization.scala:35-180, Desc.scala,
QuantSerializer.scala). ``Quantization`` is the int8 rule of the reference — ``round(v / max(|max|, |min|) * 127)`` per
row, inverse for dequantisation — kept as the oracle the fp8 path is compared with; ``Desc`` carries the geometry a
quantised layer is built from; ``QuantSerializer`` packs a quantised layer's weights (bytes + per-row scales) to and
from a plain dict, the payload the module serializer stores as a QUANT tensor."""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch


class Quantization:
    @staticmethod
    def findMax(x) -> float:
        return float(torch.as_tensor(x).abs().max()) if torch.as_tensor(x).numel() else 0.0

    @staticmethod
    def quantize(value: float, max_: float, min_: float) -> int:
        """One value: round(v / max(|max|, |min|) * 127) as a signed byte (Quantization.scala:35-41)."""
        thr = max(abs(max_), abs(min_))
        return 0 if thr == 0 else int(np.clip(np.rint(value / thr * 127.0), -128, 127))

    @staticmethod
    def dequantize(byte: int, max_: float, min_: float) -> float:
        return float(byte) / 127.0 * max(abs(max_), abs(min_))

    @staticmethod
    def quantize_tensor(x: torch.Tensor):
        """Row-wise (dim 0) symmetric int8: returns (int8 tensor, per-row max, per-row min) like the reference's
        ``quantize(src, buffer, offset)`` over 2-D / 4-D weights (:112-180)."""
        x2 = x.detach().float().reshape(x.shape[0], -1)
        mx, mn = x2.max(dim=1).values, x2.min(dim=1).values
        thr = torch.maximum(mx.abs(), mn.abs()).clamp_min(1e-30)
        q = torch.clamp(torch.round(x2 / thr[:, None] * 127.0), -128, 127).to(torch.int8)
        return q.reshape(x.shape), mx, mn

    @staticmethod
    def dequantize_tensor(q: torch.Tensor, mx: torch.Tensor, mn: torch.Tensor) -> torch.Tensor:
        thr = torch.maximum(mx.abs(), mn.abs())
        return (q.float().reshape(q.shape[0], -1) / 127.0 * thr[:, None]).reshape(q.shape)

    @staticmethod
    def loss(x: torch.Tensor) -> float:
        """Mean absolute round-trip error of the int8 rule on ``x``."""
        q, mx, mn = Quantization.quantize_tensor(x)
        return float((Quantization.dequantize_tensor(q, mx, mn) - x.float()).abs().mean())


class DescType:
    """Which of the four BigQuant descriptors a parameter record describes (Desc.scala:24-28)."""
    ConvData, ConvWeight, LinearData, LinearWeight = "ConvData", "ConvWeight", "LinearData", "LinearWeight"


ConvData, ConvWeight, LinearData, LinearWeight = (DescType.ConvData, DescType.ConvWeight, DescType.LinearData,
                                                  DescType.LinearWeight)


class DescParams:
    """Base of the parameter records: ``getType`` names the descriptor kind, ``array`` lists the values in declaration
    order (Desc.scala:30-40)."""

    _type = None

    def getType(self):
        return self._type

    def array(self):
        return [getattr(self, f) for f in self.__dataclass_fields__]


@dataclass
class ConvDataParams(DescParams):
    _type = DescType.ConvData

    nInputPlane: int
    kernelH: int
    kernelW: int
    strideH: int
    strideW: int
    padH: int
    padW: int
    dilationH: int = 1
    dilationW: int = 1
    batchSize: int = 1
    inputHeight: int = 0
    inputWidth: int = 0


@dataclass
class ConvWeightParams(DescParams):
    _type = DescType.ConvWeight

    nOutputPlane: int
    nInputPlane: int
    kernelH: int
    kernelW: int
    dataFormat: str = "NCHW"


@dataclass
class LinearDataParams(DescParams):
    _type = DescType.LinearData

    batchSize: int
    inputSize: int


@dataclass
class LinearWeightParams(DescParams):
    _type = DescType.LinearWeight

    outputSize: int
    inputSize: int


class Desc:
    """Descriptor factory (Desc.scala): the parameter records above, from a layer."""

    @staticmethod
    def of(layer):
        if hasattr(layer, "kernel_h"):
            return ConvWeightParams(layer.n_output_plane, layer.n_input_plane // getattr(layer, "n_group", 1),
                                    layer.kernel_h, layer.kernel_w)
        return LinearWeightParams(layer.output_size, layer.input_size)


class QuantSerializer:
    @staticmethod
    def serialize(layer) -> dict:
        """Weights of a quantised layer as numpy payloads: fp8 bytes, per-output-channel abs-max, calibrated input abs-max, bias."""
        out = {"type": type(layer).__name__, "desc": Desc.of(layer).__dict__}
        for name in ("weight_q", "weight_absmax", "input_absmax", "bias"):
            t = getattr(layer, name, None)
            if t is None:
                t = layer._buffers.get(name) if hasattr(layer, "_buffers") else None
            if isinstance(t, torch.Tensor):
                out[name] = t.detach().cpu().view(torch.uint8).numpy().copy() if t.dtype not in (
                    torch.float32, torch.int8, torch.uint8) else t.detach().cpu().numpy().copy()
        return out

    @staticmethod
    def deserialize(layer, payload: dict):
        for name in ("weight_q", "weight_absmax", "input_absmax", "bias"):
            if name in payload:
                cur = getattr(layer, name, None)
                if cur is None and hasattr(layer, "_buffers"):
                    cur = layer._buffers.get(name)
                if isinstance(cur, torch.Tensor):
                    src = torch.from_numpy(np.asarray(payload[name]))
                    if cur.dtype not in (torch.float32, torch.int8, torch.uint8):
                        src = src.view(cur.dtype)
                    cur.data.copy_(src.reshape(cur.shape))
        return layer
