"""Quantised weight container (S/tensor/QuantizedTensor.scala:28-305): the int8 bytes of a 2-D (rows = output
channels) view of a float tensor plus per-row ``maxOfRow / minOfRow / sumOfRow``. The reference additionally holds a
pointer to a BigQuant native descriptor that ``release()`` frees; here the payload is an ordinary tensor, and
``to_e4m3`` produces the fp8 form the sm_100a kernels consume."""
from __future__ import annotations

import torch


class QuantizedTensor:
    def __init__(self, q: torch.Tensor, max_of_row: torch.Tensor, min_of_row: torch.Tensor, sum_of_row: torch.Tensor,
                 shape):
        self.q, self.maxOfRow, self.minOfRow, self.sumOfRow, self._shape = q, max_of_row, min_of_row, sum_of_row, list(shape)
        self._released = False

    @staticmethod
    def from_float(t: torch.Tensor) -> "QuantizedTensor":
        """Symmetric per-row int8: q = round(x / max(|max|,|min|) * 127) (S/nn/quantized/Quantization.scala:35-62)."""
        shape = t.shape
        m = t.detach().float().reshape(shape[0], -1)
        mx, mn = m.amax(1), m.amin(1)
        am = torch.maximum(mx.abs(), mn.abs())
        scale = torch.where(am > 0, 127.0 / am, torch.zeros_like(am))
        q = torch.round(m * scale[:, None]).clamp(-127, 127).to(torch.int8)
        return QuantizedTensor(q, mx, mn, q.float().sum(1), shape)

    def size(self, d: int | None = None):
        return self._shape if d is None else self._shape[d - 1]

    def nElement(self) -> int:
        return self.q.numel()

    def getStorage(self) -> torch.Tensor:
        return self.q.reshape(-1)

    def dequantize(self) -> torch.Tensor:
        am = torch.maximum(self.maxOfRow.abs(), self.minOfRow.abs())
        return (self.q.float() * (am / 127.0)[:, None]).reshape(self._shape)

    def to_e4m3(self):
        """(uint8 e4m3 bytes [rows, cols padded to 16], per-row absmax) for the fp8 tcgen05 GEMM."""
        from ..ops import quant as Q

        w = self.dequantize().reshape(self._shape[0], -1)
        am = Q.row_absmax(w)
        return Q.quantize_e4m3(w, am, per_row=True), am

    def release(self):
        self._released = True
        return self

    def __repr__(self):
        return f"QuantizedTensor(shape={self._shape}, int8)"
