"""ONNX-specific layers (S/nn/onnx/{Gemm,Reshape,Shape}.scala)."""
from __future__ import annotations

import torch

from .module import Module
from .ops import Operation, _ins


class Gemm(Module):
    """Y = alpha * A' * B' + beta * C with optional transposes (Gemm.scala:30-90). Inputs: Table(A, B, C) or, when
    ``matrix_b`` / ``matrix_c`` were given at construction, just A (weights held as parameters)."""

    def __init__(self, alpha: float = 1.0, beta: float = 1.0, trans_a: bool = False, trans_b: bool = False,
                 matrix_b=None, matrix_c=None):
        super().__init__()
        self.alpha, self.beta, self.trans_a, self.trans_b = alpha, beta, trans_a, trans_b
        if matrix_b is not None:
            self.register_parameter("weight", torch.as_tensor(matrix_b, dtype=torch.float32).clone())
        if matrix_c is not None:
            self.register_parameter("bias", torch.as_tensor(matrix_c, dtype=torch.float32).clone())

    def updateOutput(self, input):
        items = _ins(input) if not isinstance(input, torch.Tensor) else [input]
        a = items[0]
        b = items[1] if len(items) > 1 else self._parameters["weight"]
        c = items[2] if len(items) > 2 else self._parameters.get("bias")
        a = a.t() if self.trans_a else a
        b = b.t() if self.trans_b else b
        y = self.alpha * (a @ b.to(a.dtype))
        if c is not None:
            y = y + self.beta * c.to(a.dtype)
        return y


class Reshape(Module):
    """Reshape to a shape given at construction or as a second input; 0 copies the input dim, −1 is inferred."""

    def __init__(self, shape=None):
        super().__init__()
        self.shape = None if shape is None else [int(s) for s in shape]

    def updateOutput(self, input):
        if isinstance(input, torch.Tensor):
            x, shape = input, self.shape
        else:
            x, s = _ins(input, 2)
            shape = [int(v) for v in torch.as_tensor(s).reshape(-1).tolist()]
        shape = [x.shape[i] if d == 0 else d for i, d in enumerate(shape)]
        return x.reshape(shape)


class Shape(Operation):
    def updateOutput(self, input):
        return torch.tensor(list(input.shape), dtype=torch.int64)
