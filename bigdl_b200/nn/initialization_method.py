"""Weight initialisers (reference: S/nn/abstractnn/InitializationMethod.scala:1-362; Python: P/nn/initialization_method.py).

``VariableFormat`` tells an initialiser how to read fan-in / fan-out from a weight shape.
"""
from __future__ import annotations

import math

import torch

from ..utils.random_generator import RNG


class VariableFormat:
    DEFAULT = "Default"
    ONE_D = "ONE_D"
    IN_OUT = "IN_OUT"
    OUT_IN = "OUT_IN"
    IN_OUT_KW_KH = "IN_OUT_KW_KH"
    OUT_IN_KW_KH = "OUT_IN_KW_KH"
    GP_OUT_IN_KW_KH = "GP_OUT_IN_KW_KH"
    GP_IN_OUT_KW_KH = "GP_IN_OUT_KW_KH"
    OUT_IN_KT_KH_KW = "OUT_IN_KT_KH_KW"
    GP_KH_KW_IN_OUT = "GP_KH_KW_IN_OUT"

    @staticmethod
    def fans(shape, fmt) -> tuple[int, int]:
        s = list(shape)
        if fmt == VariableFormat.ONE_D or len(s) == 1:
            return s[0], 1
        if fmt == VariableFormat.IN_OUT:
            return s[0], s[1]
        if fmt == VariableFormat.OUT_IN or (fmt == VariableFormat.DEFAULT and len(s) == 2):
            return s[1], s[0]
        if fmt == VariableFormat.IN_OUT_KW_KH:
            rf = s[2] * s[3]
            return s[0] * rf, s[1] * rf
        if fmt == VariableFormat.OUT_IN_KW_KH or (fmt == VariableFormat.DEFAULT and len(s) == 4):
            rf = s[2] * s[3]
            return s[1] * rf, s[0] * rf
        if fmt == VariableFormat.GP_OUT_IN_KW_KH or (fmt == VariableFormat.DEFAULT and len(s) == 5):
            rf = s[3] * s[4]
            return s[2] * rf, s[1] * rf
        if fmt == VariableFormat.GP_IN_OUT_KW_KH:
            rf = s[3] * s[4]
            return s[1] * rf, s[2] * rf
        if fmt == VariableFormat.OUT_IN_KT_KH_KW:
            rf = s[2] * s[3] * s[4]
            return s[1] * rf, s[0] * rf
        if fmt == VariableFormat.GP_KH_KW_IN_OUT:
            rf = s[1] * s[2]
            return s[3] * rf, s[4] * rf
        n = 1
        for d in s[1:]:
            n *= d
        return n, s[0]


class InitializationMethod:
    def init(self, variable: torch.Tensor, dataFormat=VariableFormat.DEFAULT) -> torch.Tensor:  # pragma: no cover
        raise NotImplementedError

    def __call__(self, variable, dataFormat=VariableFormat.DEFAULT):
        with torch.no_grad():
            return self.init(variable, dataFormat)


class RandomUniform(InitializationMethod):
    """U(lower, upper); with no bounds given, U(-1/sqrt(fanIn), 1/sqrt(fanIn)) — the reference's default for
    Linear/Conv weights *and* biases (S/nn/Linear.scala:63-68)."""

    def __init__(self, lower: float | None = None, upper: float | None = None):
        self.lower, self.upper = lower, upper

    def init(self, variable, dataFormat=VariableFormat.DEFAULT):
        if self.lower is None:
            fan_in, _ = VariableFormat.fans(variable.shape, dataFormat)
            stdv = 1.0 / math.sqrt(max(fan_in, 1))
            lo, hi = -stdv, stdv
        else:
            lo, hi = self.lower, self.upper
        return RNG.fill_uniform(variable, lo, hi)


class RandomNormal(InitializationMethod):
    def __init__(self, mean: float = 0.0, stdv: float = 1.0):
        self.mean, self.stdv = mean, stdv

    def init(self, variable, dataFormat=VariableFormat.DEFAULT):
        return RNG.fill_normal(variable, self.mean, self.stdv)


class Zeros(InitializationMethod):
    def init(self, variable, dataFormat=VariableFormat.DEFAULT):
        return variable.zero_()


class Ones(InitializationMethod):
    def init(self, variable, dataFormat=VariableFormat.DEFAULT):
        return variable.fill_(1.0)


class ConstInitMethod(InitializationMethod):
    def __init__(self, value: float):
        self.value = value

    def init(self, variable, dataFormat=VariableFormat.DEFAULT):
        return variable.fill_(self.value)


class Xavier(InitializationMethod):
    """U(±sqrt(6/(fanIn+fanOut))) (InitializationMethod.scala:280-299)."""

    def init(self, variable, dataFormat=VariableFormat.DEFAULT):
        fan_in, fan_out = VariableFormat.fans(variable.shape, dataFormat)
        stdv = math.sqrt(6.0 / (fan_in + fan_out))
        return RNG.fill_uniform(variable, -stdv, stdv)


class MsraFiller(InitializationMethod):
    """N(0, sqrt(2/n)), n = fanIn (varianceNormAverage=False → fanOut as in the reference: ``n = fanOut`` when
    ``varianceNormAverage`` is false uses fan-in; average uses (fanIn+fanOut)/2). (InitializationMethod.scala:301-330)"""

    def __init__(self, varianceNormAverage: bool = True):
        self.varianceNormAverage = varianceNormAverage

    def init(self, variable, dataFormat=VariableFormat.DEFAULT):
        fan_in, fan_out = VariableFormat.fans(variable.shape, dataFormat)
        n = (fan_in + fan_out) / 2.0 if self.varianceNormAverage else fan_in
        return RNG.fill_normal(variable, 0.0, math.sqrt(2.0 / n))


class BilinearFiller(InitializationMethod):
    """Bilinear up-sampling kernel for SpatialFullConvolution (InitializationMethod.scala:332-362)."""

    def init(self, variable, dataFormat=VariableFormat.DEFAULT):
        if variable.dim() not in (4, 5):
            raise ValueError("BilinearFiller needs a 4-D or 5-D weight")
        kh, kw = variable.shape[-2], variable.shape[-1]
        if kh != kw:
            raise ValueError("Kernel must be square")
        f = math.ceil(kw / 2.0)
        c = (2 * f - 1 - f % 2) / (2.0 * f)
        xs = torch.arange(kw, dtype=torch.float32)
        ys = torch.arange(kh, dtype=torch.float32)
        k2 = (1 - (ys / f - c).abs())[:, None] * (1 - (xs / f - c).abs())[None, :]
        variable.copy_(k2.expand_as(variable))
        return variable
