"""Element-wise activations, noise layers and in-graph penalties.

Reference files (S/nn/): ReLU6, Threshold, BinaryThreshold, LeakyReLU, PReLU, RReLU, SReLU, ELU, Sigmoid,
HardSigmoid, Tanh, HardTanh, TanhShrink, LogSigmoid, SoftPlus, SoftSign, SoftShrink, HardShrink, Clamp, Abs, Exp, Log,
Sqrt, Square, Power, Negative, AddConstant, MulConstant, GradientReversal, SpatialDropout1D/2D/3D, GaussianDropout,
GaussianNoise, GaussianSampler, L1Penalty, NegativeEntropyPenalty, ActivityRegularization.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from ..utils.random_generator import RNG
from ..utils.table import Table, to_table
from .initialization_method import ConstInitMethod, RandomUniform, VariableFormat, Xavier, Zeros
from .module import Module


def _ew_native(x, op):
    """bf16 CUDA tensors with a dense layout go through the vectorised elementwise kernel."""
    from .layers_core import native

    if native(x) and x.dtype == torch.bfloat16 and x.numel() % 8 == 0 and (
            x.is_contiguous() or (x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last))):
        return True
    return False


class ReLU6(Module):
    def __init__(self, inplace: bool = False):
        super().__init__()

    def updateOutput(self, input):
        return F.relu6(input)


class Threshold(Module):
    def __init__(self, th: float = 1e-6, v: float = 0.0, ip: bool = False):
        super().__init__()
        self.th, self.v = th, v

    def updateOutput(self, input):
        return torch.where(input > self.th, input, torch.full_like(input, self.v))


class BinaryThreshold(Module):
    def __init__(self, th: float = 1e-6, ip: bool = False):
        super().__init__()
        self.th = th

    def updateOutput(self, input):
        return (input > self.th).to(input.dtype)


class LeakyReLU(Module):
    def __init__(self, negval: float = 0.01, inplace: bool = False):
        super().__init__()
        self.negval = negval

    def updateOutput(self, input):
        return F.leaky_relu(input, self.negval)


class PReLU(Module):
    """Learned leak, one per channel (``n_output_plane`` > 0) or shared (0); init 0.25 (S/nn/PReLU.scala)."""

    def __init__(self, n_output_plane: int = 0):
        super().__init__()
        self.n_output_plane = n_output_plane
        self.register_parameter("weight", torch.full((max(n_output_plane, 1),), 0.25))

    def reset(self):
        self._init_methods.get("weight", ConstInitMethod(0.25))(self.weight.data, VariableFormat.ONE_D)
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        w = self.weight.to(input.dtype)
        if self.n_output_plane == 0:
            return torch.where(input > 0, input, input * w)
        cdim = 0 if input.dim() in (1, 3) else 1
        shape = [1] * input.dim()
        shape[cdim] = -1
        return torch.where(input > 0, input, input * w.reshape(shape))


class RReLU(Module):
    def __init__(self, lower: float = 1.0 / 8, upper: float = 1.0 / 3, inplace: bool = False):
        super().__init__()
        self.lower, self.upper = lower, upper

    def updateOutput(self, input):
        if self.train:
            a = torch.empty(input.shape, dtype=torch.float32)
            RNG.fill_uniform(a, self.lower, self.upper)
            a = a.to(device=input.device, dtype=input.dtype)
            return torch.where(input >= 0, input, input * a)
        return F.leaky_relu(input, (self.lower + self.upper) / 2)


class SReLU(Module):
    """S-shaped ReLU with learnable thresholds/slopes (t_left, a_left, t_right, a_right) (S/nn/SReLU.scala)."""

    def __init__(self, shape, share_axes=None):
        super().__init__()
        self.shape = list(shape)
        self.share_axes = share_axes
        pshape = list(self.shape)
        if share_axes is not None:
            for ax in share_axes:
                pshape[ax - 1] = 1
        self.register_parameter("tLeft", torch.zeros(pshape))
        self.register_parameter("aLeft", torch.zeros(pshape))
        self.register_parameter("tRight", torch.zeros(pshape))
        self.register_parameter("aRight", torch.ones(pshape))
        self.reset()

    def reset(self):
        Zeros()(self.tLeft.data)
        Xavier()(self.aLeft.data, VariableFormat.IN_OUT if self.aLeft.dim() == 2 else VariableFormat.ONE_D)
        Xavier()(self.tRight.data, VariableFormat.IN_OUT if self.tRight.dim() == 2 else VariableFormat.ONE_D)
        self.aRight.data.fill_(1.0)
        self.zeroGradParameters()
        return self

    def setInitMethod(self, weightInitMethod=None, biasInitMethod=None, tLeftInit=None, aLeftInit=None,
                      tRightInit=None, aRightInit=None):
        for p, init in ((self.tLeft, tLeftInit), (self.aLeft, aLeftInit), (self.tRight, tRightInit),
                        (self.aRight, aRightInit)):
            if init is not None:
                init(p.data, VariableFormat.ONE_D)
        return self

    def updateOutput(self, input):
        tl, al, tr, ar = (p.to(input.dtype) for p in (self.tLeft, self.aLeft, self.tRight, self.aRight))
        tr_abs = tl + tr.abs()
        y_left = tl + al * (input - tl)
        y_right = tr_abs + ar * (input - tr_abs)
        return torch.where(input >= tr_abs, y_right, torch.where(input <= tl, y_left, input))


class ELU(Module):
    def __init__(self, alpha: float = 1.0, inplace: bool = False):
        super().__init__()
        self.alpha = alpha

    def updateOutput(self, input):
        return F.elu(input, self.alpha)


class Sigmoid(Module):
    def updateOutput(self, input):
        return torch.sigmoid(input)


class HardSigmoid(Module):
    def updateOutput(self, input):
        return torch.clamp(0.2 * input + 0.5, 0.0, 1.0)


class Tanh(Module):
    def updateOutput(self, input):
        return torch.tanh(input)


class HardTanh(Module):
    def __init__(self, min_value: float = -1.0, max_value: float = 1.0, inplace: bool = False):
        super().__init__()
        self.min_value, self.max_value = min_value, max_value

    def updateOutput(self, input):
        return F.hardtanh(input, self.min_value, self.max_value)


class Clamp(HardTanh):
    def __init__(self, min: float, max: float):
        super().__init__(min, max)


class TanhShrink(Module):
    def updateOutput(self, input):
        return input - torch.tanh(input)


class LogSigmoid(Module):
    def updateOutput(self, input):
        return F.logsigmoid(input)


class SoftPlus(Module):
    def __init__(self, beta: float = 1.0):
        super().__init__()
        self.beta = beta

    def updateOutput(self, input):
        return F.softplus(input, self.beta, threshold=20.0)


class SoftSign(Module):
    def updateOutput(self, input):
        return input / (1 + input.abs())


class SoftShrink(Module):
    def __init__(self, the_lambda: float = 0.5):
        super().__init__()
        self.the_lambda = the_lambda

    def updateOutput(self, input):
        return F.softshrink(input, self.the_lambda)


class HardShrink(Module):
    def __init__(self, the_lambda: float = 0.5):
        super().__init__()
        self.the_lambda = the_lambda

    def updateOutput(self, input):
        return F.hardshrink(input, self.the_lambda)


class Abs(Module):
    def updateOutput(self, input):
        return input.abs()


class Exp(Module):
    def updateOutput(self, input):
        return input.exp()


class Log(Module):
    def updateOutput(self, input):
        return input.log()


class Sqrt(Module):
    def updateOutput(self, input):
        return input.sqrt()


class Square(Module):
    def updateOutput(self, input):
        return input * input


class Power(Module):
    """(shift + scale * x)^power (S/nn/Power.scala)."""

    def __init__(self, power: float, scale: float = 1.0, shift: float = 0.0):
        super().__init__()
        self.power, self.scale, self.shift = power, scale, shift

    def updateOutput(self, input):
        return (self.shift + self.scale * input).pow(self.power)


class Negative(Module):
    def __init__(self, inplace: bool = False):
        super().__init__()

    def updateOutput(self, input):
        return -input


class AddConstant(Module):
    def __init__(self, constant_scalar: float, inplace: bool = False):
        super().__init__()
        self.constant_scalar = constant_scalar

    def updateOutput(self, input):
        return input + self.constant_scalar


class MulConstant(Module):
    def __init__(self, scalar: float, inplace: bool = False):
        super().__init__()
        self.scalar = scalar

    def updateOutput(self, input):
        return input * self.scalar


class _GradReverse(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, lam):
        ctx.lam = lam
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return -ctx.lam * g, None


class GradientReversal(Module):
    def __init__(self, the_lambda: float = 1.0):
        super().__init__()
        self.the_lambda = the_lambda

    def setLambda(self, lam: float):
        self.the_lambda = lam
        return self

    def updateOutput(self, input):
        return _GradReverse.apply(input, self.the_lambda)


# ---------------------------------------------------------------- noise
class _SpatialDropoutBase(Module):
    keep_dims: tuple = ()

    def __init__(self, init_p: float = 0.5, data_format: str = "NCHW"):
        super().__init__()
        self.p = init_p
        self.data_format = data_format.upper()

    def _mask_shape(self, shape):  # pragma: no cover
        raise NotImplementedError

    def updateOutput(self, input):
        if not self.train or self.p <= 0:
            return input
        m = torch.empty(self._mask_shape(list(input.shape)), dtype=torch.float32)
        RNG.fill_bernoulli(m, 1 - self.p)
        return input * m.to(device=input.device, dtype=input.dtype)


class SpatialDropout1D(_SpatialDropoutBase):
    def __init__(self, init_p: float = 0.5):
        super().__init__(init_p)

    def _mask_shape(self, s):
        return [s[0], 1, s[2]] if len(s) == 3 else [1, s[1]]


class SpatialDropout2D(_SpatialDropoutBase):
    def _mask_shape(self, s):
        if self.data_format == "NCHW":
            return [s[0], s[1], 1, 1] if len(s) == 4 else [s[0], 1, 1]
        return [s[0], 1, 1, s[3]] if len(s) == 4 else [1, 1, s[2]]


class SpatialDropout3D(_SpatialDropoutBase):
    def _mask_shape(self, s):
        if self.data_format in ("NCHW", "NCDHW"):
            return [s[0], s[1], 1, 1, 1] if len(s) == 5 else [s[0], 1, 1, 1]
        return [s[0], 1, 1, 1, s[4]] if len(s) == 5 else [1, 1, 1, s[3]]


class GaussianDropout(Module):
    def __init__(self, rate: float):
        super().__init__()
        self.rate = rate

    def updateOutput(self, input):
        if not self.train:
            return input
        stddev = (self.rate / (1 - self.rate)) ** 0.5
        n = torch.empty(input.shape, dtype=torch.float32)
        RNG.fill_normal(n, 1.0, stddev)
        return input * n.to(device=input.device, dtype=input.dtype)


class GaussianNoise(Module):
    def __init__(self, stddev: float):
        super().__init__()
        self.stddev = stddev

    def updateOutput(self, input):
        if not self.train:
            return input
        n = torch.empty(input.shape, dtype=torch.float32)
        RNG.fill_normal(n, 0.0, self.stddev)
        return input + n.to(device=input.device, dtype=input.dtype)


class GaussianSampler(Module):
    """Sample N(mean, exp(logvar)) with the reparameterisation trick; input = table(mean, logvar)."""

    def updateOutput(self, input):
        mean, logvar = to_table(input).to_list()
        eps = torch.empty(mean.shape, dtype=torch.float32)
        RNG.fill_normal(eps, 0.0, 1.0)
        return mean + torch.exp(0.5 * logvar) * eps.to(device=mean.device, dtype=mean.dtype)


# ---------------------------------------------------------------- in-graph penalties
class _Penalty(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, grad_fn):
        ctx.grad_fn_ = grad_fn
        ctx.save_for_backward(x)
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        return g + ctx.grad_fn_(x), None


class L1Penalty(Module):
    """Identity in forward (records the L1 loss), adds l1weight*sign(x) to the gradient (S/nn/L1Penalty.scala)."""

    def __init__(self, l1weight: int, size_average: bool = False, provide_output: bool = True):
        super().__init__()
        self.l1weight, self.size_average, self.provide_output = l1weight, size_average, provide_output
        self.loss = 0.0

    def updateOutput(self, input):
        m = self.l1weight / (input.numel() if self.size_average else 1)
        self.loss = float(m * input.detach().abs().sum())
        y = _Penalty.apply(input, lambda x: m * torch.sign(x))
        return y if self.provide_output else y


class NegativeEntropyPenalty(Module):
    """Adds beta * d(sum p log p)/dp to the gradient of a probability input (S/nn/NegativeEntropyPenalty.scala)."""

    def __init__(self, beta: float = 0.01):
        super().__init__()
        self.beta = beta
        self.loss = 0.0

    def updateOutput(self, input):
        self.loss = float(self.beta * (input.detach() * input.detach().log()).sum())
        return _Penalty.apply(input, lambda x: self.beta * (x.log() + 1))


class ActivityRegularization(Module):
    def __init__(self, l1: float = 0.0, l2: float = 0.0):
        super().__init__()
        self.l1, self.l2 = l1, l2
        self.loss = 0.0

    def updateOutput(self, input):
        d = input.detach()
        self.loss = float(self.l1 * d.abs().sum() + self.l2 * (d * d).sum())
        return _Penalty.apply(input, lambda x: self.l1 * torch.sign(x) + 2 * self.l2 * x)
