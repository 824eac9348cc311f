"""Linear-family, embedding and miscellaneous normalisation layers.

Reference files (S/nn/): SparseLinear, Bilinear, Add, CAdd, Mul, CMul, Scale, Cosine, Euclidean, LookupTable,
LookupTableSparse, Maxout, Highway, LayerNormalization, Normalize, NormalizeScale, SpatialWithinChannelLRN,
SpatialContrastiveNormalization, SpatialDivisiveNormalization, SpatialSubtractiveNormalization, DenseToSparse,
SparseJoinTable.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from ..utils.random_generator import RNG
from ..utils.table import Table, to_table
from .initialization_method import ConstInitMethod, Ones, RandomNormal, RandomUniform, VariableFormat, Zeros
from .layers_core import Linear
from .module import Container, Module, Sequential


def _uniform(t, stdv):
    return RNG.fill_uniform(t, -stdv, stdv)


class SparseLinear(Linear):
    """Linear on a sparse (COO) 2-D input; backward only produces parameter gradients unless ``backward_start`` /
    ``backward_length`` select a dense slice of the input to differentiate (S/nn/SparseLinear.scala)."""

    def __init__(self, input_size, output_size, backward_start: int = -1, backward_length: int = -1, with_bias=True,
                 wRegularizer=None, bRegularizer=None, init_weight=None, init_bias=None, init_grad_weight=None,
                 init_grad_bias=None):
        super().__init__(input_size, output_size, with_bias, wRegularizer, bRegularizer, init_weight, init_bias)
        self.backward_start, self.backward_length = backward_start, backward_length

    def updateOutput(self, input):
        x = input.to_dense() if input.is_sparse else input
        if self.backward_start > 0 and self.backward_length > 0:
            s = self.backward_start - 1
            left, mid, right = x[:, :s].detach(), x[:, s:s + self.backward_length], x[:, s + self.backward_length:].detach()
            x = torch.cat([left, mid, right], dim=1)
        elif input.is_sparse:
            x = x.detach()
        return F.linear(x, self.weight, self._parameters.get("bias"))


class Bilinear(Module):
    """y_k = x1^T W_k x2 + b_k (S/nn/Bilinear.scala)."""

    def __init__(self, input_size1: int, input_size2: int, output_size: int, bias_res: bool = True,
                 wRegularizer=None, bRegularizer=None):
        super().__init__()
        self.in1, self.in2, self.out, self.bias_res = input_size1, input_size2, output_size, bias_res
        self.wRegularizer, self.bRegularizer = wRegularizer, bRegularizer
        self.register_parameter("weight", torch.zeros(output_size, input_size1, input_size2))
        if bias_res:
            self.register_parameter("bias", torch.zeros(output_size))
        self.reset()

    def reset(self):
        stdv = 1.0 / math.sqrt(self.in1)
        _uniform(self.weight.data, stdv)
        if self.bias_res:
            _uniform(self.bias.data, stdv)
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        a, b = to_table(input).to_list()
        return F.bilinear(a, b, self.weight, self._parameters.get("bias"))


class Add(Module):
    """Learned bias vector added to the input (S/nn/Add.scala)."""

    def __init__(self, input_size: int):
        super().__init__()
        self.input_size = input_size
        self.register_parameter("bias", torch.zeros(input_size))
        self.reset()

    def reset(self):
        _uniform(self.bias.data, 1.0 / math.sqrt(self.input_size))
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        return input + self.bias.to(input.dtype).reshape((1,) * (input.dim() - 1) + (-1,)) if input.dim() > 1 \
            else input + self.bias.to(input.dtype)


class CAdd(Module):
    """Learned bias of arbitrary (broadcastable) shape (S/nn/CAdd.scala)."""

    def __init__(self, size, bRegularizer=None):
        super().__init__()
        self.size = list(size)
        self.bRegularizer = bRegularizer
        self.register_parameter("bias", torch.zeros(self.size))
        self.reset()

    def reset(self):
        _uniform(self.bias.data, 1.0 / math.sqrt(self.bias.numel()))
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        b = self.bias.to(input.dtype)
        if b.dim() == input.dim() - 1:
            b = b.unsqueeze(0)
        return input + b


class Mul(Module):
    """Single learned scalar gain (S/nn/Mul.scala)."""

    def __init__(self):
        super().__init__()
        self.register_parameter("weight", torch.zeros(1))
        self.reset()

    def reset(self):
        _uniform(self.weight.data, 1.0)
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        return input * self.weight.to(input.dtype)


class CMul(Module):
    """Learned gain of arbitrary (broadcastable) shape (S/nn/CMul.scala)."""

    def __init__(self, size, wRegularizer=None):
        super().__init__()
        self.size = list(size)
        self.wRegularizer = wRegularizer
        self.register_parameter("weight", torch.zeros(self.size))
        self.reset()

    def reset(self):
        _uniform(self.weight.data, 1.0 / math.sqrt(self.weight.numel()))
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        w = self.weight.to(input.dtype)
        if w.dim() == input.dim() - 1:
            w = w.unsqueeze(0)
        return input * w


class Scale(Container):
    """CMul followed by CAdd (Caffe's Scale layer, S/nn/Scale.scala)."""

    def __init__(self, size):
        super().__init__()
        self.cmul, self.cadd = CMul(size), CAdd(size)
        self.add(self.cmul)
        self.add(self.cadd)

    def updateOutput(self, input):
        return self.cadd._call(self.cmul._call(input))


class Cosine(Module):
    """Cosine similarity between the input and each of ``output_size`` learned vectors (S/nn/Cosine.scala)."""

    def __init__(self, input_size: int, output_size: int):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.register_parameter("weight", torch.zeros(output_size, input_size))
        self.reset()

    def reset(self):
        _uniform(self.weight.data, 1.0 / math.sqrt(self.input_size))
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        x = input if input.dim() > 1 else input.unsqueeze(0)
        y = F.linear(F.normalize(x, dim=1, eps=1e-12), F.normalize(self.weight, dim=1, eps=1e-12))
        return y if input.dim() > 1 else y.squeeze(0)


class Euclidean(Module):
    """Euclidean distance between the input and each learned centre (S/nn/Euclidean.scala); weight is [in, out]."""

    def __init__(self, input_size: int, output_size: int, fast_backward: bool = True):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.register_parameter("weight", torch.zeros(input_size, output_size))
        self.reset()

    def reset(self):
        _uniform(self.weight.data, 1.0 / math.sqrt(self.input_size))
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        x = input if input.dim() > 1 else input.unsqueeze(0)
        d = (x.unsqueeze(2) - self.weight.unsqueeze(0)).pow(2).sum(dim=1).clamp_min(0).sqrt()
        return d if input.dim() > 1 else d.squeeze(0)


class LookupTable(Module):
    """Embedding with 1-based indices, optional max-norm renormalisation and padding index whose gradient is dropped
    (S/nn/LookupTable.scala)."""

    def __init__(self, n_index: int, n_output: int, padding_value: float = 0.0, max_norm: float = float("inf"),
                 norm_type: float = 2.0, should_scale_grad_by_freq: bool = False, wRegularizer=None,
                 mask_zero: bool = False):
        super().__init__()
        self.n_index, self.n_output = n_index, n_output
        self.padding_value, self.max_norm, self.norm_type = padding_value, max_norm, norm_type
        self.should_scale_grad_by_freq, self.mask_zero = should_scale_grad_by_freq, mask_zero
        self.wRegularizer = wRegularizer
        self.register_parameter("weight", torch.zeros(n_index, n_output))
        self.reset()

    def reset(self):
        self._init_methods.get("weight", RandomNormal(0.0, 1.0))(self.weight.data, VariableFormat.ONE_D)
        self.zeroGradParameters()
        self._mark_weights_dirty()
        return self

    def updateOutput(self, input):
        idx = input.long()
        if self.mask_zero:
            idx0 = torch.clamp(idx - 1, min=0)
        else:
            idx0 = idx - 1
        if self.max_norm != float("inf"):
            with torch.no_grad():
                rows = torch.unique(idx0)
                w = self.weight[rows]
                n = w.norm(p=self.norm_type, dim=1, keepdim=True)
                self.weight[rows] = torch.where(n > self.max_norm, w * (self.max_norm / (n + 1e-7)), w)
        pad = int(self.padding_value) - 1 if self.padding_value != 0 else None
        out = F.embedding(idx0, self.weight, padding_idx=pad if (pad is not None and 0 <= pad < self.n_index) else None,
                          scale_grad_by_freq=self.should_scale_grad_by_freq)
        if self.mask_zero:
            out = out * (idx != 0).unsqueeze(-1).to(out.dtype)
        return out


class LookupTableSparse(Module):
    """embedding_lookup_sparse: input = sparse ids [B, n] (or table(ids, weights)); rows are combined with
    sum / mean / sqrtn (S/nn/LookupTableSparse.scala)."""

    def __init__(self, n_index: int, n_output: int, combiner: str = "sum", max_norm: float = -1.0, wRegularizer=None):
        super().__init__()
        self.n_index, self.n_output, self.combiner, self.max_norm = n_index, n_output, combiner.lower(), max_norm
        self.wRegularizer = wRegularizer
        self.register_parameter("weight", torch.zeros(n_index, n_output))
        self.reset()

    def reset(self):
        self._init_methods.get("weight", RandomNormal(0.0, 1.0))(self.weight.data, VariableFormat.ONE_D)
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        if isinstance(input, torch.Tensor):
            ids, wts = input, None
        else:
            ids, wts = to_table(input).to_list()
        ids = ids.coalesce() if ids.is_sparse else ids.to_sparse().coalesce()
        rows = ids.indices()[0]
        vals = ids.values().long() - 1
        w = wts.coalesce().values().to(self.weight.dtype) if wts is not None else torch.ones_like(vals, dtype=self.weight.dtype)
        emb = self.weight[vals]
        if self.max_norm > 0:
            n = emb.norm(dim=1, keepdim=True)
            emb = torch.where(n > self.max_norm, emb * (self.max_norm / (n + 1e-7)), emb)
        B = ids.shape[0]
        out = torch.zeros(B, self.n_output, dtype=emb.dtype, device=emb.device).index_add(0, rows, emb * w.unsqueeze(1))
        if self.combiner == "mean":
            den = torch.zeros(B, dtype=emb.dtype, device=emb.device).index_add(0, rows, w)
            out = out / den.clamp_min(1e-12).unsqueeze(1)
        elif self.combiner == "sqrtn":
            den = torch.zeros(B, dtype=emb.dtype, device=emb.device).index_add(0, rows, w * w).sqrt()
            out = out / den.clamp_min(1e-12).unsqueeze(1)
        return out


class Maxout(Container):
    """Linear to ``output_size * maxout_number`` followed by a max over the pieces (S/nn/Maxout.scala)."""

    def __init__(self, input_size, output_size, maxout_number, with_bias=True, w_regularizer=None, b_regularizer=None,
                 init_weight=None, init_bias=None):
        super().__init__()
        self.output_size, self.maxout_number = output_size, maxout_number
        self.linear = Linear(input_size, output_size * maxout_number, with_bias, w_regularizer, b_regularizer,
                             init_weight, init_bias)
        self.add(self.linear)

    def updateOutput(self, input):
        y = self.linear._call(input)
        return y.reshape(-1, self.maxout_number, self.output_size).max(dim=1).values


def Highway(size: int, with_bias: bool = True, activation: Module | None = None, wRegularizer=None, bRegularizer=None):
    """y = T(x) * H(x) + (1 - T(x)) * x with T = sigmoid(Linear), H = activation(Linear) (S/nn/Highway.scala:28-58)."""
    from .layers_activation import Sigmoid
    from .layers_table import CAddTable, CMulTable
    from .module import Graph, Input
    from .layers_activation import AddConstant, Negative

    inp = Input()
    gate = Sigmoid().inputs(Linear(size, size, with_bias, wRegularizer, bRegularizer).inputs(inp))
    neg_gate = AddConstant(1.0).inputs(Negative().inputs(gate))
    h = Linear(size, size, with_bias, wRegularizer, bRegularizer).inputs(inp)
    if activation is not None:
        h = activation.inputs(h)
    out = CAddTable().inputs(CMulTable().inputs(gate, h), CMulTable().inputs(neg_gate, inp))
    return Graph(inp, out)


# ---------------------------------------------------------------- normalisation (non-batch)
class LayerNormalization(Module):
    """Normalise over the last dim with learned gain/bias (S/nn/LayerNormalization.scala)."""

    def __init__(self, hidden_size: int, eps: float = 1e-6):
        super().__init__()
        self.hidden_size, self.eps = hidden_size, eps
        self.register_parameter("weight", torch.ones(hidden_size))
        self.register_parameter("bias", torch.zeros(hidden_size))

    def reset(self):
        self.weight.data.fill_(1.0)
        self.bias.data.zero_()
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        x = input.float()
        mean = x.mean(dim=-1, keepdim=True)
        var = (x - mean).pow(2).mean(dim=-1, keepdim=True)
        return ((x - mean) * torch.rsqrt(var + self.eps) * self.weight + self.bias).to(input.dtype)


class Normalize(Module):
    """L_p-normalise each sample over dim 1 (S/nn/Normalize.scala)."""

    def __init__(self, p: float, eps: float = 1e-10):
        super().__init__()
        self.p, self.eps = p, eps

    def updateOutput(self, input):
        x = input if input.dim() > 1 else input.unsqueeze(0)
        if math.isinf(self.p):
            n = x.abs().amax(dim=1, keepdim=True)
        else:
            n = x.abs().pow(self.p).sum(dim=1, keepdim=True).add(self.eps).pow(1.0 / self.p)
        y = x / n
        return y if input.dim() > 1 else y.squeeze(0)


class NormalizeScale(Container):
    """Normalize followed by a learned per-channel scale (SSD's conv4_3 norm, S/nn/NormalizeScale.scala)."""

    def __init__(self, p: float, eps: float = 1e-10, scale: float = 1.0, size=None, w_regularizer=None):
        super().__init__()
        self.normalize = Normalize(p, eps)
        self.cmul = CMul(size, w_regularizer)
        with torch.no_grad():
            self.cmul.weight.fill_(scale)
        self.add(self.normalize)
        self.add(self.cmul)

    def updateOutput(self, input):
        return self.cmul._call(self.normalize._call(input))


class SpatialWithinChannelLRN(Module):
    """LRN over a spatial window inside each channel: x / (1 + alpha/size^2 * avg-window-sum(x^2))^beta
    (S/nn/SpatialWithinChannelLRN.scala)."""

    def __init__(self, size: int = 5, alpha: float = 1.0, beta: float = 0.75):
        super().__init__()
        self.size, self.alpha, self.beta = size, alpha, beta

    def updateOutput(self, input):
        x = input if input.dim() == 4 else input.unsqueeze(0)
        pad = (self.size - 1) // 2
        s = F.avg_pool2d(x * x, self.size, 1, pad, count_include_pad=True)
        y = x * (1 + self.alpha * s).pow(-self.beta)
        return y if input.dim() == 4 else y.squeeze(0)


def _gaussian_kernel1d(n: int) -> torch.Tensor:
    k = torch.arange(n, dtype=torch.float32) - (n - 1) / 2.0
    g = torch.exp(-(k / (0.25 * n)) ** 2 / 2)
    return g / g.sum()


class SpatialSubtractiveNormalization(Module):
    """Subtract the kernel-weighted neighbourhood mean (across all input planes), with border-aware normalisation
    (S/nn/SpatialSubtractiveNormalization.scala)."""

    def __init__(self, n_input_plane: int = 1, kernel=None):
        super().__init__()
        self.n_input_plane = n_input_plane
        k = torch.ones(9, 9) if kernel is None else torch.as_tensor(kernel, dtype=torch.float32)
        if k.dim() == 1:
            k = torch.outer(k, k)
        self.kernel = k / (k.sum() * n_input_plane)

    def _mean(self, x):
        kh, kw = self.kernel.shape
        w = self.kernel.to(x.device, x.dtype).expand(1, x.shape[1], kh, kw)
        num = F.conv2d(x, w, padding=(kh // 2, kw // 2))
        ones = torch.ones(1, x.shape[1], x.shape[2], x.shape[3], dtype=x.dtype, device=x.device)
        coef = F.conv2d(ones, w, padding=(kh // 2, kw // 2))
        return num / coef

    def updateOutput(self, input):
        x = input if input.dim() == 4 else input.unsqueeze(0)
        y = x - self._mean(x)
        return y if input.dim() == 4 else y.squeeze(0)


class SpatialDivisiveNormalization(SpatialSubtractiveNormalization):
    """Divide by the kernel-weighted neighbourhood standard deviation, thresholded from below
    (S/nn/SpatialDivisiveNormalization.scala)."""

    def __init__(self, n_input_plane: int = 1, kernel=None, threshold: float = 1e-4, thresval: float = 1e-4):
        super().__init__(n_input_plane, kernel)
        self.threshold, self.thresval = threshold, thresval

    def updateOutput(self, input):
        x = input if input.dim() == 4 else input.unsqueeze(0)
        std = self._mean(x * x).clamp_min(0).sqrt()
        std = torch.where(std > self.threshold, std, torch.full_like(std, self.thresval))
        y = x / std
        return y if input.dim() == 4 else y.squeeze(0)


class SpatialContrastiveNormalization(Container):
    """Subtractive followed by divisive normalisation (S/nn/SpatialContrastiveNormalization.scala)."""

    def __init__(self, n_input_plane: int = 1, kernel=None, threshold: float = 1e-4, thresval: float = 1e-4):
        super().__init__()
        self.sub = SpatialSubtractiveNormalization(n_input_plane, kernel)
        self.div = SpatialDivisiveNormalization(n_input_plane, kernel, threshold, thresval)
        self.add(self.sub)
        self.add(self.div)

    def updateOutput(self, input):
        return self.div._call(self.sub._call(input))


# ---------------------------------------------------------------- sparse helpers
class DenseToSparse(Module):
    def __init__(self, propagate_back: bool = True):
        super().__init__()
        self.propagate_back = propagate_back

    def updateOutput(self, input):
        return (input if self.propagate_back else input.detach()).to_sparse()


class SparseJoinTable(Module):
    """Concatenate sparse 2-D tensors along ``dimension`` (S/nn/SparseJoinTable.scala)."""

    def __init__(self, dimension: int):
        super().__init__()
        self.dimension = dimension

    def updateOutput(self, input):
        xs = [x if x.is_sparse else x.to_sparse() for x in to_table(input).to_list()]
        return torch.cat(xs, dim=self.dimension - 1)
