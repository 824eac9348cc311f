"""Keras-1.2.2-style layers with shape inference (S/nn/keras/*.scala, 71 files; Python mirror P/nn/keras/layer.py).

Every ``KerasLayer`` is a thin, shape-aware wrapper: ``build(input_shape)`` instantiates the underlying framework layer
("labor" in the reference, KerasLayer.scala:165-250) once the input shape — WITHOUT the batch dimension — is known, and
returns the output shape. The reference hand-derives ``computeOutputShape`` per layer; here it is measured by pushing a
two-sample zero batch through the freshly built labor on the CPU, which cannot drift from the real kernels.
Argument names, defaults, ``dim_ordering`` ("th" = NCHW default, "tf" = NHWC) and ``border_mode`` ("valid" | "same")
follow the reference."""
from __future__ import annotations

import math

import torch

from .. import layer as L
from ..initialization_method import Ones, RandomNormal, RandomUniform, Xavier, Zeros
from ..module import Container, Module, Node
from ...utils.shape import MultiShape, Shape, SingleShape


def _init_method(name: str):
    name = (name or "glorot_uniform").lower()
    if name == "glorot_uniform":
        return Xavier()
    if name == "one":
        return Ones()
    if name == "zero":
        return Zeros()
    if name == "uniform":
        return RandomUniform(-0.05, 0.05)
    if name == "normal":
        return RandomNormal(0.0, 0.05)
    raise ValueError(f"Unsupported initialization method: {name}")


def activation_module(name: str | None):
    if name is None:
        return None
    n = name.lower()
    table = {"tanh": L.Tanh, "sigmoid": L.Sigmoid, "hard_sigmoid": L.HardSigmoid, "relu": L.ReLU,
             "softplus": L.SoftPlus, "softsign": L.SoftSign, "linear": L.Identity}
    if n == "softmax":
        return L.SoftMax()
    if n in table:
        return table[n]()
    raise ValueError(f"Invalid activation: {name}")


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


class KerasLayer(Container):
    def __init__(self, input_shape=None, name=None, **kw):
        super().__init__()
        self._input_shape = None if input_shape is None else [int(v) for v in input_shape]
        self._output_shape = None
        self.labor: Module | None = None
        if name:
            self.setName(name)
        if kw:
            raise TypeError(f"Wrong arguments for {type(self).__name__}: {sorted(kw)}")

    # ---- shape plumbing
    def isBuilt(self) -> bool:
        return self.labor is not None

    def isKerasStyle(self) -> bool:
        return True

    def getInputShape(self):
        return None if self._input_shape is None else tuple([None] + list(self._input_shape))

    def getOutputShape(self):
        if self._output_shape is None:
            raise RuntimeError(f"{type(self).__name__} has not been built")
        s = self._output_shape
        if isinstance(s, list) and s and isinstance(s[0], list):
            return [tuple([None] + x) for x in s]
        return tuple([None] + list(s))

    get_input_shape, get_output_shape = getInputShape, getOutputShape

    def doBuild(self, input_shape) -> Module:  # pragma: no cover - abstract
        raise NotImplementedError

    def _example(self, shape):
        if shape and isinstance(shape[0], (list, tuple)):
            from ...utils.table import Table

            return Table(*[torch.zeros(2, *s) for s in shape])
        return torch.zeros(2, *shape)

    def build(self, input_shape):
        """``input_shape`` excludes the batch dimension; returns the output shape (also without batch)."""
        if self.labor is None:
            shp = [list(s) for s in input_shape] if input_shape and isinstance(input_shape[0], (list, tuple)) \
                else list(input_shape)
            self._input_shape = shp
            self.labor = self.doBuild(shp)
            self.modules = [self.labor]
            was = self.labor.isTraining()
            self.labor.evaluate()
            with torch.no_grad():
                out = self.labor._call(self._example(shp))
            if was:
                self.labor.training()
            self._output_shape = list(out.shape[1:]) if isinstance(out, torch.Tensor) else \
                [list(o.shape[1:]) for o in out.to_list()]
        return self._output_shape

    def computeOutputShape(self, input_shape):
        return self.build(input_shape)

    def updateOutput(self, input):
        if self.labor is None:
            if self._input_shape is None:
                shape = list(input.shape[1:]) if isinstance(input, torch.Tensor) else \
                    [list(t.shape[1:]) for t in input.to_list()]
            else:
                shape = self._input_shape
            self.build(shape)
            dev = input.device if isinstance(input, torch.Tensor) else None
            if dev is not None and dev.type != "cpu":
                self.labor.to(dev)
        return self.labor._call(input)

    # ---- functional API: layer(node)
    def __call__(self, *nodes):
        if nodes and all(isinstance(n, Node) for n in nodes) or (
                len(nodes) == 1 and isinstance(nodes[0], (list, tuple)) and nodes[0] and isinstance(nodes[0][0], Node)):
            flat = list(nodes[0]) if isinstance(nodes[0], (list, tuple)) else list(nodes)
            shapes = [getattr(n, "_keras_shape", None) for n in flat]
            if all(s is not None for s in shapes):
                out = self.build(shapes[0] if len(shapes) == 1 else shapes)
            else:
                out = None
            node = Node(self, flat)
            node._keras_shape = out
            return node
        return super().__call__(*nodes)

    def _with_activation(self, layer: Module, activation) -> Module:
        act = activation_module(activation) if isinstance(activation, (str, type(None))) else activation
        if act is None:
            return layer
        return L.Sequential().add(layer).add(act)


def Input(shape=None, name=None):
    """Placeholder node carrying its shape (Input.scala:60-80)."""
    node = L.Input(name)
    node._keras_shape = None if shape is None else [int(v) for v in shape]
    return node


class InputLayer(KerasLayer):
    def doBuild(self, s):
        return L.Identity()


# ------------------------------------------------------------------------------------------------ core
class Dense(KerasLayer):
    def __init__(self, output_dim, init="glorot_uniform", activation=None, W_regularizer=None, b_regularizer=None,
                 bias=True, input_dim=None, input_shape=None, **kw):
        if input_dim is not None and input_shape is None:
            input_shape = (input_dim,)
        super().__init__(input_shape, **kw)
        self.output_dim, self.init, self.activation, self.wr, self.br, self.bias = (output_dim, init, activation,
                                                                                   W_regularizer, b_regularizer, bias)

    def doBuild(self, s):
        lin = L.Linear(s[-1], self.output_dim, self.bias, self.wr, self.br)
        lin.setInitMethod(_init_method(self.init), Zeros())
        m = self._with_activation(lin, self.activation)
        if len(s) > 1:          # Dense on the last axis of a higher-rank input (Dense.scala:60-80 uses InferReshape)
            seq = L.Sequential().add(L.InferReshape([-1, s[-1]], False)).add(m).add(L.InferReshape([-1] + list(s[:-1]) + [self.output_dim], False))
            return seq
        return m


class Activation(KerasLayer):
    def __init__(self, activation, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.activation = activation

    def doBuild(self, s):
        if self.activation.lower() == "softmax" and len(s) >= 2:
            return L.TimeDistributed(L.SoftMax()) if len(s) == 2 else L.SoftMax()
        return activation_module(self.activation)


class SoftMax(KerasLayer):
    def doBuild(self, s):
        return L.TimeDistributed(L.SoftMax()) if len(s) == 2 else L.SoftMax()


class Dropout(KerasLayer):
    def __init__(self, p, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.p = p

    def doBuild(self, s):
        return L.Dropout(self.p)


class Flatten(KerasLayer):
    def doBuild(self, s):
        return L.Reshape([int(math.prod(s))], True)


class Reshape(KerasLayer):
    def __init__(self, target_shape, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.target_shape = list(target_shape)

    def doBuild(self, s):
        t = list(self.target_shape)
        if -1 in t:
            known = -math.prod(t)
            t[t.index(-1)] = int(math.prod(s)) // known
        return L.Reshape(t, True)


class Permute(KerasLayer):
    def __init__(self, dims, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.dims = list(dims)

    def doBuild(self, s):
        # decompose the permutation (1-based over non-batch dims) into swaps for Transpose (Permute.scala:50-75)
        cur, swaps = list(range(1, len(self.dims) + 1)), []
        for i, d in enumerate(self.dims):
            j = cur.index(d)
            if j != i:
                swaps.append((i + 2, j + 2))
                cur[i], cur[j] = cur[j], cur[i]
        return L.Sequential().add(L.Transpose(swaps)).add(L.Contiguous()) if swaps else L.Identity()


class RepeatVector(KerasLayer):
    def __init__(self, n, input_dim=None, input_shape=None, **kw):
        if input_dim is not None and input_shape is None:
            input_shape = (input_dim,)
        super().__init__(input_shape, **kw)
        self.n = n

    def doBuild(self, s):
        return L.Replicate(self.n, 2, 2)


class Masking(KerasLayer):
    def __init__(self, mask_value=0.0, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.mask_value = mask_value

    def doBuild(self, s):
        return L.Masking(self.mask_value)


class Highway(KerasLayer):
    def __init__(self, activation=None, W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.activation, self.wr, self.br, self.bias = activation, W_regularizer, b_regularizer, bias

    def doBuild(self, s):
        return L.Highway(size=s[-1], with_bias=self.bias, activation=activation_module(self.activation),
                         wRegularizer=self.wr, bRegularizer=self.br)


class MaxoutDense(KerasLayer):
    def __init__(self, output_dim, nb_feature=4, W_regularizer=None, b_regularizer=None, bias=True, input_dim=None,
                 input_shape=None, **kw):
        if input_dim is not None and input_shape is None:
            input_shape = (input_dim,)
        super().__init__(input_shape, **kw)
        self.output_dim, self.nb, self.wr, self.br, self.bias = output_dim, nb_feature, W_regularizer, b_regularizer, bias

    def doBuild(self, s):
        return L.Maxout(s[-1], self.output_dim, self.nb, self.bias, self.wr, self.br)


class Embedding(KerasLayer):
    """Indices are 0-based as in Keras; the +1 for the 1-based LookupTable is applied inside (Embedding.scala:55-75)."""

    def __init__(self, input_dim, output_dim, init="uniform", W_regularizer=None, input_length=None, input_shape=None,
                 **kw):
        if input_length is not None and input_shape is None:
            input_shape = (input_length,)
        super().__init__(input_shape, **kw)
        self.input_dim, self.output_dim, self.init, self.wr = input_dim, output_dim, init, W_regularizer

    def doBuild(self, s):
        lt = L.LookupTable(self.input_dim, self.output_dim, wRegularizer=self.wr)
        lt.setInitMethod(_init_method(self.init))
        return L.Sequential().add(L.AddConstant(1.0)).add(lt)


class BatchNormalization(KerasLayer):
    def __init__(self, epsilon=0.001, mode=0, axis=1, momentum=0.99, beta_init="zero", gamma_init="one",
                 dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if mode != 0:
            raise ValueError("For BatchNormalization, only mode=0 is supported for now")
        self.eps, self.momentum, self.beta_init, self.gamma_init, self.order = (epsilon, momentum, beta_init, gamma_init,
                                                                               dim_ordering)

    def doBuild(self, s):
        fmt = "NCHW" if self.order == "th" else "NHWC"
        n = s[0] if fmt == "NCHW" else s[-1]
        if len(s) == 3:
            bn = L.SpatialBatchNormalization(n, self.eps, 1.0 - self.momentum, data_format=fmt)
        else:
            bn = L.BatchNormalization(s[-1] if len(s) == 1 else n, self.eps, 1.0 - self.momentum)
        bn.setInitMethod(_init_method(self.gamma_init), _init_method(self.beta_init))
        return bn

    # ---- running statistics (P/nn/keras/layer.py:270-300)
    def _bn(self):
        if self.labor is None:
            raise RuntimeError("BatchNormalization has not been built: add it to a model (or call build) first")
        return self.labor

    def set_running_mean(self, running_mean):
        b = self._bn()._buffers["runningMean"]
        b.copy_(torch.as_tensor(running_mean, dtype=b.dtype).reshape(b.shape))
        return self

    def set_running_std(self, running_std):
        """(the Python API's name; the value is the running VARIANCE, as in the reference)"""
        b = self._bn()._buffers["runningVar"]
        b.copy_(torch.as_tensor(running_std, dtype=b.dtype).reshape(b.shape))
        return self

    def get_running_mean(self):
        return self._bn()._buffers["runningMean"].detach().cpu().numpy()

    def get_running_std(self):
        return self._bn()._buffers["runningVar"].detach().cpu().numpy()

    setRunningMean, setRunningStd, getRunningMean, getRunningStd = (set_running_mean, set_running_std,
                                                                    get_running_mean, get_running_std)


class Merge(KerasLayer):
    """sum | mul | concat | ave | cos | dot | max over a list of inputs (Merge.scala:60-190)."""

    def __init__(self, layers=None, mode="sum", concat_axis=-1, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.mode, self.concat_axis, self.merge_layers = mode.lower(), concat_axis, layers

    def doBuild(self, s):
        m = self.mode
        if m == "sum":
            core = L.CAddTable()
        elif m == "mul":
            core = L.CMulTable()
        elif m == "max":
            core = L.CMaxTable()
        elif m == "ave":
            core = L.CAveTable()
        elif m == "concat":
            rank = len(s[0]) + 1
            axis = self.concat_axis if self.concat_axis >= 0 else rank + self.concat_axis
            core = L.JoinTable(axis + 1, rank)
        elif m == "dot":
            core = L.Sequential().add(L.DotProduct()).add(L.Reshape([1], True))
        elif m == "cos":
            core = L.Sequential().add(L.CosineDistance()).add(L.Reshape([1, 1], True))
        else:
            raise ValueError(f"Invalid merge mode: {self.mode}")
        if self.merge_layers:
            pt = L.ParallelTable()
            for lyr, shp in zip(self.merge_layers, s):
                if isinstance(lyr, KerasLayer):
                    lyr.build(shp)
                pt.add(lyr)
            return L.Sequential().add(pt).add(core)
        return core

    def build(self, input_shape):
        if self.merge_layers and self.labor is None:
            input_shape = [list(l._input_shape) for l in self.merge_layers]
            out = super().build(input_shape)
            return out
        return super().build(input_shape)


def merge(inputs, mode="sum", concat_axis=-1, name=None):
    """Functional merge of nodes (package.scala ``merge``)."""
    return Merge(mode=mode, concat_axis=concat_axis, name=name)(list(inputs))


# ------------------------------------------------------------------------------------------------ convolution
def _same_pad(k):
    return -1


class Convolution2D(KerasLayer):
    def __init__(self, nb_filter, nb_row, nb_col, init="glorot_uniform", activation=None, border_mode="valid",
                 subsample=(1, 1), dim_ordering="th", W_regularizer=None, b_regularizer=None, bias=True,
                 input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if border_mode not in ("valid", "same"):
            raise ValueError(f"Invalid border mode: {border_mode}")
        self.a = (nb_filter, nb_row, nb_col, init, activation, border_mode, _pair(subsample), dim_ordering,
                  W_regularizer, b_regularizer, bias)

    def doBuild(self, s):
        nf, kr, kc, init, act, bm, ss, order, wr, br, bias = self.a
        fmt = "NCHW" if order == "th" else "NHWC"
        cin = s[0] if fmt == "NCHW" else s[-1]
        pad = -1 if bm == "same" else 0
        conv = L.SpatialConvolution(cin, nf, kc, kr, ss[1], ss[0], pad, pad, wRegularizer=wr, bRegularizer=br,
                                    with_bias=bias, data_format=fmt)
        conv.setInitMethod(_init_method(init), Zeros())
        return self._with_activation(conv, act)


class Convolution1D(KerasLayer):
    def __init__(self, nb_filter, filter_length, init="glorot_uniform", activation=None, border_mode="valid",
                 subsample_length=1, W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.a = (nb_filter, filter_length, init, activation, border_mode, subsample_length, W_regularizer,
                  b_regularizer, bias)

    def doBuild(self, s):
        nf, fl, init, act, bm, sl, wr, br, bias = self.a
        steps, dim = s
        seq = L.Sequential()
        if bm == "same":
            total = max((math.ceil(steps / sl) - 1) * sl + fl - steps, 0)
            if total:
                seq.add(L.Padding(2, -(total // 2), 3)) if total // 2 else None
                seq.add(L.Padding(2, total - total // 2, 3))
        conv = L.TemporalConvolution(dim, nf, fl, sl, weight_regularizer=wr, bias_regularizer=br)
        conv.setInitMethod(_init_method(init), Zeros())
        seq.add(conv)
        a = activation_module(act)
        if a is not None:
            seq.add(a)
        return seq


class Convolution3D(KerasLayer):
    def __init__(self, nb_filter, kernel_dim1, kernel_dim2, kernel_dim3, init="glorot_uniform", activation=None,
                 border_mode="valid", subsample=(1, 1, 1), dim_ordering="th", W_regularizer=None, b_regularizer=None,
                 bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if dim_ordering != "th":
            raise ValueError("Convolution3D only supports dim_ordering='th'")
        self.a = (nb_filter, kernel_dim1, kernel_dim2, kernel_dim3, init, activation, border_mode, tuple(subsample),
                  W_regularizer, b_regularizer, bias)

    def doBuild(self, s):
        nf, k1, k2, k3, init, act, bm, ss, wr, br, bias = self.a
        pad = -1 if bm == "same" else 0
        conv = L.VolumetricConvolution(s[0], nf, k1, k3, k2, ss[0], ss[2], ss[1], pad, pad, pad, bias, wr, br)
        conv.setInitMethod(_init_method(init), Zeros())
        return self._with_activation(conv, act)


class AtrousConvolution2D(KerasLayer):
    def __init__(self, nb_filter, nb_row, nb_col, init="glorot_uniform", activation=None, subsample=(1, 1),
                 atrous_rate=(1, 1), dim_ordering="th", W_regularizer=None, b_regularizer=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if dim_ordering != "th":
            raise ValueError("AtrousConvolution2D only supports dim_ordering='th'")
        self.a = (nb_filter, nb_row, nb_col, init, activation, _pair(subsample), _pair(atrous_rate), W_regularizer,
                  b_regularizer)

    def doBuild(self, s):
        nf, kr, kc, init, act, ss, ar, wr, br = self.a
        conv = L.SpatialDilatedConvolution(s[0], nf, kc, kr, ss[1], ss[0], 0, 0, ar[1], ar[0], wr, br)
        conv.setInitMethod(_init_method(init), Zeros())
        return self._with_activation(conv, act)


class AtrousConvolution1D(KerasLayer):
    def __init__(self, nb_filter, filter_length, init="glorot_uniform", activation=None, subsample_length=1,
                 atrous_rate=1, W_regularizer=None, b_regularizer=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.a = (nb_filter, filter_length, init, activation, subsample_length, atrous_rate, W_regularizer, b_regularizer)

    def doBuild(self, s):
        nf, fl, init, act, sl, ar, wr, br = self.a
        steps, dim = s
        conv = L.SpatialDilatedConvolution(dim, nf, 1, fl, 1, sl, 0, 0, 1, ar, wr, br)
        conv.setInitMethod(_init_method(init), Zeros())
        seq = (L.Sequential().add(L.Transpose([(2, 3)])).add(L.Contiguous()).add(L.Reshape([dim, steps, 1], True))
               .add(conv).add(L.Squeeze(4)).add(L.Transpose([(2, 3)])).add(L.Contiguous()))
        a = activation_module(act)
        if a is not None:
            seq.add(a)
        return seq


class Deconvolution2D(KerasLayer):
    def __init__(self, nb_filter, nb_row, nb_col, output_shape=None, init="glorot_uniform", activation=None,
                 border_mode="valid", subsample=(1, 1), dim_ordering="th", W_regularizer=None, b_regularizer=None,
                 bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if dim_ordering != "th" or border_mode != "valid":
            raise ValueError("Deconvolution2D supports dim_ordering='th' and border_mode='valid' only")
        self.a = (nb_filter, nb_row, nb_col, init, activation, _pair(subsample), W_regularizer, b_regularizer, bias)

    def doBuild(self, s):
        nf, kr, kc, init, act, ss, wr, br, bias = self.a
        conv = L.SpatialFullConvolution(s[0], nf, kc, kr, ss[1], ss[0], 0, 0, 0, 0, 1, not bias, wr, br)
        conv.setInitMethod(_init_method(init), Zeros())
        return self._with_activation(conv, act)


class SeparableConvolution2D(KerasLayer):
    def __init__(self, nb_filter, nb_row, nb_col, init="glorot_uniform", activation=None, border_mode="valid",
                 subsample=(1, 1), depth_multiplier=1, dim_ordering="th", depthwise_regularizer=None,
                 pointwise_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.a = (nb_filter, nb_row, nb_col, init, activation, border_mode, _pair(subsample), depth_multiplier,
                  dim_ordering, depthwise_regularizer, pointwise_regularizer, b_regularizer, bias)

    def doBuild(self, s):
        nf, kr, kc, init, act, bm, ss, dm, order, dr, pr, br, bias = self.a
        fmt = "NCHW" if order == "th" else "NHWC"
        cin = s[0] if fmt == "NCHW" else s[-1]
        pad = -1 if bm == "same" else 0
        conv = L.SpatialSeparableConvolution(cin, nf, dm, kc, kr, ss[1], ss[0], pad, pad, bias, fmt, dr, br, pr)
        return self._with_activation(conv, act)


class LocallyConnected2D(KerasLayer):
    def __init__(self, nb_filter, nb_row, nb_col, activation=None, border_mode="valid", subsample=(1, 1),
                 dim_ordering="th", W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.a = (nb_filter, nb_row, nb_col, activation, border_mode, _pair(subsample), dim_ordering, W_regularizer,
                  b_regularizer, bias)

    def doBuild(self, s):
        nf, kr, kc, act, bm, ss, order, wr, br, bias = self.a
        fmt = "NCHW" if order == "th" else "NHWC"
        c, h, w = (s[0], s[1], s[2]) if fmt == "NCHW" else (s[2], s[0], s[1])
        pad = -1 if bm == "same" else 0
        m = L.LocallyConnected2D(c, w, h, nf, kc, kr, ss[1], ss[0], pad, pad, wRegularizer=wr, bRegularizer=br,
                                 with_bias=bias, data_format=fmt)
        return self._with_activation(m, act)


class LocallyConnected1D(KerasLayer):
    def __init__(self, nb_filter, filter_length, activation=None, border_mode="valid", subsample_length=1,
                 W_regularizer=None, b_regularizer=None, bias=True, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if border_mode != "valid":
            raise ValueError("For LocallyConnected1D, only border_mode='valid' is supported for now")
        self.a = (nb_filter, filter_length, activation, subsample_length, W_regularizer, b_regularizer)

    def doBuild(self, s):
        nf, fl, act, sl, wr, br = self.a
        m = L.LocallyConnected1D(s[0], s[1], nf, fl, sl, weight_regularizer=wr, bias_regularizer=br)
        return self._with_activation(m, act)


# ------------------------------------------------------------------------------------------------ pooling
class _Pool2D(KerasLayer):
    kind = "max"

    def __init__(self, pool_size=(2, 2), strides=None, border_mode="valid", dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if border_mode not in ("valid", "same"):
            raise ValueError(f"Invalid border mode: {border_mode}")
        self.pool, self.strides, self.bm, self.order = _pair(pool_size), _pair(strides or pool_size), border_mode, dim_ordering

    def doBuild(self, s):
        fmt = "NCHW" if self.order == "th" else "NHWC"
        pad = -1 if self.bm == "same" else 0
        if self.kind == "max":
            return L.SpatialMaxPooling(self.pool[1], self.pool[0], self.strides[1], self.strides[0], pad, pad, format=fmt)
        return L.SpatialAveragePooling(self.pool[1], self.pool[0], self.strides[1], self.strides[0], pad, pad,
                                       count_include_pad=False, format=fmt)


class MaxPooling2D(_Pool2D):
    kind = "max"


class AveragePooling2D(_Pool2D):
    kind = "avg"


class _Pool1D(KerasLayer):
    kind = "max"

    def __init__(self, pool_length=2, stride=None, border_mode="valid", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.pl, self.stride, self.bm = pool_length, stride or pool_length, border_mode

    def doBuild(self, s):
        steps, dim = s
        pad = -1 if self.bm == "same" else 0
        pool = L.SpatialMaxPooling(1, self.pl, 1, self.stride, 0, pad, format="NHWC") if self.kind == "max" else \
            L.SpatialAveragePooling(1, self.pl, 1, self.stride, 0, pad, count_include_pad=False, format="NHWC")
        return L.Sequential().add(L.Reshape([steps, 1, dim], True)).add(pool).add(L.Squeeze(3))


class MaxPooling1D(_Pool1D):
    kind = "max"


class AveragePooling1D(_Pool1D):
    kind = "avg"


class _Pool3D(KerasLayer):
    kind = "max"

    def __init__(self, pool_size=(2, 2, 2), strides=None, border_mode="valid", dim_ordering="th", input_shape=None,
                 **kw):
        super().__init__(input_shape, **kw)
        if dim_ordering != "th" or border_mode != "valid":
            raise ValueError("3-D pooling supports dim_ordering='th' and border_mode='valid' only")
        self.pool, self.strides = tuple(pool_size), tuple(strides or pool_size)

    def doBuild(self, s):
        p, st = self.pool, self.strides
        if self.kind == "max":
            return L.VolumetricMaxPooling(p[0], p[2], p[1], st[0], st[2], st[1])
        return L.VolumetricAveragePooling(p[0], p[2], p[1], st[0], st[2], st[1], count_include_pad=False)


class MaxPooling3D(_Pool3D):
    kind = "max"


class AveragePooling3D(_Pool3D):
    kind = "avg"


class _GlobalPool(KerasLayer):
    kind, rank = "max", 2

    def __init__(self, dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.order = dim_ordering

    def doBuild(self, s):
        if self.rank == 1:
            dims = [2]
        elif self.order == "th":
            dims = list(range(3, 3 + self.rank))
        else:
            dims = list(range(2, 2 + self.rank))
        seq = L.Sequential()
        for d in reversed(dims):
            seq.add(L.Max(d, len(s) + 1) if self.kind == "max" else L.Mean(d, len(s) + 1, True))
        return seq


class GlobalMaxPooling1D(_GlobalPool):
    kind, rank = "max", 1

    def __init__(self, input_shape=None, **kw):
        super().__init__("th", input_shape, **kw)


class GlobalAveragePooling1D(_GlobalPool):
    kind, rank = "avg", 1

    def __init__(self, input_shape=None, **kw):
        super().__init__("th", input_shape, **kw)


class GlobalMaxPooling2D(_GlobalPool):
    kind, rank = "max", 2


class GlobalAveragePooling2D(_GlobalPool):
    kind, rank = "avg", 2


class GlobalMaxPooling3D(_GlobalPool):
    kind, rank = "max", 3


class GlobalAveragePooling3D(_GlobalPool):
    kind, rank = "avg", 3


# ------------------------------------------------------------------------------------------------ padding / resize
class ZeroPadding1D(KerasLayer):
    def __init__(self, padding=1, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.padding = _pair(padding)

    def doBuild(self, s):
        return L.Sequential().add(L.Padding(2, -self.padding[0], 3)).add(L.Padding(2, self.padding[1], 3))


class ZeroPadding2D(KerasLayer):
    def __init__(self, padding=(1, 1), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        p = tuple(padding)
        self.pad = (p[0], p[0], p[1], p[1]) if len(p) == 2 else p     # top, bottom, left, right
        self.order = dim_ordering

    def doBuild(self, s):
        t, b, l, r = self.pad
        hd, wd = (3, 4) if self.order == "th" else (2, 3)
        return (L.Sequential().add(L.Padding(hd, -t, 4)).add(L.Padding(hd, b, 4)).add(L.Padding(wd, -l, 4))
                .add(L.Padding(wd, r, 4)))


class ZeroPadding3D(KerasLayer):
    def __init__(self, padding=(1, 1, 1), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.padding, self.order = tuple(padding), dim_ordering

    def doBuild(self, s):
        first = 3 if self.order == "th" else 2
        seq = L.Sequential()
        for i, p in enumerate(self.padding):
            seq.add(L.Padding(first + i, -p, 5)).add(L.Padding(first + i, p, 5))
        return seq


class Cropping1D(KerasLayer):
    def __init__(self, cropping=(1, 1), input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.cropping = tuple(cropping)

    def doBuild(self, s):
        return L.Narrow(2, 1 + self.cropping[0], s[0] - self.cropping[0] - self.cropping[1])


class Cropping2D(KerasLayer):
    def __init__(self, cropping=((0, 0), (0, 0)), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.cropping, self.order = cropping, dim_ordering

    def doBuild(self, s):
        return L.Cropping2D(list(self.cropping[0]), list(self.cropping[1]), "NCHW" if self.order == "th" else "NHWC")


class Cropping3D(KerasLayer):
    def __init__(self, cropping=((1, 1), (1, 1), (1, 1)), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.cropping, self.order = cropping, dim_ordering

    def doBuild(self, s):
        c = self.cropping
        return L.Cropping3D(list(c[0]), list(c[1]), list(c[2]), "channel_first" if self.order == "th" else "channel_last")


class UpSampling1D(KerasLayer):
    def __init__(self, length=2, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.length = length

    def doBuild(self, s):
        return L.UpSampling1D(self.length)


class UpSampling2D(KerasLayer):
    def __init__(self, size=(2, 2), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.size, self.order = tuple(size), dim_ordering

    def doBuild(self, s):
        return L.UpSampling2D(list(self.size), "NCHW" if self.order == "th" else "NHWC")


class UpSampling3D(KerasLayer):
    def __init__(self, size=(2, 2, 2), dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.size = tuple(size)

    def doBuild(self, s):
        return L.UpSampling3D(list(self.size))


# ------------------------------------------------------------------------------------------------ activations / noise
class LeakyReLU(KerasLayer):
    def __init__(self, alpha=0.01, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.alpha = alpha

    def doBuild(self, s):
        return L.LeakyReLU(self.alpha)


class ELU(KerasLayer):
    def __init__(self, alpha=1.0, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.alpha = alpha

    def doBuild(self, s):
        return L.ELU(self.alpha)


class ThresholdedReLU(KerasLayer):
    def __init__(self, theta=1.0, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.theta = theta

    def doBuild(self, s):
        return L.Threshold(self.theta, 0.0)


class SReLU(KerasLayer):
    def __init__(self, t_left_init="zero", a_left_init="glorot_uniform", t_right_init="glorot_uniform",
                 a_right_init="one", shared_axes=None, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.shared_axes = shared_axes

    def doBuild(self, s):
        return L.SReLU(list(s), self.shared_axes)


class GaussianNoise(KerasLayer):
    def __init__(self, sigma, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.sigma = sigma

    def doBuild(self, s):
        return L.GaussianNoise(self.sigma)


class GaussianDropout(KerasLayer):
    def __init__(self, p, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.p = p

    def doBuild(self, s):
        return L.GaussianDropout(self.p)


class SpatialDropout1D(KerasLayer):
    def __init__(self, p=0.5, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.p = p

    def doBuild(self, s):
        return L.SpatialDropout1D(self.p)


class SpatialDropout2D(KerasLayer):
    def __init__(self, p=0.5, dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.p, self.order = p, dim_ordering

    def doBuild(self, s):
        return L.SpatialDropout2D(self.p, "NCHW" if self.order == "th" else "NHWC")


class SpatialDropout3D(KerasLayer):
    def __init__(self, p=0.5, dim_ordering="th", input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.p, self.order = p, dim_ordering

    def doBuild(self, s):
        return L.SpatialDropout3D(self.p, "NCHW" if self.order == "th" else "NHWC")


# ------------------------------------------------------------------------------------------------ recurrent
class _Recurrent(KerasLayer):
    def __init__(self, output_dim, return_sequences=False, go_backwards=False, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.output_dim, self.return_sequences, self.go_backwards = output_dim, return_sequences, go_backwards

    def buildCell(self, s) -> Module:  # pragma: no cover - abstract
        raise NotImplementedError

    def doBuild(self, s):
        seq = L.Sequential()
        if self.go_backwards:
            seq.add(L.Reverse(2))
        seq.add(L.Recurrent().add(self.buildCell(s)))
        if not self.return_sequences:
            seq.add(L.Select(2, -1))
        return seq


class SimpleRNN(_Recurrent):
    def __init__(self, output_dim, activation="tanh", return_sequences=False, go_backwards=False, W_regularizer=None,
                 U_regularizer=None, b_regularizer=None, input_shape=None, **kw):
        super().__init__(output_dim, return_sequences, go_backwards, input_shape, **kw)
        self.activation, self.regs = activation, (W_regularizer, U_regularizer, b_regularizer)

    def buildCell(self, s):
        return L.RnnCell(s[1], self.output_dim, activation_module(self.activation), True, True, *self.regs)


class LSTM(_Recurrent):
    def __init__(self, output_dim, activation="tanh", inner_activation="hard_sigmoid", return_sequences=False,
                 go_backwards=False, W_regularizer=None, U_regularizer=None, b_regularizer=None, input_shape=None, **kw):
        super().__init__(output_dim, return_sequences, go_backwards, input_shape, **kw)
        self.activation, self.inner, self.regs = activation, inner_activation, (W_regularizer, U_regularizer, b_regularizer)

    def buildCell(self, s):
        return L.LSTM(s[1], self.output_dim, 0.0, activation_module(self.activation), activation_module(self.inner),
                      *self.regs)


class GRU(_Recurrent):
    def __init__(self, output_dim, activation="tanh", inner_activation="hard_sigmoid", return_sequences=False,
                 go_backwards=False, W_regularizer=None, U_regularizer=None, b_regularizer=None, input_shape=None, **kw):
        super().__init__(output_dim, return_sequences, go_backwards, input_shape, **kw)
        self.activation, self.inner, self.regs = activation, inner_activation, (W_regularizer, U_regularizer, b_regularizer)

    def buildCell(self, s):
        return L.GRU(s[1], self.output_dim, 0.0, activation_module(self.activation), activation_module(self.inner),
                     *self.regs)


class ConvLSTM2D(_Recurrent):
    def __init__(self, nb_filter, nb_row, nb_col=None, activation="tanh", inner_activation="hard_sigmoid",
                 dim_ordering="th", subsample=(1, 1), border_mode="same", W_regularizer=None, U_regularizer=None,
                 b_regularizer=None, return_sequences=False, go_backwards=False, input_shape=None, **kw):
        super().__init__(nb_filter, return_sequences, go_backwards, input_shape, **kw)
        if dim_ordering != "th":
            raise ValueError("ConvLSTM2D only supports dim_ordering='th'")
        if nb_col is not None and nb_col != nb_row:
            raise ValueError("ConvLSTM2D only supports square kernels")
        self.k, self.activation, self.inner, self.stride = nb_row, activation, inner_activation, _pair(subsample)[0]
        self.regs = (W_regularizer, U_regularizer, b_regularizer)

    def buildCell(self, s):
        return L.ConvLSTMPeephole(s[1], self.output_dim, self.k, self.k, self.stride, -1,
                                  activation_module(self.activation), activation_module(self.inner), *self.regs,
                                  with_peephole=False)


class Bidirectional(KerasLayer):
    def __init__(self, layer, merge_mode="concat", input_shape=None, **kw):
        super().__init__(input_shape or layer._input_shape, **kw)
        if not isinstance(layer, _Recurrent):
            raise TypeError("Bidirectional wraps a recurrent layer")
        self.layer, self.merge_mode = layer, merge_mode.lower()

    def doBuild(self, s):
        merge = {"concat": L.JoinTable(3, 3), "sum": L.CAddTable(), "mul": L.CMulTable(), "ave": L.CAveTable()}[self.merge_mode]
        seq = L.Sequential().add(L.BiRecurrent(merge).add(self.layer.buildCell(s)))
        if not self.layer.return_sequences:
            seq.add(L.Select(2, -1))
        return seq


class TimeDistributed(KerasLayer):
    def __init__(self, layer, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        self.layer = layer

    def doBuild(self, s):
        inner = self.layer
        if isinstance(inner, KerasLayer):
            inner.build(s[1:])
        return L.TimeDistributed(inner)


# ------------------------------------------------------------------------------------------------ wrappers
class KerasLayerWrapper(KerasLayer):
    """Use a Torch-style layer inside a Keras-style model (KerasLayer.scala:95-111): the output shape is found by
    running the wrapped layer on a batch of ones."""

    def __init__(self, torch_layer: Module, input_shape=None, **kw):
        super().__init__(input_shape, **kw)
        if getattr(torch_layer, "isKerasStyle", lambda: False)():
            raise ValueError(f"We only accept torch layer here, but got: {torch_layer}")
        self.torchLayer = torch_layer

    def doBuild(self, s):
        return self.torchLayer


class KerasIdentityWrapper(KerasLayerWrapper):
    """The same for a layer known to keep its input shape (KerasLayer.scala:76-86)."""

    def computeOutputShape(self, input_shape):
        self.build(input_shape)
        return list(input_shape)


class KerasUtils:
    """The lookups the Keras-style layers share (S/nn/keras/KerasUtils.scala)."""

    @staticmethod
    def getInitMethod(init: str):
        return _init_method(init)

    @staticmethod
    def getActivation(activation):
        return activation_module(activation)

    getKerasActivation = getActivation

    @staticmethod
    def getPadsFromBorderMode(border_mode: str = "valid"):
        """"same" is the −1 the Torch-style layers read as TensorFlow SAME padding; "valid" is 0."""
        if border_mode not in ("valid", "same"):
            raise ValueError(f"Invalid border mode: {border_mode}")
        return (-1, -1) if border_mode == "same" else (0, 0)

    @staticmethod
    def toBigDLFormat(dim_ordering: str = "th") -> str:
        d = str(dim_ordering).lower()
        if d not in ("tf", "th"):
            raise ValueError(f"Invalid dim_ordering: {dim_ordering}")
        return "NHWC" if d == "tf" else "NCHW"

    @staticmethod
    def toBigDLFormat5D(dim_ordering: str = "th") -> str:
        d = str(dim_ordering).lower()
        if d not in ("tf", "th"):
            raise ValueError(f"Invalid dim_ordering: {dim_ordering}")
        return "CHANNEL_LAST" if d == "tf" else "CHANNEL_FIRST"

    @staticmethod
    def addBatch(shape):
        return None if shape is None else tuple([None] + list(shape))

    @staticmethod
    def removeBatch(shape):
        return None if shape is None else tuple(list(shape)[1:])


# the reference's names of the abstract pooling bases (S/nn/keras/{Pooling1D,Pooling2D,Pooling3D,GlobalPooling*}.scala)
Pooling1D, Pooling2D, Pooling3D = _Pool1D, _Pool2D, _Pool3D
GlobalPooling1D = GlobalPooling2D = GlobalPooling3D = _GlobalPool
