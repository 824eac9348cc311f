"""Containers over Tables and table/shape operators.

Reference files (S/nn/): Concat, ConcatTable, ParallelTable, MapTable, Bottle, JoinTable, SplitTable,
BifurcateSplitTable, FlattenTable, NarrowTable, SelectTable, CAddTable, CSubTable, CMulTable, CDivTable, CMaxTable,
CMinTable, CAveTable, MixtureTable, Pack, Reshape, InferReshape, View, Squeeze, Unsqueeze, Transpose, Contiguous,
Narrow, Select, Index, MaskedSelect, Masking, Replicate, Tile, ExpandSize, Reverse, Padding, SpatialZeroPadding,
Cropping2D, Cropping3D, Sum, Mean, Max, Min, DotProduct, CrossProduct, PairwiseDistance, CosineDistance, MM, MV,
TableOperation. Dimensions and indices in constructors are 1-based like the reference; negative dims count from the end.
"""
from __future__ import annotations

import torch

from ..utils import config
import torch.nn.functional as F

from ..utils.table import Table, flatten_activity, to_table
from .module import Container, Module


def _dim(d: int, ndim: int, n_input_dims: int = -1) -> int:
    """1-based (possibly negative) reference dimension → 0-based torch dim, shifting by the batch dim when the input
    has more dims than ``n_input_dims``."""
    if d < 0:
        return ndim + d
    z = d - 1
    if n_input_dims > 0 and ndim == n_input_dims + 1:
        z += 1
    return z


_SIDE: dict = {}


def _side_streams(device, n: int):
    pool = _SIDE.setdefault(device.index, [])
    while len(pool) < n:
        pool.append(torch.cuda.Stream(device=device))
    return pool[:n]


def _tl(x) -> list:
    return to_table(x).to_list()


# ---------------------------------------------------------------- containers
class Concat(Container):
    """Apply every child to the same input and concatenate along ``dimension`` (S/nn/Concat.scala)."""

    def __init__(self, dimension: int):
        super().__init__()
        self.dimension = dimension

    def _tail_conv(self, branch):
        """The convolution that produces ``branch``'s final activation (ReLUs folded into it are skipped), or None."""
        from .layers_core import ReLU, SpatialConvolution

        m = branch
        while isinstance(m, Container) and type(m).__name__ == "Sequential" and m.modules:
            last = m.modules[-1]
            if type(last) is ReLU and getattr(last, "_fused", False) and len(m.modules) >= 2:
                last = m.modules[-2]
            m = last
        return m if isinstance(m, SpatialConvolution) and m.data_format == "NCHW" else None

    def updateOutput(self, input):
        from .layers_core import native

        if (self.dimension == 2 and isinstance(input, torch.Tensor) and input.dim() == 4 and native(input)
                and input.dtype == torch.bfloat16 and getattr(self, "_zero_copy", True)):
            out = self._zero_copy_forward(input)
            if out is not None:
                return out
        outs = [m._call(input) for m in self.modules]
        return torch.cat(outs, dim=self.dimension - 1)

    def _zero_copy_forward(self, input):
        """Channel concat without the copy: every branch's last convolution writes straight into its channel slice of
        one NHWC buffer (the analogue of the reference's mkldnn JoinTable fusion, S/nn/mkldnn/Fusion.scala)."""
        from ..ops import functions as Fn

        tails = [self._tail_conv(b) for b in self.modules]
        if any(t is None for t in tails) or any(t.n_output_plane % 8 for t in tails):
            self._zero_copy = False
            return None
        key = tuple(input.shape)
        cache = self.__dict__.setdefault("_out_hw", {})
        if key not in cache:
            # first call with this input shape: run normally once to learn the output spatial size
            outs = [m._call(input) for m in self.modules]
            if len({tuple(o.shape[2:]) for o in outs}) != 1:
                self._zero_copy = False
            else:
                cache[key] = tuple(outs[0].shape[2:])
            return torch.cat(outs, dim=1)
        Ho, Wo = cache[key]
        ctot = sum(t.n_output_plane for t in tails)
        big = torch.empty((input.shape[0], Ho, Wo, ctot), dtype=torch.bfloat16, device=input.device)
        outs, c = [], 0
        multi = bool(config.get("bigdl.concat.streams")) and len(self.modules) > 1
        if multi:
            # The branches are independent and, in Inception, individually too small to fill 148 SMs: run them on side
            # streams (forked from / joined to the current stream, which also works under CUDA-graph capture). Autograd
            # replays each branch's backward on the stream its forward ran on.
            main = torch.cuda.current_stream(input.device)
            side = _side_streams(input.device, len(self.modules) - 1)
        # one fused gradient sum for the fan-out instead of autograd's chain of binary adds
        if input.requires_grad and len(self.modules) > 1:
            gf = input.grad_fn      # the previous zero-copy Concat, if this input is its fused-ReLU output
            prod = gf if (type(gf).__name__ == "ConcatViewFnBackward" and getattr(gf, "relu_all", False)) else None
            xs = Fn.FanOutFn.apply(input, len(self.modules), prod)
        else:
            xs = (input,) * len(self.modules)
        for i, (branch, tail) in enumerate(zip(self.modules, tails)):
            tail._next_out_hint = big[..., c:c + tail.n_output_plane]
            if multi and i > 0:
                st = side[i - 1]
                st.wait_stream(main)
                with torch.cuda.stream(st):
                    outs.append(branch._call(xs[i]))
            else:
                outs.append(branch._call(xs[i]))
            c += tail.n_output_plane
        if multi:
            for st in side:
                main.wait_stream(st)
        relu_all = all(bool(getattr(t, "_fuse_relu", False)) for t in tails)
        return Fn.ConcatViewFn.apply(big, relu_all, tails, *outs)


class ConcatTable(Container):
    def updateOutput(self, input):
        tail = self.__dict__.get("_res_tail")
        if (tail is not None and not self.isTraining() and isinstance(input, torch.Tensor) and input.is_cuda
                and not torch.is_grad_enabled()):       # the fused conv has no backward for the shortcut operand
            fused = self._residual_forward(input, tail)
            if fused is not None:
                return fused
        bn = self.__dict__.get("_bn_tail")
        if bn is not None and isinstance(input, torch.Tensor) and input.is_cuda:
            bn.__dict__["_defer_now"] = True      # consumed (popped) by that BN's next call, i.e. inside this forward
            pair = self.__dict__.get("_skip_pair")
            token = None
            if pair is not None and torch.is_grad_enabled() and input.requires_grad:
                from ..ops.functions import SkipToken

                token = SkipToken()               # the block input's two gradients meet in a dgrad epilogue
                for conv in pair:
                    if conv is not None:
                        conv.__dict__["_skip_token"] = token
            try:
                outs = [m._call(input) for m in self.modules]
            finally:
                bn.__dict__.pop("_defer_now", None)
                if pair is not None:
                    for conv in pair:
                        if conv is not None:
                            conv.__dict__.pop("_skip_token", None)
            if token is not None and pair[1] is None:         # identity shortcut: the fused bn + add holds the other end
                for o in outs:
                    if getattr(o, "_bdl_bn", None) is not None:
                        o._bdl_skip_token = token
            return Table(*outs)
        return Table(*[m._call(input) for m in self.modules])

    def _residual_forward(self, input, tail):
        """Inference residual block with the sum folded into the main branch's last convolution (fusion.py marks
        ``_res_tail`` / ``_res_main`` / ``_res_relu``): the shortcut runs first, the tail conv adds it (and applies the
        ReLU that follows the CAddTable) in its GEMM epilogue. Returns a one-element table whose tensor is tagged as
        already summed, or None when the convolution could not take the residual (then the plain path runs)."""
        from ..ops.functions import take_residual_hint

        main = self.__dict__["_res_main"]
        shortcut = self.modules[1 - main]._call(input)
        if not (isinstance(shortcut, torch.Tensor) and shortcut.dtype == torch.bfloat16):
            return None
        tail.__dict__["_next_residual"] = (shortcut, bool(self.__dict__.get("_res_relu", False)))
        y = self.modules[main]._call(input)
        tail.__dict__.pop("_next_residual", None)
        if take_residual_hint() is not None:          # the conv did not consume it: finish with an ordinary add
            return Table(y, shortcut) if main == 0 else Table(shortcut, y)
        y._bdl_sum_done = True
        return Table(y)


class ParallelTable(Container):
    def updateOutput(self, input):
        xs = _tl(input)
        return Table(*[m._call(x) for m, x in zip(self.modules, xs)])


class MapTable(Container):
    """Apply ONE module (shared weights) to every element of the input table (S/nn/MapTable.scala)."""

    def __init__(self, module: Module | None = None):
        super().__init__()
        if module is not None:
            self.add(module)

    def updateOutput(self, input):
        m = self.modules[0]
        return Table(*[m._call(x) for x in _tl(input)])


class Bottle(Container):
    """Collapse leading dims so an n_input_dim module sees a batch, then restore (S/nn/Bottle.scala)."""

    def __init__(self, module: Module, n_input_dim: int = 2, n_output_dim1: int = -1):
        super().__init__()
        self.n_input_dim = n_input_dim
        self.n_output_dim = n_input_dim if n_output_dim1 == -1 else n_output_dim1
        self.add(module)

    def updateOutput(self, input):
        if input.dim() <= self.n_input_dim:
            return self.modules[0]._call(input)
        lead = input.shape[:input.dim() - self.n_input_dim + 1]
        x = input.reshape(-1, *input.shape[input.dim() - self.n_input_dim + 1:])
        y = self.modules[0]._call(x)
        return y.reshape(*lead, *y.shape[1:])


# ---------------------------------------------------------------- table <-> tensor
class JoinTable(Module):
    def __init__(self, dimension: int, n_input_dims: int = -1):
        super().__init__()
        self.dimension, self.n_input_dims = dimension, n_input_dims

    def updateOutput(self, input):
        xs = _tl(input)
        return torch.cat(xs, dim=_dim(self.dimension, xs[0].dim(), self.n_input_dims))


class SplitTable(Module):
    def __init__(self, dimension: int, n_input_dims: int = -1):
        super().__init__()
        self.dimension, self.n_input_dims = dimension, n_input_dims

    def updateOutput(self, input):
        d = _dim(self.dimension, input.dim(), self.n_input_dims)
        return Table(*torch.unbind(input, dim=d))


class BifurcateSplitTable(Module):
    def __init__(self, dimension: int):
        super().__init__()
        self.dimension = dimension

    def updateOutput(self, input):
        d = _dim(self.dimension, input.dim())
        half = input.shape[d] // 2
        return Table(input.narrow(d, 0, half), input.narrow(d, half, input.shape[d] - half))


class FlattenTable(Module):
    def updateOutput(self, input):
        return Table(*flatten_activity(input))


class NarrowTable(Module):
    def __init__(self, offset: int, length: int = 1):
        super().__init__()
        self.offset, self.length = offset, length

    def updateOutput(self, input):
        xs = _tl(input)
        n = self.length if self.length >= 0 else len(xs) - self.offset + 2 + self.length
        return Table(*xs[self.offset - 1:self.offset - 1 + n])


class SelectTable(Module):
    def __init__(self, index: int):
        super().__init__()
        self.index = index

    def updateOutput(self, input):
        xs = _tl(input)
        return xs[self.index - 1] if self.index > 0 else xs[self.index]


class Pack(Module):
    def __init__(self, dimension: int):
        super().__init__()
        self.dimension = dimension

    def updateOutput(self, input):
        xs = _tl(input) if not isinstance(input, torch.Tensor) else [input]
        return torch.stack(xs, dim=self.dimension - 1)


class _CReduce(Module):
    def __init__(self, inplace: bool = False):
        super().__init__()
        self.inplace = inplace

    def _op(self, a, b):  # pragma: no cover
        raise NotImplementedError

    def updateOutput(self, input):
        xs = _tl(input)
        y = xs[0]
        for x in xs[1:]:
            y = self._op(y, x)
        return y


class CAddTable(_CReduce):
    """Element-wise sum of a table of tensors; broadcasts a scalar/smaller operand like the reference
    (S/nn/CAddTable.scala). On GPU two same-layout bf16 operands go through the fused add(+ReLU) kernel."""

    _fuse_relu = False

    def updateOutput(self, input):
        xs = _tl(input)
        if len(xs) == 1 and getattr(xs[0], "_bdl_sum_done", False):
            return xs[0]          # a fused ConcatTable already added the shortcut (and applied the ReLU) in the conv
        tags = [getattr(x, "_bdl_bn", None) for x in xs]
        if any(t is not None for t in tags):
            from .layers_core import bn_add_relu

            k = 0 if tags[0] is not None else 1
            if (len(xs) == 2 and tags[1 - k] is None and xs[1 - k].dtype == torch.bfloat16
                    and xs[0].shape == xs[1].shape):
                return bn_add_relu(tags[k], xs[k], xs[1 - k], self._fuse_relu, getattr(xs[k], "_bdl_skip_token", None))
            # not the shape the fused kernel takes: run the deferred normalisations now and add as usual
            xs = [x if t is None else t._call(x) for x, t in zip(xs, tags)]
        if (len(xs) == 2 and xs[0].is_cuda and xs[0].dtype == torch.bfloat16 and xs[1].dtype == torch.bfloat16
                and xs[0].shape == xs[1].shape):
            from .layers_core import native

            if native(xs[0]):
                from ..ops.functions import AddFn

                return AddFn.apply(xs[0], xs[1], self._fuse_relu)
        y = xs[0]
        for x in xs[1:]:
            y = y + x
        return F.relu(y) if self._fuse_relu else y


class CSubTable(_CReduce):
    def _op(self, a, b):
        return a - b


class CMulTable(_CReduce):
    def _op(self, a, b):
        return a * b


class CDivTable(_CReduce):
    def _op(self, a, b):
        return a / b


class CMaxTable(_CReduce):
    def _op(self, a, b):
        return torch.maximum(a, b)


class CMinTable(_CReduce):
    def _op(self, a, b):
        return torch.minimum(a, b)


class CAveTable(Module):
    def __init__(self, inplace: bool = False):
        super().__init__()

    def updateOutput(self, input):
        xs = _tl(input)
        return sum(xs[1:], xs[0]) / len(xs)


class MixtureTable(Module):
    """Mixture of experts: gater [B, n] weights experts given as a table of [B, ...] or a tensor [B, n, ...]
    (S/nn/MixtureTable.scala)."""

    def __init__(self, dim: int = 2 ** 31 - 1):
        super().__init__()
        self.dim = dim

    def updateOutput(self, input):
        gater, experts = _tl(input)
        if not isinstance(experts, torch.Tensor):
            experts = torch.stack(_tl(experts), dim=1)
        elif self.dim != 2 ** 31 - 1 and self.dim - 1 != 1:
            experts = experts.transpose(1, self.dim - 1)
        g = gater.reshape(*gater.shape, *([1] * (experts.dim() - gater.dim())))
        return (experts * g).sum(dim=1)


class TableOperation(Container):
    """Wrap a two-input table op so that a smaller second operand is expanded to the first's shape
    (S/nn/TableOperation.scala)."""

    def __init__(self, operation_layer: Module):
        super().__init__()
        self.add(operation_layer)

    def updateOutput(self, input):
        a, b = _tl(input)
        if a.shape != b.shape:
            small, big = (b, a) if b.numel() <= a.numel() else (a, b)
            small = small.expand_as(big)
            a, b = (big, small) if big is a else (small, big)
        return self.modules[0]._call(Table(a, b))


def CMulTableExpand():
    return TableOperation(CMulTable())


def CSubTableExpand():
    return TableOperation(CSubTable())


# ---------------------------------------------------------------- shape ops
class Reshape(Module):
    """Reshape the non-batch dims to ``size``; ``batch_mode`` None = infer (treat dim 0 as batch when the element
    count does not match) (S/nn/Reshape.scala)."""

    def __init__(self, size, batch_mode=None):
        super().__init__()
        self.size = list(size)
        self.batch_mode = batch_mode
        self.n_element = 1
        for s in self.size:
            self.n_element *= s

    def updateOutput(self, input):
        if self.batch_mode is False or (self.batch_mode is None and input.numel() == self.n_element
                                        and input.shape[0] != 1):
            return input.reshape(self.size)
        return input.reshape(input.shape[0], *self.size)


class InferReshape(Module):
    """Reshape where -1 infers a dim and 0 copies the input dim (S/nn/InferReshape.scala)."""

    def __init__(self, size, batch_mode: bool = False):
        super().__init__()
        self.size, self.batch_mode = list(size), batch_mode

    def updateOutput(self, input):
        shape = list(input.shape)
        off = 1 if self.batch_mode else 0
        out = [shape[i + off] if s == 0 else s for i, s in enumerate(self.size)]
        if self.batch_mode:
            out = [shape[0]] + out
        return input.reshape(out)


class View(Module):
    def __init__(self, *sizes, num_input_dims: int = 0):
        super().__init__()
        if len(sizes) == 1 and isinstance(sizes[0], (list, tuple)):
            sizes = tuple(sizes[0])
        self.sizes = list(sizes)
        self.num_input_dims = num_input_dims
        self.num_elements = 1
        for s in self.sizes:
            if s >= 0:
                self.num_elements *= s

    def setNumInputDims(self, n: int):
        self.num_input_dims = n
        return self

    set_num_input_dims = setNumInputDims

    def updateOutput(self, input):
        # View.scala batchSize(): with numInputDims set, one extra leading dim is the batch (even a batch of 1);
        # otherwise the batch is inferred from the element count
        if self.num_input_dims > 0 and input.dim() == self.num_input_dims + 1:
            batch = True
        elif -1 in self.sizes:
            batch = False
        else:
            batch = input.numel() != self.num_elements
        if batch:
            return input.reshape(input.shape[0], *self.sizes)
        return input.reshape(self.sizes)


class Squeeze(Module):
    def __init__(self, dim=None, num_input_dims: int = -2 ** 31, dims=None, batch_mode: bool = False):
        super().__init__()
        if dims is None and dim is not None:
            dims = list(dim) if isinstance(dim, (list, tuple)) else [dim]
        self.dims = dims
        self.batch_mode = batch_mode or (num_input_dims != -2 ** 31 and num_input_dims > 0)
        self.num_input_dims = num_input_dims

    def updateOutput(self, input):
        if self.dims is None:
            if self.batch_mode:
                y = input
                for d in range(input.dim() - 1, 0, -1):
                    if y.shape[d] == 1:
                        y = y.squeeze(d)
                return y
            return input.squeeze()
        y = input
        shift = 1 if (self.batch_mode and self.num_input_dims != -2 ** 31 and input.dim() == self.num_input_dims + 1) \
            or (self.batch_mode and self.num_input_dims == -2 ** 31) else 0
        for d in sorted([_dim(d, input.dim()) + shift for d in self.dims], reverse=True):
            if y.shape[d] == 1:
                y = y.squeeze(d)
        return y


class Unsqueeze(Module):
    def __init__(self, pos, num_input_dims: int = -2 ** 31):
        super().__init__()
        self.pos = list(pos) if isinstance(pos, (list, tuple)) else [pos]
        self.num_input_dims = num_input_dims

    def updateOutput(self, input):
        y = input
        for p in sorted(self.pos):
            shift = 1 if (self.num_input_dims != -2 ** 31 and input.dim() == self.num_input_dims + 1) else 0
            y = y.unsqueeze(p - 1 + shift)
        return y


class Transpose(Module):
    def __init__(self, permutations):
        super().__init__()
        self.permutations = [tuple(p) for p in permutations]

    def updateOutput(self, input):
        y = input
        for a, b in self.permutations:
            y = y.transpose(a - 1, b - 1)
        return y


class Contiguous(Module):
    def updateOutput(self, input):
        return input.contiguous()


class Narrow(Module):
    def __init__(self, dimension: int, offset: int, length: int = 1):
        super().__init__()
        self.dimension, self.offset, self.length = dimension, offset, length

    def updateOutput(self, input):
        d = _dim(self.dimension, input.dim())
        n = self.length if self.length >= 0 else input.shape[d] - self.offset + 2 + self.length
        return input.narrow(d, self.offset - 1, n)


class Select(Module):
    def __init__(self, dimension: int, index: int):
        super().__init__()
        self.dimension, self.index = dimension, index

    def updateOutput(self, input):
        d = _dim(self.dimension, input.dim())
        i = self.index - 1 if self.index > 0 else input.shape[d] + self.index
        return input.select(d, i)


class Index(Module):
    def __init__(self, dimension: int):
        super().__init__()
        self.dimension = dimension

    def updateOutput(self, input):
        t, idx = _tl(input)
        return t.index_select(self.dimension - 1, idx.long().reshape(-1) - 1)


class MaskedSelect(Module):
    def updateOutput(self, input):
        t, mask = _tl(input)
        return torch.masked_select(t, mask.bool())


class Masking(Module):
    """Zero every timestep whose features all equal ``mask_value`` (S/nn/Masking.scala)."""

    def __init__(self, mask_value: float = 0.0):
        super().__init__()
        self.mask_value = mask_value

    def updateOutput(self, input):
        keep = (input != self.mask_value).any(dim=-1, keepdim=True)
        return input * keep.to(input.dtype)


class Replicate(Module):
    def __init__(self, n_features: int, dim: int = 1, n_dim: int = 2 ** 31 - 1):
        super().__init__()
        self.n_features, self.dim, self.n_dim = n_features, dim, n_dim

    def updateOutput(self, input):
        d = self.dim - 1 + (1 if (self.n_dim != 2 ** 31 - 1 and input.dim() > self.n_dim) else 0)
        shape = list(input.shape)
        shape.insert(d, self.n_features)
        return input.unsqueeze(d).expand(shape)


class Tile(Module):
    def __init__(self, dim: int = 1, copies: int = 2):
        super().__init__()
        self.dim, self.copies = dim, copies

    def updateOutput(self, input):
        reps = [1] * input.dim()
        reps[self.dim - 1] = self.copies
        return input.repeat(*reps)


class ExpandSize(Module):
    def __init__(self, sizes):
        super().__init__()
        self.sizes = list(sizes)

    def updateOutput(self, input):
        tgt = [s if s != -1 else input.shape[i] for i, s in enumerate(self.sizes)]
        return input.expand(tgt)


class Reverse(Module):
    def __init__(self, dimension: int = 1, is_inplace: bool = False):
        super().__init__()
        self.dimension = dimension

    def updateOutput(self, input):
        return torch.flip(input, dims=[self.dimension - 1])


class Padding(Module):
    """Pad ``pad`` entries (negative = before, positive = after) along ``dim`` with ``value`` (S/nn/Padding.scala)."""

    def __init__(self, dim: int, pad: int, n_input_dim: int, value: float = 0.0, n_index: int = 1):
        super().__init__()
        self.dim, self.pad, self.n_input_dim, self.value, self.n_index = dim, pad, n_input_dim, value, n_index

    def updateOutput(self, input):
        d = self.dim - 1 + (1 if input.dim() != self.n_input_dim else 0)
        shape = list(input.shape)
        shape[d] = abs(self.pad)
        filler = torch.full(shape, self.value, dtype=input.dtype, device=input.device)
        if self.pad < 0:
            idx = self.n_index - 1
            return torch.cat([input.narrow(d, 0, idx), filler, input.narrow(d, idx, input.shape[d] - idx)], dim=d)
        idx = input.shape[d] - (self.n_index - 1)
        return torch.cat([input.narrow(d, 0, idx), filler, input.narrow(d, idx, input.shape[d] - idx)], dim=d)


class SpatialZeroPadding(Module):
    def __init__(self, pad_left: int, pad_right: int | None = None, pad_top: int | None = None,
                 pad_bottom: int | None = None):
        super().__init__()
        self.pl = pad_left
        self.pr = pad_left if pad_right is None else pad_right
        self.pt = pad_left if pad_top is None else pad_top
        self.pb = pad_left if pad_bottom is None else pad_bottom

    def updateOutput(self, input):
        return F.pad(input, (self.pl, self.pr, self.pt, self.pb))


class Cropping2D(Module):
    def __init__(self, heightCrop, widthCrop, data_format: str = "NCHW"):
        super().__init__()
        self.hc, self.wc, self.data_format = list(heightCrop), list(widthCrop), data_format.upper()

    def updateOutput(self, input):
        hd, wd = (2, 3) if self.data_format == "NCHW" else (1, 2)
        y = input.narrow(hd, self.hc[0], input.shape[hd] - self.hc[0] - self.hc[1])
        return y.narrow(wd, self.wc[0], input.shape[wd] - self.wc[0] - self.wc[1])


class Cropping3D(Module):
    def __init__(self, dim1Crop, dim2Crop, dim3Crop, data_format: str = "channel_first"):
        super().__init__()
        self.crops = [list(dim1Crop), list(dim2Crop), list(dim3Crop)]
        self.first = data_format.lower() in ("channel_first", "ncdhw")

    def updateOutput(self, input):
        base = 2 if self.first else 1
        y = input
        for i, (a, b) in enumerate(self.crops):
            y = y.narrow(base + i, a, y.shape[base + i] - a - b)
        return y


# ---------------------------------------------------------------- reductions
class Sum(Module):
    def __init__(self, dimension: int = 1, n_input_dims: int = -1, size_average: bool = False, squeeze: bool = True):
        super().__init__()
        self.dimension, self.n_input_dims, self.size_average, self.squeeze = dimension, n_input_dims, size_average, squeeze

    def updateOutput(self, input):
        d = _dim(self.dimension, input.dim(), self.n_input_dims)
        y = input.sum(dim=d, keepdim=not self.squeeze)
        return y / input.shape[d] if self.size_average else y


class Mean(Sum):
    def __init__(self, dimension: int = 1, n_input_dims: int = -1, squeeze: bool = True):
        super().__init__(dimension, n_input_dims, True, squeeze)


class Max(Module):
    def __init__(self, dim: int = 1, num_input_dims: int = -2 ** 31):
        super().__init__()
        self.dim, self.num_input_dims = dim, num_input_dims

    def updateOutput(self, input):
        d = _dim(self.dim, input.dim(), self.num_input_dims if self.num_input_dims > 0 else -1)
        return input.max(dim=d).values


class Min(Max):
    def updateOutput(self, input):
        d = _dim(self.dim, input.dim(), self.num_input_dims if self.num_input_dims > 0 else -1)
        return input.min(dim=d).values


# ---------------------------------------------------------------- pairwise products / distances
class DotProduct(Module):
    def updateOutput(self, input):
        a, b = _tl(input)
        if a.dim() == 1:
            return (a * b).sum().reshape(1)
        return (a * b).sum(dim=1)


class CrossProduct(Module):
    """Pairwise dot products of all table elements (S/nn/CrossProduct.scala)."""

    def __init__(self, num_tensor: int = 0, embedding_size: int = 0):
        super().__init__()
        self.num_tensor, self.embedding_size = num_tensor, embedding_size

    def updateOutput(self, input):
        xs = [x if x.dim() > 1 else x.unsqueeze(0) for x in _tl(input)]
        outs = []
        for i in range(len(xs)):
            for j in range(i + 1, len(xs)):
                outs.append((xs[i] * xs[j]).sum(dim=1, keepdim=True))
        return torch.cat(outs, dim=1)


class PairwiseDistance(Module):
    def __init__(self, norm: int = 2):
        super().__init__()
        self.norm = norm

    def updateOutput(self, input):
        a, b = _tl(input)
        d = (a - b).abs()
        if a.dim() == 1:
            return d.pow(self.norm).sum().pow(1.0 / self.norm).reshape(1)
        return d.pow(self.norm).sum(dim=1).pow(1.0 / self.norm)


class CosineDistance(Module):
    def updateOutput(self, input):
        a, b = _tl(input)
        if a.dim() == 1:
            a, b = a.unsqueeze(0), b.unsqueeze(0)
        return F.cosine_similarity(a, b, dim=1, eps=1e-12)


class MM(Module):
    def __init__(self, trans_a: bool = False, trans_b: bool = False):
        super().__init__()
        self.trans_a, self.trans_b = trans_a, trans_b

    def updateOutput(self, input):
        a, b = _tl(input)
        if self.trans_a:
            a = a.transpose(-1, -2)
        if self.trans_b:
            b = b.transpose(-1, -2)
        return torch.matmul(a, b)


class MV(Module):
    def __init__(self, trans: bool = False):
        super().__init__()
        self.trans = trans

    def updateOutput(self, input):
        m, v = _tl(input)
        if self.trans:
            m = m.transpose(-1, -2)
        return torch.matmul(m, v.unsqueeze(-1)).squeeze(-1)
