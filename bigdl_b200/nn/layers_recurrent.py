"""Recurrent layers: Recurrent / BiRecurrent / RecurrentDecoder containers over Cells (RnnCell, LSTM, LSTMPeephole, GRU,
ConvLSTMPeephole, ConvLSTMPeephole3D, MultiRNNCell), TimeDistributed, and tree LSTMs.

Reference: S/nn/Recurrent.scala:243-527 (time-unrolled loop over cell clones that share weights), Cell.scala,
RNN.scala (RnnCell), LSTM.scala:44-175, LSTMPeephole.scala, GRU.scala:72-160, ConvLSTMPeephole.scala,
ConvLSTMPeephole3D.scala, MultiRNNCell.scala, BiRecurrent.scala, RecurrentDecoder.scala, TimeDistributed.scala,
TreeLSTM.scala, BinaryTreeLSTM.scala; MKL-DNN fused twin S/nn/mkldnn/RNN.scala.

Design: like the reference's ``preTopology`` trick, the input projection of ALL time steps is one GEMM
(``[B*T, in] x [in, G*H]``) before the time loop; each step then does one recurrent GEMM plus a fused gate kernel.
Gate order (chunks of the 4H axis) follows the reference: LSTM = [input, cell-candidate, forget, output],
GRU = [reset, update, candidate]. Input layout is batch-first ``[B, T, F]``.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from ..utils import config
from ..utils.random_generator import RNG
from ..utils.table import Table, to_table
from .initialization_method import RandomUniform, VariableFormat
from .layers_core import Dropout, Linear
from .module import Container, Module


class Cell(Module):
    """One time step: ``step(x_proj_t, hidden) -> (output_t, new_hidden)``; ``pre(x)`` projects the whole sequence."""

    hidden_count = 1

    def pre(self, x):
        return x

    # ---- reference-named pieces (Cell.scala:85-240)
    @property
    def hiddenSizeOfPreTopo(self) -> int:
        """Width of the projected input one step consumes (4·hidden for an LSTM, 3·hidden for a GRU, …)."""
        with torch.no_grad():
            return int(self.pre(torch.zeros(1, 1, self.input_size)).shape[-1])

    def hidResize(self, hidden, batchSize: int, stepShape=None):
        """Zero hidden state for ``batchSize`` rows (a tensor, or a Table when the cell carries several)."""
        hs = self.init_hidden(batchSize, torch.zeros(1), tuple(stepShape) if stepShape else None)
        return hs[0] if len(hs) == 1 else Table(*hs)

    def regluarized(self, isRegularized: bool):
        """(sic) switch every regulariser of the cell's layers on or off."""
        for m in self.all_modules():
            for name in ("wRegularizer", "bRegularizer", "uRegularizer"):
                r = getattr(m, name, None)
                if r is not None:
                    (r.enable if isRegularized else r.disable)()

    regularized = regluarized

    def init_hidden(self, batch: int, like: torch.Tensor, spatial=None):
        raise NotImplementedError

    def step(self, xt, hidden):
        raise NotImplementedError

    def updateOutput(self, input):
        """Cell as a stand-alone module: input = table(x_t, hidden...) → table(out, hidden...) (reference Cell API)."""
        xs = to_table(input).to_list()
        x, hidden = xs[0], xs[1:]
        if len(hidden) == 1 and isinstance(hidden[0], Table):
            hidden = hidden[0].to_list()
        out, new_h = self.step(self.pre(x.unsqueeze(1)).squeeze(1), list(hidden))
        return Table(out, *new_h)


def _lin_init(mod: Linear, fan_in: int):
    stdv = 1.0 / math.sqrt(fan_in)
    RNG.fill_uniform(mod.weight.data, -stdv, stdv)
    if mod.with_bias:
        RNG.fill_uniform(mod.bias.data, -stdv, stdv)


class RnnCell(Cell):
    """h_t = act(W x_t + b + U h_{t-1}) (S/nn/RNN.scala)."""

    def __init__(self, input_size, hidden_size, activation=None, isInputWithBias=True, isHiddenWithBias=True,
                 wRegularizer=None, uRegularizer=None, bRegularizer=None):
        super().__init__()
        from .layers_activation import Tanh

        self.input_size, self.hidden_size = input_size, hidden_size
        self.activation = activation or Tanh()
        self.i2h = Linear(input_size, hidden_size, isInputWithBias, wRegularizer, bRegularizer)
        self.h2h = Linear(hidden_size, hidden_size, isHiddenWithBias, uRegularizer)
        self.modules = [self.i2h, self.h2h, self.activation]

    def pre(self, x):
        return self.i2h._call(x)

    def init_hidden(self, batch, like, spatial=None):
        return [like.new_zeros(batch, self.hidden_size)]

    def step(self, xt, hidden):
        h = self.activation._call(xt + self.h2h._call(hidden[0]))
        return h, [h]


class LSTM(Cell):
    """LSTM cell, gate chunks [input, candidate, forget, output] (S/nn/LSTM.scala:44-175)."""

    hidden_count = 2

    def __init__(self, input_size, hidden_size, p: float = 0.0, activation=None, inner_activation=None,
                 wRegularizer=None, uRegularizer=None, bRegularizer=None):
        super().__init__()
        self.input_size, self.hidden_size, self.p = input_size, hidden_size, p
        self.activation, self.inner_activation = activation, inner_activation
        self.i2g = Linear(input_size, 4 * hidden_size, True, wRegularizer, bRegularizer)
        self.h2g = Linear(hidden_size, 4 * hidden_size, False, uRegularizer)
        self.drop = Dropout(p) if p else None
        # the dropout is a child too: evaluate() / training() must reach it (it is parameter-free, so the order of the
        # stored tensors does not change)
        self.modules = [self.i2g, self.h2g] + ([activation] if activation else []) + \
                       ([inner_activation] if inner_activation else []) + ([self.drop] if self.drop is not None else [])

    def pre(self, x):
        if self.drop is not None:
            x = self.drop._call(x)
        return self.i2g._call(x)

    def init_hidden(self, batch, like, spatial=None):
        return [like.new_zeros(batch, self.hidden_size), like.new_zeros(batch, self.hidden_size)]

    def _act(self, v):
        return torch.tanh(v) if self.activation is None else self.activation._call(v)

    def _inner(self, v):
        return torch.sigmoid(v) if self.inner_activation is None else self.inner_activation._call(v)

    def step(self, xt, hidden):
        h, c = hidden
        g = xt + self.h2g._call(h)
        if g.is_cuda and self.activation is None and self.inner_activation is None and g.dtype == torch.bfloat16:
            from ..ops.recurrent import lstm_gates

            h2, c2 = lstm_gates(g, c)
            return h2, [h2, c2]
        H = self.hidden_size
        i, z, f, o = self._inner(g[:, :H]), self._act(g[:, H:2 * H]), self._inner(g[:, 2 * H:3 * H]), self._inner(g[:, 3 * H:])
        c2 = i * z + f * c
        h2 = self._act(c2) * o
        return h2, [h2, c2]


class LSTMPeephole(LSTM):
    """LSTM whose gates also see the cell state through diagonal (CMul) peepholes (S/nn/LSTMPeephole.scala);
    gate chunks [input, forget, candidate, output] as in the reference."""

    def __init__(self, input_size, hidden_size, p: float = 0.0, wRegularizer=None, uRegularizer=None, bRegularizer=None):
        super().__init__(input_size, hidden_size, p, None, None, wRegularizer, uRegularizer, bRegularizer)
        for n in ("peepI", "peepF", "peepO"):
            self.register_parameter(n, torch.zeros(hidden_size))
            RNG.fill_uniform(self._parameters[n].data, -1.0 / math.sqrt(hidden_size), 1.0 / math.sqrt(hidden_size))

    def step(self, xt, hidden):
        h, c = hidden
        H = self.hidden_size
        g = xt + self.h2g._call(h)
        i = torch.sigmoid(g[:, :H] + c * self.peepI.to(g.dtype))
        f = torch.sigmoid(g[:, H:2 * H] + c * self.peepF.to(g.dtype))
        z = torch.tanh(g[:, 2 * H:3 * H])
        c2 = i * z + f * c
        o = torch.sigmoid(g[:, 3 * H:] + c2 * self.peepO.to(g.dtype))
        h2 = torch.tanh(c2) * o
        return h2, [h2, c2]


class GRU(Cell):
    """GRU cell, chunks [reset, update, candidate]; h' = (1-z)*h_hat + z*h (S/nn/GRU.scala:72-160)."""

    def __init__(self, input_size, output_size, p: float = 0.0, activation=None, inner_activation=None,
                 wRegularizer=None, uRegularizer=None, bRegularizer=None):
        super().__init__()
        self.input_size, self.hidden_size, self.p = input_size, output_size, p
        self.activation, self.inner_activation = activation, inner_activation
        self.i2g = Linear(input_size, 3 * output_size, True, wRegularizer, bRegularizer)
        self.h2g = Linear(output_size, 2 * output_size, False, uRegularizer)
        self.h2c = Linear(output_size, output_size, False, uRegularizer)
        self.drop = Dropout(p) if p else None
        self.modules = [self.i2g, self.h2g, self.h2c] + ([self.drop] if self.drop is not None else [])

    def pre(self, x):
        if self.drop is not None:
            x = self.drop._call(x)
        return self.i2g._call(x)

    def init_hidden(self, batch, like, spatial=None):
        return [like.new_zeros(batch, self.hidden_size)]

    def step(self, xt, hidden):
        h = hidden[0]
        H = self.hidden_size
        if xt.is_cuda and self.activation is None and self.inner_activation is None and xt.dtype == torch.bfloat16:
            # two fused pointwise kernels around the two recurrent GEMMs (csrc/recurrent.cu)
            from ..ops.recurrent import gru_out, gru_rz

            hb = h if h.dtype == torch.bfloat16 else h.to(torch.bfloat16)
            z, hr = gru_rz(xt[:, :2 * H], self.h2g._call(hb), hb)
            h2 = gru_out(xt[:, 2 * H:], self.h2c._call(hr), z, hb)
            return h2, [h2]
        inner = torch.sigmoid if self.inner_activation is None else self.inner_activation._call
        act = torch.tanh if self.activation is None else self.activation._call
        rz = inner(xt[:, :2 * H] + self.h2g._call(h))
        r, z = rz[:, :H], rz[:, H:]
        h_hat = act(xt[:, 2 * H:] + self.h2c._call(h * r))
        h2 = (1 - z) * h_hat + z * h
        return h2, [h2]


class ConvLSTMPeephole(Cell):
    """Convolutional LSTM with peepholes over [B, C, H, W] frames (S/nn/ConvLSTMPeephole.scala)."""

    hidden_count = 2
    ndim = 2

    def __init__(self, input_size, output_size, kernel_i, kernel_c, stride=1, padding=-1, activation=None,
                 inner_activation=None, wRegularizer=None, uRegularizer=None, bRegularizer=None, cRegularizer=None,
                 with_peephole=True):
        super().__init__()
        self.input_size, self.hidden_size = input_size, output_size
        self.kernel_i, self.kernel_c, self.stride, self.with_peephole = kernel_i, kernel_c, stride, with_peephole
        shape_i = (4 * output_size, input_size) + (kernel_i,) * self.ndim
        shape_c = (4 * output_size, output_size) + (kernel_c,) * self.ndim
        self.register_parameter("wi", torch.zeros(shape_i))
        self.register_parameter("bi", torch.zeros(4 * output_size))
        self.register_parameter("wc", torch.zeros(shape_c))
        if with_peephole:
            for n in ("peepI", "peepF", "peepO"):
                self.register_parameter(n, torch.zeros(output_size))
        self.reset()

    def reset(self):
        for n, p in self._parameters.items():
            fan = p[0].numel() if p.dim() > 1 else p.numel()
            RNG.fill_uniform(p.data, -1.0 / math.sqrt(fan), 1.0 / math.sqrt(fan))
        self.zeroGradParameters()
        return self

    def _conv(self, x, w, b, k, stride):
        fn = F.conv2d if self.ndim == 2 else F.conv3d
        return fn(x, w.to(x.dtype), None if b is None else b.to(x.dtype), stride, k // 2)

    def pre(self, x):   # x [B, T, C, *spatial]
        B, T = x.shape[:2]
        y = self._conv(x.reshape(B * T, *x.shape[2:]), self.wi, self.bi, self.kernel_i, self.stride)
        return y.reshape(B, T, *y.shape[1:])

    def init_hidden(self, batch, like, spatial=None):
        z = like.new_zeros(batch, self.hidden_size, *spatial)
        return [z, z.clone()]

    def step(self, xt, hidden):
        h, c = hidden
        H = self.hidden_size
        g = xt + self._conv(h, self.wc, None, self.kernel_c, 1)
        sh = (1, -1) + (1,) * self.ndim
        pi = self.peepI.reshape(sh).to(g.dtype) * c if self.with_peephole else 0
        pf = self.peepF.reshape(sh).to(g.dtype) * c if self.with_peephole else 0
        i = torch.sigmoid(g[:, :H] + pi)
        f = torch.sigmoid(g[:, H:2 * H] + pf)
        z = torch.tanh(g[:, 2 * H:3 * H])
        c2 = i * z + f * c
        po = self.peepO.reshape(sh).to(g.dtype) * c2 if self.with_peephole else 0
        o = torch.sigmoid(g[:, 3 * H:] + po)
        h2 = torch.tanh(c2) * o
        return h2, [h2, c2]


class ConvLSTMPeephole3D(ConvLSTMPeephole):
    ndim = 3


class MultiRNNCell(Cell):
    """Stack of cells applied within one time step (S/nn/MultiRNNCell.scala)."""

    def __init__(self, cells):
        super().__init__()
        self.cells = list(cells)
        self.modules = list(self.cells)
        self.hidden_count = sum(c.hidden_count for c in self.cells)

    def pre(self, x):
        return x

    def init_hidden(self, batch, like, spatial=None):
        out = []
        for c in self.cells:
            out.extend(c.init_hidden(batch, like, spatial))
        return out

    def step(self, xt, hidden):
        new_hidden, x, k = [], xt, 0
        for c in self.cells:
            h = hidden[k:k + c.hidden_count]
            k += c.hidden_count
            x, nh = c.step(c.pre(x.unsqueeze(1)).squeeze(1), h)
            new_hidden.extend(nh)
        return x, new_hidden


class BatchNormParams:
    """Settings of the batch normalisation ``Recurrent(batchNormParams = …)`` puts between the input projection and
    the recurrence (S/nn/Recurrent.scala:850-857)."""

    def __init__(self, eps: float = 1e-5, momentum: float = 0.1, initWeight=None, initBias=None, initGradWeight=None,
                 initGradBias=None, affine: bool = True, **kw):
        self.eps, self.momentum, self.affine = eps, momentum, affine
        self.initWeight = kw.get("init_weight", initWeight)
        self.initBias = kw.get("init_bias", initBias)
        self.initGradWeight = kw.get("init_grad_weight", initGradWeight)
        self.initGradBias = kw.get("init_grad_bias", initGradBias)


class Recurrent(Container):
    """Run a Cell over the time dimension of ``[B, T, ...]`` (S/nn/Recurrent.scala:243-527). ``setHiddenState`` /
    ``getHiddenState`` expose the carried state; ``mask_zero`` stops state updates on all-zero (padded) steps.
    With ``batch_norm_params`` the projected input of ALL time steps ([B·T, gates]) goes through one
    BatchNormalization before the recurrence (Recurrent.scala:106-140)."""

    def __init__(self, batch_norm_params=None, mask_zero: bool = False):
        super().__init__()
        if batch_norm_params is True:
            batch_norm_params = BatchNormParams()
        elif isinstance(batch_norm_params, dict):
            batch_norm_params = BatchNormParams(**batch_norm_params)
        self.batch_norm_params, self.mask_zero = batch_norm_params, mask_zero
        self.bnorm = None
        self._init_hidden = None
        self._last_hidden = None

    def add(self, module):
        if isinstance(module, TimeDistributed) and self.modules:
            # a saved model lists [cell, TimeDistributed(BatchNormalization)]: the loader hands the second one back here
            bn = module.modules[0]
            self.bnorm = module
            self.batch_norm_params = BatchNormParams(bn.eps, bn.momentum, affine=bn.affine)
            self.modules = [self.modules[0], module]
            return self
        if not isinstance(module, Cell):
            raise TypeError("Recurrent: added module should be Cell type")
        self.modules = [module]
        self.bnorm = None
        bp = self.batch_norm_params
        if bp is not None:
            if type(module).pre is Cell.pre:
                raise ValueError(f"{module.getName()} does not support BatchNormalization: it has no input projection")
            from .layers_core import BatchNormalization

            with torch.no_grad():
                width = module.pre(torch.zeros(1, 1, module.input_size)).shape[-1]
            self.bnorm = TimeDistributed(BatchNormalization(width, bp.eps, bp.momentum, bp.affine, bp.initWeight,
                                                            bp.initBias, bp.initGradWeight, bp.initGradBias))
            self.modules = [module, self.bnorm]
        return self

    def _project(self, cell, x):
        proj = cell.pre(x)
        return self.bnorm._call(proj) if self.bnorm is not None else proj

    @property
    def cell(self) -> Cell:
        return self.modules[0]

    def getCell(self) -> Cell:
        return self.modules[0]

    def findDropouts(self):
        """Every Dropout inside the cell (the reference re-seeds them per time step; here one mask per call)."""
        from .layers_core import Dropout

        return [m for m in self.cell.all_modules() if isinstance(m, Dropout)]

    def setHiddenState(self, hidden):
        self._init_hidden = to_table(hidden).to_list() if not isinstance(hidden, torch.Tensor) else [hidden]
        return self

    set_hidden_state = setHiddenState

    def getHiddenState(self):
        h = self._last_hidden
        return h[0] if h is not None and len(h) == 1 else (Table(*h) if h else None)

    get_hidden_state = getHiddenState

    def _fused_lstm(self, cell, x, reverse: bool = False):
        """The whole sequence in one persistent kernel (csrc/lstm_seq.cu) when the cell is a plain LSTM on the GPU."""
        from ..ops.recurrent import LstmSeqFn, lstm_seq_supported

        if not (type(cell) is LSTM and cell.activation is None and cell.inner_activation is None and not self.mask_zero
                and x.is_cuda and x.dim() == 3 and config.get("bigdl.engineType") == "b200"
                and bool(config.get("bigdl.rnn.persistent"))
                and lstm_seq_supported(cell.hidden_size, x.device)):
            return None
        proj = self._project(cell, x)
        if proj.dtype != torch.bfloat16:
            return None
        h0 = c0 = None
        if self._init_hidden is not None:
            h0, c0 = self._init_hidden[0], self._init_hidden[1]
        w = cell.h2g.weight
        w16 = cell.h2g.weight_for_compute("weight")
        if w16.dtype != torch.bfloat16:
            w16 = w16.detach().to(torch.bfloat16)
        hs, h_last, c_last = LstmSeqFn.apply(proj, w, w16, h0, c0, reverse)
        self._last_hidden = [h_last.detach(), c_last.detach()]
        return hs

    def updateOutput(self, input):
        cell = self.cell
        x = input
        fused = self._fused_lstm(cell, x)
        if fused is not None:
            return fused
        B, T = x.shape[0], x.shape[1]
        proj = self._project(cell, x)
        spatial = tuple(proj.shape[3:]) if proj.dim() > 3 else None
        hidden = self._init_hidden if self._init_hidden is not None else cell.init_hidden(B, proj, spatial)
        hidden = [h.to(proj.dtype) for h in hidden]
        outs = []
        for t in range(T):
            out, new_hidden = cell.step(proj[:, t], hidden)
            if self.mask_zero:
                keep = (x[:, t].reshape(B, -1) != 0).any(dim=1)
                shape = (B,) + (1,) * (out.dim() - 1)
                k = keep.reshape(shape).to(out.dtype)
                out = out * k
                new_hidden = [nh * k + h * (1 - k) for nh, h in zip(new_hidden, hidden)]
            hidden = new_hidden
            outs.append(out)
        self._last_hidden = [h.detach() for h in hidden]
        return torch.stack(outs, dim=1)


class BiRecurrent(Container):
    """Forward and time-reversed Recurrent over the same input, merged by ``merge`` (default CAddTable), or
    concatenated when ``is_split_input`` is false and merge is a JoinTable (S/nn/BiRecurrent.scala)."""

    def __init__(self, merge: Module | None = None, batch_norm_params=None, is_split_input: bool = False):
        super().__init__()
        from .layers_table import CAddTable

        self.merge = merge or CAddTable()
        self.is_split_input = is_split_input
        self.fwd, self.bwd = Recurrent(batch_norm_params), Recurrent(batch_norm_params)
        self.modules = [self.fwd, self.bwd, self.merge]

    def add(self, cell: Cell):
        self.fwd.add(cell)
        c2 = cell.cloneModule()
        c2.reset() if hasattr(c2, "reset") else None
        self.bwd.add(c2)
        return self

    # serializer hooks: the two cells are the only state that the constructor cannot rebuild
    def _ser_children(self):
        return [self.fwd.cell, self.bwd.cell] if self.fwd.modules else []

    def _deser_children(self, mods):
        if len(mods) == 2:
            self.fwd.add(mods[0])
            self.bwd.add(mods[1])

    def updateOutput(self, input):
        if self.is_split_input:
            half = input.shape[-1] // 2
            xf, xb = input[..., :half], input[..., half:]
        else:
            xf = xb = input
        yf = self.fwd._call(xf)
        yb = torch.flip(self.bwd._call(torch.flip(xb, dims=[1])), dims=[1])
        return self.merge._call(Table(yf, yb))


class RecurrentDecoder(Recurrent):
    """Feed each step's output back as the next step's input for ``output_length`` steps
    (S/nn/RecurrentDecoder.scala). Input = the first step's input [B, F]."""

    def __init__(self, output_length: int):
        super().__init__()
        self.output_length = output_length

    def updateOutput(self, input):
        cell = self.cell
        x = input
        B = x.shape[0]
        first = cell.pre(x.unsqueeze(1))
        spatial = tuple(first.shape[3:]) if first.dim() > 3 else None
        hidden = self._init_hidden if self._init_hidden is not None else cell.init_hidden(B, first, spatial)
        outs = []
        cur = x
        for _ in range(self.output_length):
            out, hidden = cell.step(cell.pre(cur.unsqueeze(1)).squeeze(1), hidden)
            outs.append(out)
            cur = out
        self._last_hidden = [h.detach() for h in hidden]
        return torch.stack(outs, dim=1)


class TimeDistributed(Container):
    """Apply a module to every time step by folding time into batch (S/nn/TimeDistributed.scala)."""

    def __init__(self, layer: Module, mask_zero: bool = False):
        super().__init__()
        self.mask_zero = mask_zero
        self.add(layer)

    def updateOutput(self, input):
        B, T = input.shape[0], input.shape[1]
        y = self.modules[0]._call(input.reshape(B * T, *input.shape[2:]))
        y = y.reshape(B, T, *y.shape[1:])
        if self.mask_zero:
            keep = (input.reshape(B, T, -1) != 0).any(dim=2)
            y = y * keep.reshape((B, T) + (1,) * (y.dim() - 2)).to(y.dtype)
        return y


class TreeLSTM(Module):
    """Base class of tree-structured LSTMs (S/nn/TreeLSTM.scala)."""

    def __init__(self, input_size: int, hidden_size: int):
        super().__init__()
        self.input_size, self.hidden_size = input_size, hidden_size


class BinaryTreeLSTM(TreeLSTM):
    """Constituency tree LSTM (Tai et al.) over binary trees (S/nn/BinaryTreeLSTM.scala). Input =
    table(embeddings [B, nWords, in], trees [B, nNodes, 3]) where each tree row is (left child, right child, leaf word
    index), 1-based, 0 = none, and the root is marked by -1 in the third column of its row in the reference encoding;
    output [B, nNodes, hidden]."""

    def __init__(self, input_size, hidden_size, gate_output: bool = True, with_graph: bool = True):
        super().__init__(input_size, hidden_size)
        self.gate_output = gate_output
        self.leaf_c = Linear(input_size, hidden_size)
        self.leaf_o = Linear(input_size, hidden_size) if gate_output else None
        self.comp_l = Linear(hidden_size, 5 * hidden_size if gate_output else 4 * hidden_size)
        self.comp_r = Linear(hidden_size, 5 * hidden_size if gate_output else 4 * hidden_size, False)
        self.modules = [m for m in (self.leaf_c, self.leaf_o, self.comp_l, self.comp_r) if m is not None]

    def _leaf(self, x):
        c = self.leaf_c._call(x)
        if self.gate_output:
            o = torch.sigmoid(self.leaf_o._call(x))
            return o * torch.tanh(c), c
        return torch.tanh(c), c

    def _compose(self, lh, lc, rh, rc):
        H = self.hidden_size
        g = self.comp_l._call(lh) + self.comp_r._call(rh)
        i, lf, rf, u = torch.sigmoid(g[..., :H]), torch.sigmoid(g[..., H:2 * H]), torch.sigmoid(g[..., 2 * H:3 * H]), \
            torch.tanh(g[..., 3 * H:4 * H])
        c = i * u + lf * lc + rf * rc
        if self.gate_output:
            o = torch.sigmoid(g[..., 4 * H:])
            return o * torch.tanh(c), c
        return torch.tanh(c), c

    def updateOutput(self, input):
        emb, trees = to_table(input).to_list()
        B, n_nodes = trees.shape[0], trees.shape[1]
        outs = []
        for b in range(B):
            tree = trees[b].long()
            memo = {}

            def rec(node):   # node is 1-based
                if node in memo:
                    return memo[node]
                l, r, w = int(tree[node - 1, 0]), int(tree[node - 1, 1]), int(tree[node - 1, 2])
                if l == 0 and r == 0:
                    res = self._leaf(emb[b, max(w, 1) - 1].unsqueeze(0))
                else:
                    lh, lc = rec(l)
                    rh, rc = rec(r)
                    res = self._compose(lh, lc, rh, rc)
                memo[node] = res
                return res

            hs = []
            for node in range(1, n_nodes + 1):
                l, r, w = int(tree[node - 1, 0]), int(tree[node - 1, 1]), int(tree[node - 1, 2])
                if l == 0 and r == 0 and w == 0:
                    hs.append(emb.new_zeros(1, self.hidden_size))      # padding row
                else:
                    hs.append(rec(node)[0])
            outs.append(torch.cat(hs, dim=0))
        return torch.stack(outs, dim=0)


class TensorTree:
    """A tree stored as a [nodes, width] tensor (S/nn/BinaryTreeLSTM.scala:513-572): row i holds the 1-based child
    indices of node i (0 = none, −1 in the first column = padding row) and, in the LAST column, the leaf's word index
    or −1 for the root. All indices are 1-based like the reference."""

    def __init__(self, content: torch.Tensor):
        if content.dim() != 2:
            raise ValueError(f"The content of TensorTree should be a two-dimensional tensor, content dim({content.dim()})")
        self.content = content

    @property
    def size(self):
        return list(self.content.shape)

    @property
    def nodeNumber(self) -> int:
        return self.content.shape[0]

    def children(self, index: int):
        return [int(v) for v in self.content[index - 1].tolist()]

    def addChild(self, parent: int, child):
        row = self.content[parent - 1]
        for i in range(row.numel() - 1):
            if float(row[i]) == 0:
                row[i] = child
                return

    def markAsRoot(self, index: int):
        self.content[index - 1, -1] = -1

    def getRoot(self) -> int:
        for i in range(self.content.shape[0]):
            if int(self.content[i, -1]) == -1:
                return i + 1
        raise RuntimeError("There is no root in the tensor tree")

    def markAsLeaf(self, index: int, leafIndex: int):
        self.content[index - 1, -1] = leafIndex

    def leafIndex(self, index: int) -> int:
        return int(self.content[index - 1, -1])

    def hasChild(self, index: int) -> bool:
        return int(self.content[index - 1, 0]) > 0

    def noChild(self, index: int) -> bool:
        return int(self.content[index - 1, 0]) == 0

    def exists(self, index: int) -> bool:
        return 1 <= index <= self.content.shape[0]

    def isPadding(self, index: int) -> bool:
        return int(self.content[index - 1, 0]) == -1
