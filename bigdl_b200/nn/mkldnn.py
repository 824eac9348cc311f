"""The "dnn" engine surface (reference: S/nn/mkldnn/*, 35 files — MklDnnModule/MklDnnContainer with
``compile(phase)``, MemoryData layout descriptors, ReorderMemory, Fusion, DnnGraph, BlasWrapper and dnn-native layer
classes selected by ``-Dbigdl.engineType=mkldnn``).

On B200 there is ONE engine: the sm_100a kernels behind the ordinary layers. What the MKL-DNN engine adds in the
reference — blocked layouts negotiated at compile time, primitive fusion, inference-only weight preparation — maps to:

* layout: activations are bf16 NHWC (physically channels-last) everywhere; ``MemoryData`` records shape/layout/dtype and
  ``ReorderMemory`` converts between NCHW/NHWC and fp32/bf16 at the graph boundary, which is the only place a reorder
  is ever needed;
* ``compile(phase)``: moves the model to the GPU (flat parameters + bf16 compute copy), runs the fusion pass
  (conv/linear+ReLU, BN+ReLU, add+ReLU epilogues, zero-copy Concat) and, for ``Phase.Inference``, folds BN into the
  preceding convolution;
* the layer classes below are the framework's layers under their mkldnn names, so model-definition code written
  against ``com.intel.analytics.bigdl.nn.mkldnn`` ports by changing the import.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch

from . import layer as L
from .fusion import fold_bn, fuse, unfuse
from .module import Container, Graph, Module
from .module import Sequential as _Sequential


class Phase:
    TrainingPhase, InferencePhase = "training", "inference"
    Training, Inference = TrainingPhase, InferencePhase


class Memory:
    class Format:
        nchw, nhwc, nc, x, oihw, ohwi, any = "nchw", "nhwc", "nc", "x", "oihw", "ohwi", "any"

    class DataType:
        F32, BF16, S32, S8, U8, F8E4M3 = "f32", "bf16", "s32", "s8", "u8", "f8e4m3"


@dataclass
class MemoryData:
    """Shape + layout + dtype (+ quantisation scales/mask) of an activation (S/nn/mkldnn/MemoryData.scala:22-299)."""

    shape: list
    layout: str = Memory.Format.nchw
    dtype: str = Memory.DataType.F32
    scales: list = field(default_factory=list)
    mask: int = 0

    def setMask(self, m):
        self.mask = m
        return self

    def setScales(self, s):
        self.scales = list(s)
        return self

    def cloneFormat(self):
        return type(self)(list(self.shape), self.layout, self.dtype, list(self.scales), self.mask)


class HeapData(MemoryData):
    """Host/plain-layout tensor (fp32, logical order)."""


class NativeData(MemoryData):
    """Device tensor in the engine's preferred layout (bf16 NHWC)."""


_TORCH_DT = {"f32": torch.float32, "bf16": torch.bfloat16, "s32": torch.int32, "s8": torch.int8, "u8": torch.uint8}


class ReorderMemory(Module):
    """Convert between two MemoryData descriptions (ReorderMemory.scala): dtype cast and NCHW↔NHWC physical order.
    The logical shape stays NCHW; "nhwc" means channels-last strides, which is what the conv/BN/pool kernels consume
    without a copy."""

    def __init__(self, input_format: MemoryData | None = None, output_format: MemoryData | None = None,
                 grad_input_format=None, grad_output_format=None):
        super().__init__()
        self.input_format, self.output_format = input_format, output_format

    def updateOutput(self, input):
        out = self.output_format
        if out is None:
            return input
        y = input
        dt = _TORCH_DT.get(out.dtype)
        if dt is not None and y.dtype != dt:
            y = y.to(dt)
        if y.dim() == 4:
            fmt = torch.channels_last if out.layout == Memory.Format.nhwc else torch.contiguous_format
            y = y.contiguous(memory_format=fmt)
        return y


class _DnnContainerMixin:
    def compile(self, phase: str = Phase.TrainingPhase, input_formats=None, device=None):
        """Prepare the model for ``phase`` on the GPU (or stay on CPU when there is none). Returns self."""
        dev = torch.device(device) if device is not None else \
            torch.device("cuda" if torch.cuda.is_available() else "cpu")
        if phase == Phase.InferencePhase:
            self.evaluate()
            fold_bn(self)
        else:
            self.training()
        self.to(dev)
        fuse(self)
        self._compiled_phase = phase
        return self

    def fusion(self):
        return fuse(self)

    def setRuntime(self, runtime=None):
        return self


class ReorderManager:
    """Decides where layout conversions are needed and caches them (ReorderManager.scala): ``register(from, to)``
    creates the ReorderMemory for a (producer format, consumer format) pair once; ``infer`` applies it."""

    def __init__(self):
        self._reorders: dict = {}

    @staticmethod
    def _key(a: MemoryData, b: MemoryData):
        return (type(a).__name__, tuple(a.shape), a.layout, a.dtype, type(b).__name__, tuple(b.shape), b.layout, b.dtype)

    def register(self, from_: MemoryData, to: MemoryData):
        if self.needReorder(from_, to) and self._key(from_, to) not in self._reorders:
            self._reorders[self._key(from_, to)] = ReorderMemory(from_, to)

    @staticmethod
    def needReorder(from_: MemoryData, to: MemoryData) -> bool:
        return from_.layout != to.layout or from_.dtype != to.dtype

    def infer(self, from_: MemoryData, to: MemoryData, output):
        r = self._reorders.get(self._key(from_, to))
        return output if r is None else r.forward(output)

    def release(self):
        self._reorders.clear()


class TensorMMap:
    """A parameter with a plain fp32 master ("dense") and the engine's compute copy ("native": bf16 on the GPU), kept
    in step by ``sync()`` — the heap/native weight mirror of TensorMMap.scala:34-131. The flat parameter buffer of a
    model does this for all layers at once; this class is the single-tensor form of the same contract."""

    def __init__(self, size, device=None):
        self.dense = torch.zeros(tuple(size))
        dev = torch.device(device) if device is not None else torch.device(
            "cuda" if torch.cuda.is_available() else "cpu")
        self._native = torch.zeros(tuple(size), dtype=torch.bfloat16 if dev.type == "cuda" else torch.float32, device=dev)

    def native(self) -> torch.Tensor:
        return self._native

    def sync(self):
        """dense → native (called every training forward in the reference)."""
        self._native.copy_(self.dense)
        return self

    def syncToHeap(self):
        self.dense.copy_(self._native)
        return self

    def release(self):
        self._native = None


class Sequential(_Sequential, _DnnContainerMixin):
    """mkldnn.Sequential (S/nn/mkldnn/Sequential.scala:29-445)."""


class DnnGraph(Graph, _DnnContainerMixin):
    """mkldnn.DnnGraph (S/nn/mkldnn/DnnGraph.scala): a static graph with ``compile(phase)`` and ``fusion()``."""

    @staticmethod
    def from_graph(g: Graph) -> "DnnGraph":
        return DnnGraph(g.input_nodes, g.output_nodes)


class BlasWrapper(Container):
    """Run any ordinary layer inside a dnn container (BlasWrapper.scala:100). All layers share one engine here, so the
    wrapper only forwards."""

    def __init__(self, module: Module):
        super().__init__()
        self.module = module
        self.modules = [module]

    def updateOutput(self, input):
        return self.module._call(input)


class Input(Module):
    """Graph/sequence entry that declares the input layout (mkldnn/Input.scala)."""

    def __init__(self, shape=None, layout: str = Memory.Format.nchw):
        super().__init__()
        self.format = HeapData(list(shape or []), layout)

    def updateOutput(self, input):
        return input


class Output(Module):
    """Sequence exit converting back to the plain fp32 layout (mkldnn/Output.scala)."""

    def __init__(self, out_layout: str = Memory.Format.nchw, grad_out_layout=None):
        super().__init__()
        self.layout = out_layout

    def updateOutput(self, input):
        if isinstance(input, torch.Tensor) and input.dtype == torch.bfloat16:
            return input.float().contiguous()
        return input


def RNN(mode: str, input_size: int, hidden_size: int, f=None, direction: str = "unidirectional_left2right",
        layers: int = 1):
    """Fused multi-layer (bi)directional RNN/LSTM/GRU (S/nn/mkldnn/RNN.scala:31-100; AlgKind vanilla_lstm | vanilla_gru |
    vanilla_rnn). Built from Recurrent/BiRecurrent cells whose gate math runs in csrc/recurrent.cu."""
    mode = mode.lower()

    def cell(n_in):
        if "lstm" in mode:
            return L.LSTM(n_in, hidden_size)
        if "gru" in mode:
            return L.GRU(n_in, hidden_size)
        return L.RnnCell(n_in, hidden_size, L.Tanh())

    bidir = direction.startswith("bidirectional")
    seq = L.Sequential()
    n_in = input_size
    for _ in range(layers):
        if bidir:
            merge = L.CAddTable() if direction.endswith("sum") else L.JoinTable(3, 3)
            seq.add(L.BiRecurrent(merge).add(cell(n_in)))
            n_in = hidden_size if direction.endswith("sum") else 2 * hidden_size
        else:
            seq.add(L.Recurrent().add(cell(n_in)))
            n_in = hidden_size
    return seq


# the dnn-named layers are the framework's layers
SpatialConvolution = L.SpatialConvolution
Linear = L.Linear
SpatialBatchNormalization = L.SpatialBatchNormalization
ReLU = L.ReLU
MaxPooling = L.SpatialMaxPooling
AvgPooling = L.SpatialAveragePooling
LRN = L.SpatialCrossMapLRN
SoftMax = L.SoftMax
Dropout = L.Dropout
CAddTable = L.CAddTable
ConcatTable = L.ConcatTable
JoinTable = L.JoinTable
SelectTable = L.SelectTable
Identity = L.Identity


class Fusion:
    """Graph rewrites of S/nn/mkldnn/Fusion.scala:34-332 as flag-setting passes (nn/fusion.py)."""

    fuse = staticmethod(fuse)
    unfuse = staticmethod(unfuse)
    fusionConvBn = staticmethod(fold_bn)


# the reference's base classes of dnn layers / containers: every layer here already satisfies their contract
MklDnnLayer = Module
MklDnnContainer = _DnnContainerMixin


class Perf:
    """Local perf harness entry point (S/nn/mkldnn/Perf.scala:40-140): ``Perf.main(["-m", "resnet50", ...])``."""

    @staticmethod
    def main(argv=None):
        from ..cli import perf

        return perf.main(argv)
