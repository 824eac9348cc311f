"""The small shared types of ``com.intel.analytics.bigdl.nn.abstractnn`` that have no home of their own:
``DataFormat`` (DataFormat.scala:31-68), ``Activity`` / ``EmptyGradInput`` (Activity.scala:30-120), ``InferShape`` /
``IdentityOutputShape`` / ``InvalidLayer`` (InferShape.scala:25-120), ``Initializable`` (Initializable.scala:21-50) and
``SizeAverageStatus`` (AbstractCriterion.scala:135-138). ``AbstractModule`` / ``AbstractCriterion`` themselves are
``nn.module.Module`` and ``nn.criterion.Criterion``."""
from __future__ import annotations

import torch

from ..utils.table import Table
from .criterion import Criterion as AbstractCriterion  # noqa: F401
from .module import Module as AbstractModule  # noqa: F401

TensorModule = AbstractModule          # a module whose input and output are tensors (AbstractModule.scala:38)
TensorCriterion = AbstractCriterion


class _Format:
    def __init__(self, value: str):
        self.value = value

    def getHWCDims(self, inputDims: int):
        """1-based (height, width, channel) dimension indices of a 3-D (no batch) or 4-D input."""
        if self.value == "NCHW":
            return (2, 3, 1) if inputDims == 3 else (3, 4, 2)
        return (1, 2, 3) if inputDims == 3 else (2, 3, 4)

    def __eq__(self, o):
        return (o.value if isinstance(o, _Format) else str(o).upper()) == self.value

    def __hash__(self):
        return hash(self.value)

    def __str__(self):
        return self.value

    __repr__ = __str__


class DataFormat:
    """``DataFormat("nhwc")`` → ``DataFormat.NHWC``; layers here take the plain strings "NCHW" / "NHWC", which these
    objects compare equal to."""

    NCHW, NHWC = _Format("NCHW"), _Format("NHWC")

    def __new__(cls, formatString: str):
        s = str(formatString).upper()
        if s == "NCHW":
            return cls.NCHW
        if s == "NHWC":
            return cls.NHWC
        raise ValueError(f"unknown data format {formatString}")


class Activity:
    """A layer input / output is a Tensor or a Table. torch Tensors and ``utils.table.Table`` play that role directly;
    these helpers are the reference's ``isTensor`` / ``isTable`` / ``toTensor`` / ``toTable`` / ``allocate``."""

    @staticmethod
    def isTensor(a) -> bool:
        return isinstance(a, torch.Tensor)

    @staticmethod
    def isTable(a) -> bool:
        return isinstance(a, Table)

    @staticmethod
    def toTensor(a) -> torch.Tensor:
        if isinstance(a, EmptyGradInput):
            return a.toTensor()
        if not isinstance(a, torch.Tensor):
            raise TypeError(f"{type(a).__name__} is not a Tensor")
        return a

    @staticmethod
    def toTable(a) -> Table:
        if isinstance(a, EmptyGradInput):
            return a.toTable()
        if not isinstance(a, Table):
            raise TypeError(f"{type(a).__name__} is not a Table")
        return a

    @staticmethod
    def allocate(kind="tensor", dtype=torch.float32):
        """An empty activity of the asked kind: ``"table"`` / ``Table`` or a tensor of ``dtype``."""
        if kind in ("table", Table):
            return Table()
        return torch.empty(0, dtype=dtype)


class EmptyGradInput(Activity):
    """Placeholder gradInput of a layer that does not propagate (``propagateBack = false``): any use is an error."""

    def __init__(self, moduleName: str):
        self.moduleName = moduleName

    def _fail(self, *a, **k):
        raise RuntimeError(f"The gradInput of {self.moduleName} is empty. You should not use it anywhere")

    toTensor = toTable = isTensor = isTable = _fail


class InvalidLayer(RuntimeError):
    """Raised when Keras-style and Torch-style layers are mixed in one container."""


class InferShape:
    """Shape inference protocol of the Keras-style layers (``nn.keras.layers.KerasLayer`` implements it): ``build``
    records the input shape and computes the output shape; Torch-style layers answer ``isKerasStyle() == False`` and
    cannot report shapes."""

    _inputShapeValue = None
    _outputShapeValue = None

    def isKerasStyle(self) -> bool:
        return False

    def isBuilt(self) -> bool:
        return self._outputShapeValue is not None

    def allowRebuilt(self) -> bool:
        return False

    def computeOutputShape(self, inputShape):
        raise RuntimeError("Haven't been implemented yet. Do not use it with Keras Layer")

    def build(self, inputShape):
        out = self.computeOutputShape(inputShape)
        self._inputShapeValue, self._outputShapeValue = inputShape, out
        return out

    def getInputShape(self):
        if not self.isKerasStyle():
            raise RuntimeError("Torch style definition doesn't support getInputShape for now.")
        return self._inputShapeValue

    def getOutputShape(self):
        if not self.isKerasStyle():
            raise RuntimeError("Torch style definition doesn't support getOutputShape for now.")
        if not self.isBuilt():
            raise RuntimeError("This module hasn't been built.")
        return self._outputShapeValue

    def excludeInvalidLayers(self, modules):
        mine = self.isKerasStyle()
        bad = [m for m in modules if bool(getattr(m, "isKerasStyle", lambda: False)()) != mine]
        if bad:
            raise InvalidLayer(f"Do not mix {self}(isKerasStyle={mine}) with Layer (isKerasStyle={not mine}): "
                               + ",".join(str(m) for m in bad))

    def validateInput(self, modules):
        if self.isKerasStyle() and not modules:
            raise ValueError("Empty input is not allowed")
        self.excludeInvalidLayers(modules)


class IdentityOutputShape(InferShape):
    """Layers whose output has the shape of their input (activations, dropout, normalisation)."""

    def computeOutputShape(self, inputShape):
        return inputShape


class Initializable:
    """``setInitMethod(weightInitMethod, biasInitMethod)`` + ``reset()``; every parameterised layer of ``nn.layer``
    has both through ``Module``. Mixing this into a layer of your own gives the same behaviour."""

    weightInitMethod = None
    biasInitMethod = None

    def setInitMethod(self, weightInitMethod=None, biasInitMethod=None):
        if isinstance(weightInitMethod, (list, tuple)):
            raise NotImplementedError(f"setInitMethod with an array of InitializationMethod is not supported for {self}")
        if weightInitMethod is not None:
            self.weightInitMethod = weightInitMethod
        if biasInitMethod is not None:
            self.biasInitMethod = biasInitMethod
        self.reset()
        return self

    def reset(self):  # pragma: no cover - interface
        raise NotImplementedError


class SizeAverageStatus:
    """Whether a criterion wrapped by TimeDistributedCriterion / ParallelCriterion averages over the batch."""
    True_, False_, None_ = "True", "False", "None"


setattr(SizeAverageStatus, "True", SizeAverageStatus.True_)
setattr(SizeAverageStatus, "False", SizeAverageStatus.False_)
setattr(SizeAverageStatus, "None", SizeAverageStatus.None_)
