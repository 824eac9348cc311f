"""The rest of ``com.intel.analytics.bigdl.nn.tf`` / ``nn.ops`` as constructible layers: the backward-half ops of a
TensorFlow training graph (``ReluGrad`` … ``FusedBatchNormGrad``), the TF-layout ``Conv2D`` / ``Conv3D`` /
``MaxPool`` / ``FusedBatchNorm`` operations and the image decoders (S/nn/tf/{NNOps,MathOps,ImageOps,ParsingOps}.scala,
S/nn/ops/{Conv*Backprop*,Dilation2DBackprop*,DepthwiseConv2DBackprop*,ResizeBilinearGrad,MaxPoolGrad…}.scala).

One implementation serves both entry points: every class here builds the NodeDef a GraphDef would carry (from the
reference's constructor arguments) and hands it to the op loader of ``interop/tensorflow*.py``, so a layer made by hand
and a layer made by ``loadTF`` are the same module. Padding arguments follow the reference: ``-1`` means TensorFlow
"SAME", anything else "VALID"."""
from __future__ import annotations

from .module import Module


def _pad(p) -> str:
    if isinstance(p, str):
        return p.upper()
    return "SAME" if int(p) == -1 else "VALID"


def _fmt(f) -> str:
    return str(f).upper()


class _TFOp(Module):
    """Module backed by the TensorFlow op loader named ``op_name`` (default: the class name)."""

    op_name: str | None = None

    def __init__(self, **attrs):
        super().__init__()
        from ..interop import tensorflow as TF

        self.attrs = attrs
        b = TF._GraphBuilder()
        b.node(self.op_name or type(self).__name__, self.getName(), [], **attrs)
        self.impl = TF._LOADERS[self.op_name or type(self).__name__](b.gd.node[0])

    def updateOutput(self, input):
        return self.impl.updateOutput(input)


# ---------------------------------------------------------------------------------------------- element-wise gradients
class ReluGrad(_TFOp):
    """(gradients, features) → gradients · [features > 0]"""


class Relu6Grad(_TFOp):
    """(gradients, features) → gradients · [0 < features < 6]"""


class EluGrad(_TFOp):
    """(gradients, outputs) → gradients · (outputs > 0 ? 1 : outputs + 1)"""


class SigmoidGrad(_TFOp):
    """(y, dy) → dy · y · (1 − y)"""


class TanhGrad(_TFOp):
    """(y, dy) → dy · (1 − y²)"""


class SoftplusGrad(_TFOp):
    """(gradients, features) → gradients · sigmoid(features)"""


class SoftsignGrad(_TFOp):
    """(gradients, features) → gradients / (1 + |features|)²"""


class SqrtGrad(_TFOp):
    """(y, dy) → dy · 0.5 / y"""


class RsqrtGrad(_TFOp):
    """(y, dy) → −0.5 · dy · y³"""


class InvGrad(_TFOp):
    """(y, dy) → −dy · y²"""


class ReciprocalGrad(_TFOp):
    """(y, dy) → −dy · y²"""


class BiasAddGrad(_TFOp):
    """Sum of the back-propagated gradient over every dimension but the channel one."""

    def __init__(self, dataFormat="NHWC"):
        super().__init__(data_format=_fmt(dataFormat))


# ---------------------------------------------------------------------------------------------- pooling / LRN
class MaxPool(_TFOp):
    def __init__(self, ksize, strides, padding: str, format="NHWC"):
        super().__init__(ksize=list(ksize), strides=list(strides), padding=_pad(padding), data_format=_fmt(format))


def _pool_attrs(kH, kW, strideW, strideH, padH, padW, fmt):
    f = _fmt(fmt)
    k = [1, kH, kW, 1] if f == "NHWC" else [1, 1, kH, kW]
    s = [1, strideH, strideW, 1] if f == "NHWC" else [1, 1, strideH, strideW]
    return dict(ksize=k, strides=s, padding=_pad(padH if padH == -1 else padW), data_format=f)


class MaxPoolGrad(_TFOp):
    """(original input, original output, gradient) → gradient w.r.t. the input."""

    def __init__(self, kH, kW, strideW, strideH, padH, padW, format="NHWC"):
        super().__init__(**_pool_attrs(kH, kW, strideW, strideH, padH, padW, format))


class AvgPoolGrad(_TFOp):
    """(shape of the original input, gradient) → gradient w.r.t. the input."""

    def __init__(self, kH, kW, strideW, strideH, padH, padW, dataFormat="NHWC"):
        super().__init__(**_pool_attrs(kH, kW, strideW, strideH, padH, padW, dataFormat))


class LRNGrad(_TFOp):
    """(gradient, original input, original output) → gradient w.r.t. the input."""

    def __init__(self, depthRadius: int = 5, bias: float = 1.0, alpha: float = 1.0, beta: float = 0.5):
        super().__init__(depth_radius=int(depthRadius), bias=float(bias), alpha=float(alpha), beta=float(beta))


class ResizeBilinearGrad(_TFOp):
    """(gradient, original image) → gradient w.r.t. the image."""

    def __init__(self, alignCorners: bool = False):
        super().__init__(align_corners=bool(alignCorners))


# ---------------------------------------------------------------------------------------------- convolutions
def _conv2d_attrs(strideH, strideW, padH, padW, fmt):
    f = _fmt(fmt)
    s = [1, strideH, strideW, 1] if f == "NHWC" else [1, 1, strideH, strideW]
    return dict(strides=s, padding=_pad(padH if padH == -1 else padW), data_format=f)


class Conv2D(_TFOp):
    """(input, filter [kH, kW, in, out]) → convolution in TensorFlow's layout."""

    def __init__(self, strideH, strideW, padH, padW, format="NHWC"):
        super().__init__(**_conv2d_attrs(strideH, strideW, padH, padW, format))


class Conv2DTranspose(_TFOp):
    """(input sizes, filter, gradient) → gradient w.r.t. the convolution input (TF ``Conv2DBackpropInput``)."""
    op_name = "Conv2DBackpropInput"

    def __init__(self, strideW, strideH, padW=-1, padH=-1, format="NCHW"):
        super().__init__(**_conv2d_attrs(strideH, strideW, padH, padW, format))


class Conv2DBackFilter(_TFOp):
    """(input, filter sizes, gradient) → gradient w.r.t. the filter (TF ``Conv2DBackpropFilter``)."""
    op_name = "Conv2DBackpropFilter"

    def __init__(self, strideW, strideH, padW=-1, padH=-1, format="NCHW"):
        super().__init__(**_conv2d_attrs(strideH, strideW, padH, padW, format))


class DepthwiseConv2DBackpropInput(_TFOp):
    op_name = "DepthwiseConv2dNativeBackpropInput"

    def __init__(self, strideW, strideH, padW, padH, dataFormat="NHWC"):
        super().__init__(**_conv2d_attrs(strideH, strideW, padH, padW, dataFormat))


class DepthwiseConv2DBackpropFilter(_TFOp):
    op_name = "DepthwiseConv2dNativeBackpropFilter"

    def __init__(self, strideW, strideH, padW, padH, dataFormat="NHWC"):
        super().__init__(**_conv2d_attrs(strideH, strideW, padH, padW, dataFormat))


def _conv3d_attrs(dT, dH, dW, padT, padH, padW, fmt):
    f = _fmt(fmt)
    f3 = {"NHWC": "NDHWC", "NCHW": "NCDHW"}.get(f, f)
    s = [1, dT, dH, dW, 1] if f3 == "NDHWC" else [1, 1, dT, dH, dW]
    return dict(strides=s, padding=_pad(-1 if -1 in (padT, padH, padW) else 0), data_format=f3)


class Conv3D(_TFOp):
    def __init__(self, dT, dH, dW, padT, padH, padW, format="NHWC"):
        super().__init__(**_conv3d_attrs(dT, dH, dW, padT, padH, padW, format))


class Conv3DBackpropInput(Conv3D):
    """(original input, filter, gradient) → gradient w.r.t. the input."""


class Conv3DBackpropInputV2(Conv3D):
    """(input sizes, filter, gradient) → gradient w.r.t. the input."""


class Conv3DBackpropFilter(Conv3D):
    """(input, original filter, gradient) → gradient w.r.t. the filter."""


class Conv3DBackpropFilterV2(Conv3D):
    """(input, filter sizes, gradient) → gradient w.r.t. the filter."""


class _Dilation(_TFOp):
    def __init__(self, strides, rates, padding: str):
        super().__init__(strides=list(strides), rates=list(rates), padding=_pad(padding))


class Dilation2DBackpropInput(_Dilation):
    """(input, filter, gradient) → gradient w.r.t. the input of the grey-scale morphological dilation."""


class Dilation2DBackpropFilter(_Dilation):
    """(input, filter, gradient) → gradient w.r.t. the structuring element."""


# ---------------------------------------------------------------------------------------------- batch norm
class FusedBatchNorm(_TFOp):
    """(x, scale, offset, mean, variance) → (y, batch mean, batch variance, saved mean, saved inverse std)."""

    def __init__(self, epsilon: float = 1e-4, isTraining: bool = True, momentum: float = 0.1, dataFormat="NHWC"):
        super().__init__(epsilon=float(epsilon), is_training=bool(isTraining), data_format=_fmt(dataFormat))
        self.momentum = momentum


class FusedBatchNormGrad(_TFOp):
    """(dy, x, scale, saved mean, saved inverse std) → (dx, dscale, doffset)."""

    def __init__(self, epsilon: float, dataFormat="NHWC", isTrain: bool = False):
        super().__init__(epsilon=float(epsilon), is_training=bool(isTrain), data_format=_fmt(dataFormat))


# ---------------------------------------------------------------------------------------------- decoding / parsing
class DecodeJpeg(_TFOp):
    def __init__(self, channels: int = 0, ratio: int = 1):
        if ratio != 1:
            raise ValueError("currently not supported sub-sampling")
        super().__init__(channels=int(channels))


class DecodePng(_TFOp):
    def __init__(self, channels: int = 0):
        super().__init__(channels=int(channels))


class DecodeBmp(_TFOp):
    def __init__(self, channels: int = 0):
        super().__init__(channels=int(channels))


class DecodeGif(_TFOp):
    def __init__(self):
        super().__init__()


def ParseSingleExample(tDense=None, denseKeys=None, denseShape=None):
    """``ParseSingleExample(tDense, denseKeys, denseShape)`` (ParsingOps.scala:93): types are "float" / "int64" /
    "string" (or torch dtypes / type tags), keys are the feature names, shapes the per-feature output shapes."""
    from .tf import ParseSingleExample as _P

    def kind(t):
        s = str(t).lower()
        return "float" if "float" in s or "double" in s else ("string" if "str" in s else "int64")
    keys = list(denseKeys or [])
    types = [kind(t) for t in (tDense or [])] + [None] * max(0, len(keys) - len(tDense or []))
    return _P(types, keys, list(denseShape or []))


# ---------------------------------------------------------------------------------------------- small state ops
class AssignGrad(Module):
    """Copies its input into the gradient buffer of a variable (StateOps.scala ``AssignGrad``): used by imported
    training graphs to hand a computed gradient to the optimizer."""

    def __init__(self, grad):
        super().__init__()
        self.grad = grad

    def updateOutput(self, input):
        self.grad.detach().copy_(input.detach().reshape(self.grad.shape))
        return input


class SelectTensor(Module):
    """(S/nn/ops/SelectTensor.scala) picks the entry ``index`` of a Table by key — the key may itself be a tensor —
    and optionally applies ``transformer`` to it."""

    def __init__(self, index, transformer=None):
        super().__init__()
        self.index, self.transformer = index, transformer

    def updateOutput(self, input):
        key = self.index
        item = None
        for k, v in input.items() if hasattr(input, "items") else enumerate(input, 1):
            if k is key or (not hasattr(k, "shape") and not hasattr(key, "shape") and k == key):
                item = v
                break
            if hasattr(k, "shape") and hasattr(key, "shape") and k.shape == key.shape and bool((k == key).all()):
                item = v
                break
        if item is None:
            raise KeyError(f"no entry {key} in the input table")
        return self.transformer.forward(item) if self.transformer is not None else item


class TensorModuleWrapper(Module):
    """Lets a tensor → tensor layer sit where an Operation is expected (TensorModuleWrapper.scala): forward only."""

    def __init__(self, module):
        super().__init__()
        self.module = module

    def updateOutput(self, input):
        return self.module.forward(input)
