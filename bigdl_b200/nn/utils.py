"""Shape / padding arithmetic and shared error texts of the nn package, as public helpers (reference: S/nn/Utils.scala —
``getSAMEOutSizeAndPadding`` :307-332, ``getOutSizeAndPadding`` :340-394, ``getOutputShape`` :444-460,
``getPaddingAndOutputSize`` :524-565, ``isLayerwiseScaled`` :248-257 — and S/nn/ErrorInfo.scala). The layers themselves use
``layers_core.conv_out_size`` / ``same_padding``; these wrappers keep the reference's argument order and return layout
for code that calls them directly."""
from __future__ import annotations

import math

from .layers_core import conv_out_size, same_padding


def getSAMEOutSizeAndPadding(inputHeight, inputWidth, dH, dW, kH, kW, inputDepth=-1, dT=-1, kT=-1):
    """TF "SAME": [padTop, padBottom, padLeft, padRight, oH, oW] (with depth: [padFront, padBack, …, oT, oH, oW])."""
    pt, pb, oh = same_padding(inputHeight, kH, dH)
    pl, pr, ow = same_padding(inputWidth, kW, dW)
    if inputDepth != -1:
        if dT <= 0 or kT <= 0:
            raise ValueError("kernel size and strideSize cannot be smaller than 0")
        pf, pk, ot = same_padding(inputDepth, kT, dT)
        return [pf, pk, pt, pb, pl, pr, ot, oh, ow]
    return [pt, pb, pl, pr, oh, ow]


def getOutSizeAndPadding(inputHeight, inputWidth, dH, dW, kH, kW, padH, padW, ceilMode, dilationHeight=1,
                         dilationWidth=1, inputdepth=-1, dt=-1, kt=-1, padt=0, dilationDepth=1):
    """Torch rule with symmetric padding: [padH, padH, padW, padW, oH, oW] (3-D: [padT, padT, padH, padH, padW, padW,
    oT, oH, oW]). The "last window must start inside the input" correction applies only when some padding is set."""
    any_pad = padH != 0 or padW != 0 or padt != 0

    def out(size, k, stride, pad, dil):
        dk = dil * (k - 1) + 1
        num = size - dk + 2 * pad
        o = (math.ceil(num / stride) if ceilMode else math.floor(num / stride)) + 1
        if any_pad and (o - 1) * stride >= size + pad:
            o -= 1
        return int(o)

    oh, ow = out(inputHeight, kH, dH, padH, dilationHeight), out(inputWidth, kW, dW, padW, dilationWidth)
    if inputdepth > 0:
        if dt <= 0 or kt <= 0 or padt < 0:
            raise ValueError("kernel size, stride size, padding size cannot be smaller than 0")
        return [padt, padt, padH, padH, padW, padW, out(inputdepth, kt, dt, padt, dilationDepth), oh, ow]
    return [padH, padH, padW, padW, oh, ow]


def getPaddingAndOutputSize(inputHeight, inputWidth, dH, dW, kH, kW, padH, padW, ceilMode=False):
    """(padTop, padBottom, padLeft, padRight, oH, oW) with the bottom/right padding grown until the last (ceil-mode)
    window fits — what the dnn pooling layers need."""
    def out(size, k, stride, pad):
        num = size - k + 2 * pad
        return int((math.ceil(num / stride) if ceilMode else math.floor(num / stride)) + 1)

    oh, ow = out(inputHeight, kH, dH, padH), out(inputWidth, kW, dW, padW)
    if padH != 0 or padW != 0 or kH == 1 or kW == 1:
        if (oh - 1) * dH >= inputHeight + padH:
            oh -= 1
        if (ow - 1) * dW >= inputWidth + padW:
            ow -= 1
    pad_b = max(padH, dH * (oh - 1) + kH - inputHeight - padH)
    pad_r = max(padW, dW * (ow - 1) + kW - inputWidth - padW)
    return padH, pad_b, padW, pad_r, oh, ow


def getOutputShape(outputHeight, outputWidth, nOutputPlane, batchSize=-1, format="NCHW"):
    fmt = getattr(format, "name", format)
    core = [nOutputPlane, outputHeight, outputWidth] if fmt == "NCHW" else [outputHeight, outputWidth, nOutputPlane]
    return core if batchSize == -1 else [batchSize] + core


def isLayerwiseScaled(model) -> bool:
    """True when any module of the tree has a scaleW / scaleB different from 1."""
    if getattr(model, "modules", None):
        return any(isLayerwiseScaled(m) for m in model.modules)
    return model.getScaleW() != 1 or model.getScaleB() != 1


class ErrorInfo:
    constrainInputAsVectorOrBatch = (
        "The input to the layer needs to be a vector (or a mini-batch of vectors); please use the Reshape module to "
        "convert multi-dimensional input into vectors if appropriate")
    constrainInputAs3DOrBatch = (
        "The input to the layer needs to be a 3D tensor (or a mini-batch of 3D tensors); please use the Reshape module "
        "to convert multi-dimensional input into 3D tensors if appropriate")
    constrainEachInputAsVectorOrBatch = (
        "Each tensor contained in the input to the layer needs to be a vector (or a mini-batch of vectors); please use "
        "the Reshape module to convert multi-dimensional input into vectors if appropriate")
    constrainInputDimSameAsTarget = (
        "The dimensions of input and target to the criterion layer need to be the same; please use the Reshape module "
        "to convert if appropriate")
    constrainInputSizeSameAsTarget = (
        "The size of input and target to the criterion layer need to be the same; please use the Reshape module to "
        "convert if appropriate")


__all__ = ["getSAMEOutSizeAndPadding", "getOutSizeAndPadding", "getPaddingAndOutputSize", "getOutputShape",
           "isLayerwiseScaled", "ErrorInfo", "conv_out_size", "same_padding"]
