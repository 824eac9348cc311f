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


# ------------------------------------------------------------------------------------------------ activity helpers
def _is_table(x):
    from ..utils.table import Table

    return isinstance(x, Table)


def recursiveTensorApply1(x, func):
    """Apply ``func`` to every tensor of an activity (tensor, or arbitrarily nested Table); returns the same structure."""
    from ..utils.table import map_activity

    return map_activity(func, x)


def recursiveTensorApply2(x, y, func):
    """``func(x_i, y_i)`` over two activities of the same structure; the results replace the tensors of ``x``."""
    import torch

    if isinstance(x, torch.Tensor):
        if not isinstance(y, torch.Tensor):
            raise ValueError("x and y should have the same structure")
        return func(x, y)
    if not (_is_table(x) and _is_table(y)) or len(x) != len(y):
        raise ValueError("x and y should have the same structure")
    for k, v in list(x.items()):
        x[k] = recursiveTensorApply2(v, y[k], func)
    return x


def recursiveResizeAs(target, src):
    """A zero-filled activity with the structure and shapes of ``src``; tensors of ``target`` that already have the
    right shape are reused."""
    import torch

    from ..utils.table import Table

    if isinstance(src, torch.Tensor):
        if isinstance(target, torch.Tensor) and target.shape == src.shape and target.dtype == src.dtype:
            return target
        return torch.zeros_like(src)
    out = target if _is_table(target) else Table()
    for k, v in src.items():
        out[k] = recursiveResizeAs(out[k] if k in out else None, v)
    for k in [k for k in out.keys() if k not in src]:
        del out[k]
    return out


def recursiveAdd(y, alpha=1.0, x=None):
    """y += alpha · x, tensor by tensor."""
    return recursiveTensorApply2(y, x, lambda a, b: a.add_(b, alpha=alpha))


def recursiveCopy(y, x):
    """y ← x, tensor by tensor (shapes follow ``x``)."""
    return recursiveTensorApply2(y, x, lambda a, b: a.resize_(b.shape).copy_(b))


def recursiveFill(x, value: float):
    return recursiveTensorApply1(x, lambda t: t.fill_(value))


def zeroTableCopy(t1, t2):
    """Give ``t1`` zero tensors shaped like the tensor entries of ``t2``; non-tensor entries are carried over."""
    import torch

    for k, v in t2.items():
        if isinstance(v, torch.Tensor):
            t1[k] = torch.zeros_like(v)
        elif _is_table(v):
            from ..utils.table import Table

            t1[k] = zeroTableCopy(t1[k] if k in t1 and _is_table(t1[k]) else Table(), v)
        else:
            t1[k] = v
    for k in [k for k in t1.keys() if k not in t2]:
        del t1[k]
    return t1


def getNamedModules(model) -> dict:
    """{name: module} over the whole tree (named or generated names alike)."""
    return {m.getName(): m for m in model.all_modules()}


def copyModule(src, dst):
    """Copy parameters (and names / scales) from ``src`` into the structurally identical ``dst``."""
    import torch

    ps, pd = src.parameters()[0], dst.parameters()[0]
    if len(ps) != len(pd):
        raise ValueError("the two modules do not have the same number of parameters")
    with torch.no_grad():
        for a, b in zip(pd, ps):
            a.copy_(b.reshape(a.shape))
    dst.setScaleW(src.getScaleW())
    dst.setScaleB(src.getScaleB())
    if src.hasName():
        dst.setName(src.getName())
    dst._mark_weights_dirty()
    return dst


def shuffle(src, permutation, buffer=None):
    """Permute the dimensions of ``src`` (0-based permutation) into a contiguous tensor."""
    t = getattr(src, "t", src) if not hasattr(src, "permute") else src
    perm = [int(p) for p in permutation]
    if len(perm) != t.dim() or sorted(perm) != list(range(t.dim())):
        raise ValueError("permutation should list every dimension exactly once")
    out = t.permute(*perm).contiguous()
    if buffer is not None:
        buffer.resize_(out.shape).copy_(out)
        return buffer
    return out


def calculateFwdBwdTime(times):
    """Sum of the forward and of the backward times of ``model.getTimes()`` → (forward ns, backward ns)."""
    f = sum(t[1] for t in times)
    b = sum(t[2] for t in times)
    return f, b


class Utils:
    """Object form (S/nn/Utils.scala) of the helpers in this module."""


for _n in ("getSAMEOutSizeAndPadding", "getOutSizeAndPadding", "getPaddingAndOutputSize", "getOutputShape",
           "isLayerwiseScaled", "recursiveTensorApply1", "recursiveTensorApply2", "recursiveResizeAs", "recursiveAdd",
           "recursiveCopy", "recursiveFill", "zeroTableCopy", "getNamedModules", "copyModule", "shuffle",
           "calculateFwdBwdTime"):
    setattr(Utils, _n, staticmethod(globals()[_n]))


class NNPrimitive:
    """Patch-matrix primitives (S/nn/NNPrimitive.scala): ``im2col`` / ``col2im`` for NCHW tensors. The convolution
    layers do not use them — they run the implicit-GEMM kernel (TMA im2col gathers, no patch matrix in memory) — these
    exist for code that builds its own unfold-based layers."""

    @staticmethod
    def im2col(input, kH, kW, dH=1, dW=1, padH=0, padW=0, dilationH=1, dilationW=1):
        """[N, C, H, W] → [N, C·kH·kW, Ho·Wo]"""
        import torch.nn.functional as F

        return F.unfold(input, (kH, kW), dilation=(dilationH, dilationW), padding=(padH, padW), stride=(dH, dW))

    @staticmethod
    def col2im(cols, outputSize, kH, kW, dH=1, dW=1, padH=0, padW=0, dilationH=1, dilationW=1):
        """The adjoint: [N, C·kH·kW, L] → [N, C, H, W] with overlapping patches summed."""
        import torch.nn.functional as F

        return F.fold(cols, tuple(outputSize), (kH, kW), dilation=(dilationH, dilationW), padding=(padH, padW),
                      stride=(dH, dW))


__all__ = ["recursiveTensorApply1", "recursiveTensorApply2", "recursiveResizeAs", "recursiveAdd", "recursiveCopy",
           "recursiveFill", "zeroTableCopy", "getNamedModules", "copyModule", "shuffle", "calculateFwdBwdTime", "Utils",
           "NNPrimitive", "getSAMEOutSizeAndPadding", "getOutSizeAndPadding", "getPaddingAndOutputSize", "getOutputShape",
           "isLayerwiseScaled", "ErrorInfo", "conv_out_size", "same_padding"]
