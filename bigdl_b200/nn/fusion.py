"""Graph-level fusion pass (reference: S/nn/mkldnn/Fusion.scala:34-332 and the per-flag switches
``bigdl.mkldnn.fusion.{convbn,bnrelu,convrelu,convsum}`` in S/nn/mkldnn/DnnGraph.scala:412-463).

* conv/linear + ReLU  → ReLU runs in the GEMM epilogue (and its backward is fused with the bias-gradient reduction)
* BN + ReLU           → ReLU runs inside the batch-norm kernels
* CAddTable + ReLU    → one add+ReLU kernel (residual blocks)
* conv + BN (inference only, ``fold_bn``) → BN folded into the conv weights:
  ``w *= gamma/sqrt(var+eps)``, ``b = gamma/sqrt(var+eps)*(b - mean) + beta``  (Fusion.scala:173-217)
* channel Concat → zero-copy (handled inside ``Concat``; needs the ReLU folding done here first)

The pass only sets flags on the modules; ``unfuse`` clears them. Numerics are unchanged (same math, fewer passes).
"""
from __future__ import annotations

import torch

from ..utils import config
from .layers_core import BatchNormalization, Linear, ReLU, SpatialConvolution
from .layers_table import CAddTable
from .module import Graph, Module, Sequential


def _pairs(model: Module):
    """(producer, consumer) pairs where consumer is the only user of producer's output."""
    for m in model.all_modules():
        if isinstance(m, Sequential):
            for a, b in zip(m.modules, m.modules[1:]):
                yield a, b
        elif isinstance(m, Graph):
            for n in m.exec_order:
                if len(n.next) == 1 and len(n.next[0].prev) == 1 and n not in m.output_nodes:
                    yield n.element, n.next[0].element


def fuse(model: Module) -> Module:
    if not config.get("bigdl.mkldnn.fusion"):
        return model
    conv_relu = config.get("bigdl.mkldnn.fusion.convrelu")
    bn_relu = config.get("bigdl.mkldnn.fusion.bnrelu")
    conv_sum = config.get("bigdl.mkldnn.fusion.convsum")
    for a, b in _pairs(model):
        if type(b) is not ReLU or getattr(b, "_fused", False):
            continue
        if conv_relu and isinstance(a, (SpatialConvolution, Linear)) and not a._fuse_relu:
            a._fuse_relu, b._fused = True, True
        elif bn_relu and isinstance(a, BatchNormalization) and not a._fuse_relu and not getattr(a, "_folded", False):
            a._fuse_relu, b._fused = True, True
        elif conv_sum and isinstance(a, CAddTable) and not a._fuse_relu:
            a._fuse_relu, b._fused = True, True
    return model


def unfuse(model: Module) -> Module:
    for m in model.all_modules():
        if isinstance(m, (SpatialConvolution, Linear, BatchNormalization, CAddTable)):
            m._fuse_relu = False
        if type(m) is ReLU:
            m._fused = False
    return model


def fold_bn(model: Module) -> Module:
    """Inference-time conv+BN folding (Fusion.scala:173-217). The BN module becomes an identity."""
    if not config.get("bigdl.mkldnn.fusion.convbn"):
        return model
    for a, b in _pairs(model):
        if isinstance(a, SpatialConvolution) and isinstance(b, BatchNormalization) and not getattr(b, "_folded", False):
            with torch.no_grad():
                var, mean = b._buffers["runningVar"], b._buffers["runningMean"]
                inv = 1.0 / torch.sqrt(var + b.eps)
                gamma = b.weight.detach() if b.affine else torch.ones_like(var)
                beta = b.bias.detach() if b.affine else torch.zeros_like(var)
                scale = (gamma * inv).to(a.weight.device)
                a.weight.mul_(scale.reshape(-1, 1, 1, 1))
                if not a.with_bias:
                    a.with_bias = True
                    a.register_parameter("bias", torch.zeros(a.n_output_plane, device=a.weight.device))
                a.bias.copy_(scale * (a.bias.detach() - mean.to(a.weight.device)) + beta.to(a.weight.device))
            b._folded = True
            a._mark_weights_dirty()
    return model
