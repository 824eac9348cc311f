"""ResNet for CIFAR-10 (depth 6n+2) and ImageNet (18/34/50/101/152/200), shortcut types A/B/C.

Reference: S/models/resnet/ResNet.scala:56-148 (Convolution = MSRA-initialised bias-free conv, Sbn = BN with eps 1e-3,
modelInit), :150-260 (builder). ``ResNet.modelInit`` re-initialises conv weights N(0, sqrt(2/(kW*kW*nOut))), BN
gamma=1/beta=0 and linear bias 0, which overwrites the zero-gamma init of each bottleneck's last BN exactly as in the
reference.
"""
from __future__ import annotations

import math

import torch

from ..nn.layer import (CAddTable, Concat, ConcatTable, Identity, Linear, LogSoftMax, MsraFiller, MulConstant,
                        RandomNormal, ReLU, Sequential, SpatialAveragePooling, SpatialBatchNormalization,
                        SpatialConvolution, SpatialMaxPooling, View, Zeros)
from ..optim.regularizer import L2Regularizer
from ..utils.random_generator import RNG


class ShortcutType:
    A, B, C = "A", "B", "C"


class DatasetType:
    CIFAR10, ImageNet = "cifar10", "imagenet"


def Convolution(n_in, n_out, kw, kh, sw=1, sh=1, pw=0, ph=0, n_group=1, propagate_back=True):
    conv = SpatialConvolution(n_in, n_out, kw, kh, sw, sh, pw, ph, n_group, propagate_back,
                              wRegularizer=L2Regularizer(1e-4), bRegularizer=L2Regularizer(1e-4), with_bias=False)
    conv.setInitMethod(MsraFiller(False), Zeros())
    return conv


def Sbn(n_output, eps=1e-3, momentum=0.1, affine=True):
    return SpatialBatchNormalization(n_output, eps, momentum, affine)


def ResNet(class_num: int = 1000, depth: int = 50, shortcut_type: str = ShortcutType.B,
           dataset: str = DatasetType.ImageNet, with_log_softmax: bool = False):
    state = {"ich": 0}

    def shortcut(n_in, n_out, stride):
        use_conv = shortcut_type == ShortcutType.C or (shortcut_type == ShortcutType.B and n_in != n_out)
        if use_conv:
            return Sequential().add(Convolution(n_in, n_out, 1, 1, stride, stride)).add(Sbn(n_out))
        if n_in != n_out:
            return Sequential().add(SpatialAveragePooling(1, 1, stride, stride)) \
                .add(Concat(2).add(Identity()).add(MulConstant(0.0)))
        return Identity()

    def basic_block(n, stride):
        n_in = state["ich"]
        state["ich"] = n
        s = Sequential()
        s.add(Convolution(n_in, n, 3, 3, stride, stride, 1, 1)).add(Sbn(n)).add(ReLU(True))
        s.add(Convolution(n, n, 3, 3, 1, 1, 1, 1)).add(Sbn(n))
        return Sequential().add(ConcatTable().add(s).add(shortcut(n_in, n, stride))).add(CAddTable(True)).add(ReLU(True))

    def bottleneck(n, stride):
        n_in = state["ich"]
        state["ich"] = n * 4
        s = Sequential()
        s.add(Convolution(n_in, n, 1, 1, 1, 1, 0, 0)).add(Sbn(n)).add(ReLU(True))
        s.add(Convolution(n, n, 3, 3, stride, stride, 1, 1)).add(Sbn(n)).add(ReLU(True))
        s.add(Convolution(n, n * 4, 1, 1, 1, 1, 0, 0)).add(Sbn(n * 4).setInitMethod(Zeros(), Zeros()))
        return Sequential().add(ConcatTable().add(s).add(shortcut(n_in, n * 4, stride))).add(CAddTable(True)).add(ReLU(True))

    def layer(block, features, count, stride=1):
        s = Sequential()
        for i in range(count):
            s.add(block(features, stride if i == 0 else 1))
        return s

    model = Sequential()
    if dataset == DatasetType.ImageNet:
        cfg = {18: ((2, 2, 2, 2), 512, basic_block), 34: ((3, 4, 6, 3), 512, basic_block),
               50: ((3, 4, 6, 3), 2048, bottleneck), 101: ((3, 4, 23, 3), 2048, bottleneck),
               152: ((3, 8, 36, 3), 2048, bottleneck), 200: ((3, 24, 36, 3), 2048, bottleneck)}
        if depth not in cfg:
            raise ValueError(f"Invalid depth {depth}")
        loops, n_features, block = cfg[depth]
        state["ich"] = 64
        model.add(Convolution(3, 64, 7, 7, 2, 2, 3, 3, propagate_back=False)).add(Sbn(64)).add(ReLU(True))
        model.add(SpatialMaxPooling(3, 3, 2, 2, 1, 1))
        model.add(layer(block, 64, loops[0])).add(layer(block, 128, loops[1], 2))
        model.add(layer(block, 256, loops[2], 2)).add(layer(block, 512, loops[3], 2))
        model.add(SpatialAveragePooling(7, 7, 1, 1)).add(View(n_features).setNumInputDims(3))
        model.add(Linear(n_features, class_num, True, L2Regularizer(1e-4), L2Regularizer(1e-4))
                  .setInitMethod(RandomNormal(0.0, 0.01), Zeros()))
    elif dataset == DatasetType.CIFAR10:
        if (depth - 2) % 6:
            raise ValueError("depth should be one of 20, 32, 44, 56, 110, 1202")
        n = (depth - 2) // 6
        state["ich"] = 16
        model.add(Convolution(3, 16, 3, 3, 1, 1, 1, 1, propagate_back=False)).add(SpatialBatchNormalization(16))
        model.add(ReLU(True))
        model.add(layer(basic_block, 16, n)).add(layer(basic_block, 32, n, 2)).add(layer(basic_block, 64, n, 2))
        model.add(SpatialAveragePooling(8, 8, 1, 1)).add(View(64).setNumInputDims(3)).add(Linear(64, 10))
    else:
        raise ValueError(f"Invalid dataset {dataset}")
    if with_log_softmax:
        model.add(LogSoftMax())
    return model


def model_init(model):
    """ResNet.modelInit (S/models/resnet/ResNet.scala:118-148)."""
    with torch.no_grad():
        for m in model.all_modules():
            if isinstance(m, SpatialConvolution):
                n = m.kernel_w * m.kernel_w * m.n_output_plane
                tmp = torch.zeros(m.weight.shape)
                RNG.fill_normal(tmp, 0.0, math.sqrt(2.0 / n))
                m.weight.copy_(tmp)
                if m.with_bias:
                    m.bias.zero_()
            elif isinstance(m, SpatialBatchNormalization):
                m.weight.fill_(1.0)
                m.bias.zero_()
            elif isinstance(m, Linear) and m.with_bias:
                m.bias.zero_()
    model._mark_weights_dirty()
    return model


modelInit = model_init


def _resnet_from_opt(class_num, opt=None, **kw):
    """``ResNet(classNum, T("shortcutType" -> …, "depth" -> …, "dataSet" -> …, "optnet" -> …))`` (ResNet.scala:150)."""
    o = dict(opt.items()) if hasattr(opt, "items") else dict(opt or {})
    o.update(kw)
    return ResNet(class_num, depth=int(o.get("depth", 18)), shortcut_type=str(o.get("shortcutType", ShortcutType.B)),
                  dataset=str(o.get("dataSet", o.get("dataset", DatasetType.CIFAR10))).lower())


ResNet.apply = _resnet_from_opt
ResNet.graph = lambda class_num, opt=None, **kw: _resnet_from_opt(class_num, opt, **kw).toGraph()     # ResNet.scala:284
ResNet.modelInit = model_init


def shareGradInput(model):
    """``ResNet.shareGradInput`` (ResNet.scala:78-116) makes layers of the same kind share one gradInput buffer to cut
    the JVM heap. Here activations and their gradients live in the autograd graph and are freed as soon as their last
    consumer ran, which is the effect that call buys; kept as a no-op so training scripts port unchanged."""
    return model


ResNet.shareGradInput = shareGradInput
