"""Inception-v2 / BN-Inception (S/models/inception/Inception_v2.scala:27-420): every convolution is followed by
SpatialBatchNormalization(eps 1e-3) + ReLU, 5x5 branches are replaced by double 3x3, reduction modules ("max", 0) use
stride-2 3x3 convolutions and a stride-2 max pool instead of a 1x1 branch."""
from __future__ import annotations

from ..nn.layer import (Concat, ConcatTable, Linear, LogSoftMax, ReLU, Sequential, SpatialAveragePooling,
                        SpatialBatchNormalization, SpatialConvolution, SpatialMaxPooling, View)


def _cbr(seq, n_in, n_out, k, s, p, name, bias=True):
    seq.add(SpatialConvolution(n_in, n_out, k, k, s, s, p, p, 1, with_bias=bias).setName(name))
    seq.add(SpatialBatchNormalization(n_out, 1e-3).setName(name + "/bn"))
    seq.add(ReLU(True).setName(name + "/bn/sc/relu"))
    return seq


def Inception_Layer_v2(input_size: int, config, name_prefix: str = ""):
    """config = ((n1x1,), (r3, n3), (rd3, nd3), (pool_kind, n_proj)) with n1x1 = 0 / n_proj = 0 for reduction modules."""
    (c1,), (r3, n3), (rd3, nd3), (pool_kind, n_proj) = config
    reduce = pool_kind == "max" and n_proj == 0
    stride = 2 if reduce else 1
    concat = Concat(2)
    if c1 != 0:
        concat.add(_cbr(Sequential(), input_size, c1, 1, 1, 0, name_prefix + "1x1"))
    b3 = _cbr(Sequential(), input_size, r3, 1, 1, 0, name_prefix + "3x3_reduce")
    concat.add(_cbr(b3, r3, n3, 3, stride, 1, name_prefix + "3x3"))
    bd = _cbr(Sequential(), input_size, rd3, 1, 1, 0, name_prefix + "double3x3_reduce")
    _cbr(bd, rd3, nd3, 3, 1, 1, name_prefix + "double3x3a")
    concat.add(_cbr(bd, nd3, nd3, 3, stride, 1, name_prefix + "double3x3b"))
    pool = Sequential()
    if pool_kind == "max":
        pool.add((SpatialMaxPooling(3, 3, 1, 1, 1, 1) if n_proj != 0 else SpatialMaxPooling(3, 3, 2, 2)).ceil()
                 .setName(name_prefix + "pool"))
    elif pool_kind == "avg":
        pool.add(SpatialAveragePooling(3, 3, 1, 1, 1, 1).ceil().setName(name_prefix + "pool"))
    else:
        raise ValueError(pool_kind)
    if n_proj != 0:
        _cbr(pool, input_size, n_proj, 1, 1, 0, name_prefix + "pool_proj")
    concat.add(pool)
    return concat.setName(name_prefix + "output")


_CFG = [
    (192, ((64,), (64, 64), (64, 96), ("avg", 32)), "inception_3a/"),
    (256, ((64,), (64, 96), (64, 96), ("avg", 64)), "inception_3b/"),
    (320, ((0,), (128, 160), (64, 96), ("max", 0)), "inception_3c/"),
    (576, ((224,), (64, 96), (96, 128), ("avg", 128)), "inception_4a/"),
    (576, ((192,), (96, 128), (96, 128), ("avg", 128)), "inception_4b/"),
    (576, ((160,), (128, 160), (128, 160), ("avg", 96)), "inception_4c/"),
    (576, ((96,), (128, 192), (160, 192), ("avg", 96)), "inception_4d/"),
    (576, ((0,), (128, 192), (192, 256), ("max", 0)), "inception_4e/"),
    (1024, ((352,), (192, 320), (160, 224), ("avg", 128)), "inception_5a/"),
    (1024, ((352,), (192, 320), (192, 224), ("max", 128)), "inception_5b/"),
]


def _stem(model):
    _cbr(model, 3, 64, 7, 2, 3, "conv1/7x7_s2", bias=False)
    model.add(SpatialMaxPooling(3, 3, 2, 2).ceil().setName("pool1/3x3_s2"))
    _cbr(model, 64, 64, 1, 1, 0, "conv2/3x3_reduce")
    _cbr(model, 64, 192, 3, 1, 1, "conv2/3x3")
    model.add(SpatialMaxPooling(3, 3, 2, 2).ceil().setName("pool2/3x3_s2"))
    return model


def _classifier(seq, class_num, name):
    seq.add(SpatialAveragePooling(7, 7, 1, 1).ceil().setName("pool5/7x7_s1"))
    seq.add(View(1024).setNumInputDims(3))
    seq.add(Linear(1024, class_num).setName(name + "/classifier"))
    seq.add(LogSoftMax().setName(name + "/loss"))
    return seq


def Inception_v2_NoAuxClassifier(class_num: int = 1000):
    model = _stem(Sequential())
    for n_in, cfg, name in _CFG:
        model.add(Inception_Layer_v2(n_in, cfg, name))
    return _classifier(model, class_num, "loss3")


def _aux(n_in, spatial, class_num, name, pool_name):
    out = Sequential()
    out.add(SpatialAveragePooling(5, 5, 3, 3).ceil().setName(pool_name))
    _cbr(out, n_in, 128, 1, 1, 0, name + "/conv")
    out.add(View(128 * spatial * spatial).setNumInputDims(3))
    out.add(Linear(128 * spatial * spatial, 1024).setName(name + "/fc"))
    out.add(ReLU(True).setName(name + "/fc/bn/sc/relu"))
    out.add(Linear(1024, class_num).setName(name + "/classifier"))
    out.add(LogSoftMax().setName(name + "/loss"))
    return out


def Inception_v2(class_num: int = 1000):
    """With the two auxiliary heads; output = concat [loss3 | loss2 | loss1] along dim 2 (Inception_v2.scala:272-360)."""
    features1 = _stem(Sequential())
    for n_in, cfg, name in _CFG[:3]:
        features1.add(Inception_Layer_v2(n_in, cfg, name))
    output1 = _aux(576, 4, class_num, "loss1", "pool3/5x5_s3")
    features2 = Sequential()
    for n_in, cfg, name in _CFG[3:8]:
        features2.add(Inception_Layer_v2(n_in, cfg, name))
    output2 = _aux(1024, 2, class_num, "loss2", "pool4/5x5_s3")
    output3 = Sequential()
    for n_in, cfg, name in _CFG[8:]:
        output3.add(Inception_Layer_v2(n_in, cfg, name))
    _classifier(output3, class_num, "loss3")
    split2 = Concat(2).add(output3).add(output2)
    main_branch = Sequential().add(features2).add(split2)
    split1 = Concat(2).add(main_branch).add(output1)
    return Sequential().add(features1).add(split1)


# graph forms (Inception_v2.scala:229,364)
Inception_v2_NoAuxClassifier.graph = lambda class_num=1000: Inception_v2_NoAuxClassifier(class_num).toGraph()
Inception_v2.graph = lambda class_num=1000, has_dropout=True: Inception_v2(class_num).toGraph()
