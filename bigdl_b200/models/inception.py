"""GoogLeNet / Inception-v1 (with and without the two auxiliary classifiers) and Inception-v2 (BN-Inception).

Reference: S/models/inception/Inception_v1.scala:25-193 (layer block :26-104, NoAux :106-141, with aux :195-290),
Inception_v2.scala. Layer names, init methods (Xavier weights, bias 0.1) and hyper-parameters match so that
weights saved by the reference load by name.
"""
from ..nn.layer import (Concat, ConstInitMethod, Dropout, Linear, LogSoftMax, ReLU, Sequential,
                        SpatialAveragePooling, SpatialBatchNormalization, SpatialConvolution, SpatialCrossMapLRN,
                        SpatialMaxPooling, View, Xavier, Zeros)


def _conv(n_in, n_out, kw, kh, sw=1, sh=1, pw=0, ph=0, name="", propagate_back=True):
    return (SpatialConvolution(n_in, n_out, kw, kh, sw, sh, pw, ph, 1, propagate_back)
            .setInitMethod(Xavier(), ConstInitMethod(0.1)).setName(name))


def Inception_Layer_v1(input_size, config, name_prefix=""):
    """config = ((c1x1,), (c3x3_reduce, c3x3), (c5x5_reduce, c5x5), (pool_proj,))"""
    concat = Concat(2)
    conv1 = Sequential()
    conv1.add(_conv(input_size, config[0][0], 1, 1, name=name_prefix + "1x1"))
    conv1.add(ReLU(True).setName(name_prefix + "relu_1x1"))
    concat.add(conv1)
    conv3 = Sequential()
    conv3.add(_conv(input_size, config[1][0], 1, 1, name=name_prefix + "3x3_reduce"))
    conv3.add(ReLU(True).setName(name_prefix + "relu_3x3_reduce"))
    conv3.add(_conv(config[1][0], config[1][1], 3, 3, 1, 1, 1, 1, name=name_prefix + "3x3"))
    conv3.add(ReLU(True).setName(name_prefix + "relu_3x3"))
    concat.add(conv3)
    conv5 = Sequential()
    conv5.add(_conv(input_size, config[2][0], 1, 1, name=name_prefix + "5x5_reduce"))
    conv5.add(ReLU(True).setName(name_prefix + "relu_5x5_reduce"))
    conv5.add(_conv(config[2][0], config[2][1], 5, 5, 1, 1, 2, 2, name=name_prefix + "5x5"))
    conv5.add(ReLU(True).setName(name_prefix + "relu_5x5"))
    concat.add(conv5)
    pool = Sequential()
    pool.add(SpatialMaxPooling(3, 3, 1, 1, 1, 1).ceil().setName(name_prefix + "pool"))
    pool.add(_conv(input_size, config[3][0], 1, 1, name=name_prefix + "pool_proj"))
    pool.add(ReLU(True).setName(name_prefix + "relu_pool_proj"))
    concat.add(pool)
    concat.setName(name_prefix + "output")
    return concat


def _stem(model):
    model.add(_conv(3, 64, 7, 7, 2, 2, 3, 3, name="conv1/7x7_s2", propagate_back=False))
    model.add(ReLU(True).setName("conv1/relu_7x7"))
    model.add(SpatialMaxPooling(3, 3, 2, 2).ceil().setName("pool1/3x3_s2"))
    model.add(SpatialCrossMapLRN(5, 0.0001, 0.75).setName("pool1/norm1"))
    model.add(_conv(64, 64, 1, 1, name="conv2/3x3_reduce"))
    model.add(ReLU(True).setName("conv2/relu_3x3_reduce"))
    model.add(_conv(64, 192, 3, 3, 1, 1, 1, 1, name="conv2/3x3"))
    model.add(ReLU(True).setName("conv2/relu_3x3"))
    model.add(SpatialCrossMapLRN(5, 0.0001, 0.75).setName("conv2/norm2"))
    model.add(SpatialMaxPooling(3, 3, 2, 2).ceil().setName("pool2/3x3_s2"))
    return model


def Inception_v1_NoAuxClassifier(class_num: int = 1000, has_dropout: bool = True):
    model = _stem(Sequential())
    model.add(Inception_Layer_v1(192, ((64,), (96, 128), (16, 32), (32,)), "inception_3a/"))
    model.add(Inception_Layer_v1(256, ((128,), (128, 192), (32, 96), (64,)), "inception_3b/"))
    model.add(SpatialMaxPooling(3, 3, 2, 2).ceil().setName("pool3/3x3_s2"))
    model.add(Inception_Layer_v1(480, ((192,), (96, 208), (16, 48), (64,)), "inception_4a/"))
    model.add(Inception_Layer_v1(512, ((160,), (112, 224), (24, 64), (64,)), "inception_4b/"))
    model.add(Inception_Layer_v1(512, ((128,), (128, 256), (24, 64), (64,)), "inception_4c/"))
    model.add(Inception_Layer_v1(512, ((112,), (144, 288), (32, 64), (64,)), "inception_4d/"))
    model.add(Inception_Layer_v1(528, ((256,), (160, 320), (32, 128), (128,)), "inception_4e/"))
    model.add(SpatialMaxPooling(3, 3, 2, 2).ceil().setName("pool4/3x3_s2"))
    model.add(Inception_Layer_v1(832, ((256,), (160, 320), (32, 128), (128,)), "inception_5a/"))
    model.add(Inception_Layer_v1(832, ((384,), (192, 384), (48, 128), (128,)), "inception_5b/"))
    model.add(SpatialAveragePooling(7, 7, 1, 1).setName("pool5/7x7_s1"))
    if has_dropout:
        model.add(Dropout(0.4).setName("pool5/drop_7x7_s1"))
    model.add(View(1024).setNumInputDims(3))
    model.add(Linear(1024, class_num).setInitMethod(Xavier(), Zeros()).setName("loss3/classifier"))
    model.add(LogSoftMax().setName("loss3/loss3"))
    return model


def _aux_head(n_in, class_num, prefix, has_dropout, ceil):
    out = Sequential()
    pool = SpatialAveragePooling(5, 5, 3, 3)
    out.add((pool.ceil() if ceil else pool).setName(prefix + "/ave_pool"))
    out.add(SpatialConvolution(n_in, 128, 1, 1, 1, 1).setName(prefix + "/conv"))
    out.add(ReLU(True).setName(prefix + "/relu_conv"))
    out.add(View(128 * 4 * 4).setNumInputDims(3))
    out.add(Linear(128 * 4 * 4, 1024).setName(prefix + "/fc"))
    out.add(ReLU(True).setName(prefix + "/relu_fc"))
    if has_dropout:
        out.add(Dropout(0.7).setName(prefix + "/drop_fc"))
    out.add(Linear(1024, class_num).setName(prefix + "/classifier"))
    out.add(LogSoftMax().setName(prefix + "/loss"))
    return out


def Inception_v1(class_num: int = 1000, has_dropout: bool = True):
    """With the two auxiliary heads; output is the concatenation [loss3 | loss2 | loss1] along dim 2."""
    feature1 = _stem(Sequential())
    feature1.add(Inception_Layer_v1(192, ((64,), (96, 128), (16, 32), (32,)), "inception_3a/"))
    feature1.add(Inception_Layer_v1(256, ((128,), (128, 192), (32, 96), (64,)), "inception_3b/"))
    feature1.add(SpatialMaxPooling(3, 3, 2, 2).ceil().setName("pool3/3x3_s2"))
    feature1.add(Inception_Layer_v1(480, ((192,), (96, 208), (16, 48), (64,)), "inception_4a/"))
    output1 = _aux_head(512, class_num, "loss1", has_dropout, ceil=True)
    feature2 = Sequential()
    feature2.add(Inception_Layer_v1(512, ((160,), (112, 224), (24, 64), (64,)), "inception_4b/"))
    feature2.add(Inception_Layer_v1(512, ((128,), (128, 256), (24, 64), (64,)), "inception_4c/"))
    feature2.add(Inception_Layer_v1(512, ((112,), (144, 288), (32, 64), (64,)), "inception_4d/"))
    output2 = _aux_head(528, class_num, "loss2", has_dropout, ceil=False)
    output3 = Sequential()
    output3.add(Inception_Layer_v1(528, ((256,), (160, 320), (32, 128), (128,)), "inception_4e/"))
    output3.add(SpatialMaxPooling(3, 3, 2, 2).ceil().setName("pool4/3x3_s2"))
    output3.add(Inception_Layer_v1(832, ((256,), (160, 320), (32, 128), (128,)), "inception_5a/"))
    output3.add(Inception_Layer_v1(832, ((384,), (192, 384), (48, 128), (128,)), "inception_5b/"))
    output3.add(SpatialAveragePooling(7, 7, 1, 1).setName("pool5/7x7_s1"))
    if has_dropout:
        output3.add(Dropout(0.4).setName("pool5/drop_7x7_s1"))
    output3.add(View(1024).setNumInputDims(3))
    output3.add(Linear(1024, class_num).setInitMethod(Xavier(), Zeros()).setName("loss3/classifier"))
    output3.add(LogSoftMax().setName("loss3/loss3"))
    split2 = Concat(2).setName("split2")
    split2.add(output3)
    split2.add(output2)
    main_branch = Sequential()
    main_branch.add(feature2)
    main_branch.add(split2)
    split1 = Concat(2).setName("split1")
    split1.add(main_branch)
    split1.add(output1)
    model = Sequential()
    model.add(feature1)
    model.add(split1)
    return model


# graph forms (Inception_v1.scala:144,279)
Inception_v1_NoAuxClassifier.graph = lambda class_num=1000, has_dropout=True: \
    Inception_v1_NoAuxClassifier(class_num, has_dropout).toGraph()
Inception_v1.graph = lambda class_num=1000, has_dropout=True: Inception_v1(class_num, has_dropout).toGraph()
