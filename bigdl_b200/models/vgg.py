"""VGG: VggForCifar10, Vgg_16, Vgg_19 (S/models/vgg/VggForCifar10.scala:26-131: conv-BN-ReLU blocks for CIFAR; plain
conv-ReLU stacks + 4096-4096 classifier for ImageNet, 138M parameters for Vgg_16)."""
from ..nn.layer import (BatchNormalization, Dropout, Linear, LogSoftMax, ReLU, Sequential, SpatialBatchNormalization,
                        SpatialConvolution, SpatialMaxPooling, Threshold, View)


def VggForCifar10(class_num: int = 10, has_dropout: bool = True):
    model = Sequential()

    def conv_bn_relu(n_in, n_out):
        model.add(SpatialConvolution(n_in, n_out, 3, 3, 1, 1, 1, 1))
        model.add(SpatialBatchNormalization(n_out, 1e-3))
        model.add(ReLU(True))

    conv_bn_relu(3, 64)
    if has_dropout:
        model.add(Dropout(0.3))
    conv_bn_relu(64, 64)
    model.add(SpatialMaxPooling(2, 2, 2, 2).ceil())
    for n_in, n_out, reps in ((64, 128, 2), (128, 256, 3), (256, 512, 3), (512, 512, 3)):
        for r in range(reps):
            conv_bn_relu(n_in if r == 0 else n_out, n_out)
            if has_dropout and r < reps - 1:
                model.add(Dropout(0.4))
        model.add(SpatialMaxPooling(2, 2, 2, 2).ceil())
    model.add(View(512))
    classifier = Sequential()
    if has_dropout:
        classifier.add(Dropout(0.5))
    classifier.add(Linear(512, 512)).add(BatchNormalization(512)).add(ReLU(True))
    if has_dropout:
        classifier.add(Dropout(0.5))
    classifier.add(Linear(512, class_num)).add(LogSoftMax())
    model.add(classifier)
    return model


def _vgg(cfg, class_num, has_dropout):
    model = Sequential()
    n_in = 3
    for v in cfg:
        if v == "M":
            model.add(SpatialMaxPooling(2, 2, 2, 2))
        else:
            model.add(SpatialConvolution(n_in, v, 3, 3, 1, 1, 1, 1))
            model.add(ReLU(True))
            n_in = v
    model.add(View(512 * 7 * 7))
    model.add(Linear(512 * 7 * 7, 4096)).add(Threshold(0, 1e-6))
    if has_dropout:
        model.add(Dropout(0.5))
    model.add(Linear(4096, 4096)).add(Threshold(0, 1e-6))
    if has_dropout:
        model.add(Dropout(0.5))
    model.add(Linear(4096, class_num)).add(LogSoftMax())
    return model


def Vgg_16(class_num: int = 1000, has_dropout: bool = True):
    return _vgg([64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"], class_num,
                has_dropout)


def Vgg_19(class_num: int = 1000, has_dropout: bool = True):
    return _vgg([64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"],
                class_num, has_dropout)


# ``X.graph(classNum, hasDropout)`` of the reference's companion objects (VggForCifar10.scala:79,183,293): the same
# network as a Graph
VggForCifar10.graph = lambda class_num=10, has_dropout=True: VggForCifar10(class_num, has_dropout).toGraph()
Vgg_16.graph = lambda class_num=1000, has_dropout=True: Vgg_16(class_num, has_dropout).toGraph()
Vgg_19.graph = lambda class_num=1000, has_dropout=True: Vgg_19(class_num, has_dropout).toGraph()
