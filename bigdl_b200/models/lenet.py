"""LeNet-5 (S/models/lenet/LeNet5.scala:26-60): conv 1→6 5x5, conv 6→12 5x5, fc 192→100, fc 100→classNum; 22,278
parameters for 10 classes."""
from ..nn.layer import (Graph, Linear, LogSoftMax, Reshape, Sequential, SpatialConvolution, SpatialMaxPooling, Tanh)


def LeNet5(class_num: int = 10):
    model = Sequential()
    model.add(Reshape([1, 28, 28]))
    model.add(SpatialConvolution(1, 6, 5, 5).setName("conv1_5x5"))
    model.add(Tanh())
    model.add(SpatialMaxPooling(2, 2, 2, 2))
    model.add(SpatialConvolution(6, 12, 5, 5).setName("conv2_5x5"))
    model.add(Tanh())
    model.add(SpatialMaxPooling(2, 2, 2, 2))
    model.add(Reshape([12 * 4 * 4]))
    model.add(Linear(12 * 4 * 4, 100).setName("fc1"))
    model.add(Tanh())
    model.add(Linear(100, class_num).setName("fc2"))
    model.add(LogSoftMax())
    return model


def LeNet5_graph(class_num: int = 10):
    inp = Reshape([1, 28, 28]).inputs()
    conv1 = SpatialConvolution(1, 6, 5, 5).setName("conv1_5x5").inputs(inp)
    tanh1 = Tanh().inputs(conv1)
    pool1 = SpatialMaxPooling(2, 2, 2, 2).inputs(tanh1)
    conv2 = SpatialConvolution(6, 12, 5, 5).setName("conv2_5x5").inputs(pool1)
    tanh2 = Tanh().inputs(conv2)
    pool2 = SpatialMaxPooling(2, 2, 2, 2).inputs(tanh2)
    reshape = Reshape([12 * 4 * 4]).inputs(pool2)
    fc1 = Linear(12 * 4 * 4, 100).setName("fc1").inputs(reshape)
    tanh3 = Tanh().inputs(fc1)
    fc2 = Linear(100, class_num).setName("fc2").inputs(tanh3)
    out = LogSoftMax().inputs(fc2)
    return Graph(inp, out)
