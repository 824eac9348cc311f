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


def LeNet5_keras(class_num: int = 10):
    """The Keras-style Sequential form (LeNet5.scala:59-72)."""
    from ..nn.keras import layers as K
    from ..nn.keras.topology import Sequential as KSequential

    m = KSequential()
    m.add(K.Reshape([1, 28, 28], input_shape=(28, 28, 1)))
    m.add(K.Convolution2D(6, 5, 5, activation="tanh").setName("conv1_5x5"))
    m.add(K.MaxPooling2D())
    m.add(K.Convolution2D(12, 5, 5, activation="tanh").setName("conv2_5x5"))
    m.add(K.MaxPooling2D())
    m.add(K.Flatten())
    m.add(K.Dense(100, activation="tanh").setName("fc1"))
    m.add(K.Dense(class_num, activation="softmax").setName("fc2"))
    return m


def LeNet5_kerasGraph(class_num: int = 10):
    """The Keras-style functional form (LeNet5.scala:74-88)."""
    from ..nn.keras import layers as K
    from ..nn.keras.topology import Model as KModel

    inp = K.Input(shape=(28, 28, 1))
    x = K.Reshape([1, 28, 28]).inputs(inp)
    x = K.Convolution2D(6, 5, 5, activation="tanh").setName("conv1_5x5").inputs(x)
    x = K.MaxPooling2D().inputs(x)
    x = K.Convolution2D(12, 5, 5, activation="tanh").setName("conv2_5x5").inputs(x)
    x = K.MaxPooling2D().inputs(x)
    x = K.Flatten().inputs(x)
    x = K.Dense(100, activation="tanh").setName("fc1").inputs(x)
    out = K.Dense(class_num, activation="softmax").setName("fc2").inputs(x)
    return KModel(inp, out)


def _lenet_dnn(batch_size: int, class_num: int = 10):
    """``LeNet5.dnn`` / ``dnnGraph`` (LeNet5.scala:90-125) build the same network from MKL-DNN primitives with a
    fixed batch size and a bare soft-max-free head (trained with CrossEntropyCriterion). The device kernels here are
    chosen per layer at run time, so these are the ordinary builders minus the LogSoftMax."""
    m = LeNet5(class_num)
    m.modules.pop()
    return m


# companion-object spellings: LeNet5.graph(10), LeNet5.keras(10), …
LeNet5.graph = LeNet5_graph
LeNet5.keras = LeNet5_keras
LeNet5.kerasGraph = LeNet5_kerasGraph
LeNet5.dnn = _lenet_dnn
LeNet5.dnnGraph = lambda batch_size, class_num=10: _lenet_dnn(batch_size, class_num).toGraph()
