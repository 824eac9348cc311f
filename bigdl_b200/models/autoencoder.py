"""MNIST auto-encoder (S/models/autoencoder/Autoencoder.scala): 784 → classNum → 784 with ReLU / Sigmoid."""
from ..nn.layer import Linear, ReLU, Reshape, Sequential, Sigmoid


def Autoencoder(class_num: int = 32):
    model = Sequential()
    model.add(Reshape([28 * 28]))
    model.add(Linear(28 * 28, class_num)).add(ReLU(True))
    model.add(Linear(class_num, 28 * 28)).add(Sigmoid())
    return model


Autoencoder.graph = lambda class_num=32: Autoencoder(class_num).toGraph()      # Autoencoder.scala:38
