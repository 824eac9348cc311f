"""P/models/local_lenet/local_lenet.py: LeNet-5 trained from in-memory ndarrays with the LocalOptimizer."""
import numpy as np

from bigdl_b200.dataset import mnist
from bigdl_b200.models.lenet import LeNet5


def build_model(class_num):
    return LeNet5(class_num)


def get_mnist(data_type="train", location="/tmp/mnist"):
    X, Y = mnist.read_data_sets(location, data_type)
    return X.astype(np.float32).reshape(-1, 1, 28, 28), Y + 1           # labels are 1-based


def main(location="/tmp/mnist", max_epoch=20, batch_size=128):
    from bigdl.nn.criterion import ClassNLLCriterion
    from bigdl.optim.optimizer import MaxEpoch, Optimizer, SGD

    X, Y = get_mnist("train", location)
    X = (X - mnist.TRAIN_MEAN) / mnist.TRAIN_STD
    model = build_model(10)
    opt = Optimizer.create(model=model, training_set=(X, Y), criterion=ClassNLLCriterion(),
                           optim_method=SGD(learningrate=0.01, learningrate_decay=0.0002),
                           end_trigger=MaxEpoch(max_epoch), batch_size=batch_size)
    opt.optimize()
    return model


if __name__ == "__main__":
    main()
