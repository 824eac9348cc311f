"""P/models/lenet/utils.py: MNIST preprocessing helpers."""
import numpy as np

from bigdl_b200.dataset import mnist
from bigdl_b200.dataset.base import Sample


def get_mnist(sc=None, data_type="train", location="/tmp/mnist"):
    """→ list of (image ndarray [28,28,1], label + 1) records (an RDD in the reference)."""
    images, labels = mnist.read_data_sets(location, data_type)
    return [(images[i], int(labels[i]) + 1) for i in range(len(images))]


def preprocess_mnist(sc=None, options=None, location="/tmp/mnist"):
    def norm(records, mean, std):
        return [Sample((img.astype(np.float32).reshape(1, 28, 28) - mean) / std, np.array([lab], dtype=np.float32))
                for img, lab in records]
    train = norm(get_mnist(sc, "train", location), mnist.TRAIN_MEAN, mnist.TRAIN_STD)
    test = norm(get_mnist(sc, "test", location), mnist.TEST_MEAN, mnist.TEST_STD)
    return train, test
