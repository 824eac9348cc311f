from bigdl_b200.dataset.mnist import *  # noqa: F401,F403

from bigdl_b200.dataset import mnist as _m  # noqa: E402

TRAIN_MEAN, TRAIN_STD = 0.13066047740239506 * 255, 0.3081078 * 255
TEST_MEAN, TEST_STD = 0.13251460696903547 * 255, 0.31048024 * 255


def load_data(location="/tmp/mnist"):
    """((X_train, Y_train), (X_test, Y_test)) normalised with the training statistics, 1-based labels."""
    (tr_x, tr_y), (te_x, te_y) = _m.read_data_sets(location, "train"), _m.read_data_sets(location, "test")
    norm = lambda a: (a.astype("float32") - TRAIN_MEAN) / TRAIN_STD
    return (norm(tr_x), tr_y + 1), (norm(te_x), te_y + 1)
