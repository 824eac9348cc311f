"""``with_bigdl_backend`` (P/keras/backend.py) and the helper translators (P/keras/{ToBigDLHelper,optimization}.py) on a
duck-typed Keras-1.2.2 model: definition from ``to_json()``, weights from ``layers[i].get_weights()``, loss / optimizer /
metrics from the compiled attributes."""
import json

import numpy as np
import pytest
import torch

from bigdl.keras.backend import with_bigdl_backend
from bigdl.keras.optimization import OptimConverter
from bigdl.keras.ToBigDLHelper import (to_bigdl_2d_ordering, to_bigdl_2d_padding, to_bigdl_3d_padding, to_bigdl_init,
                                       to_bigdl_reg)


class _Layer:
    def __init__(self, name, weights):
        self.name, self._w = name, weights

    def get_weights(self):
        return self._w


class _KModel:
    """What Keras 1.2.2 exposes for Sequential([Dense(8, relu, input_dim=4), Dense(3, softmax)]) after compile."""

    def __init__(self, rs):
        self.w1, self.b1 = rs.randn(4, 8).astype(np.float32), rs.randn(8).astype(np.float32)
        self.w2, self.b2 = rs.randn(8, 3).astype(np.float32), rs.randn(3).astype(np.float32)
        self.layers = [_Layer("dense_1", [self.w1, self.b1]), _Layer("dense_2", [self.w2, self.b2])]
        self.loss = "categorical_crossentropy"
        self.optimizer = {"class_name": "SGD", "config": {"lr": 0.5, "momentum": 0.0, "decay": 0.0, "nesterov": False}}
        self.metrics = ["accuracy"]

    def to_json(self):
        def dense(name, out, act, **extra):
            cfg = {"name": name, "output_dim": out, "activation": act, "init": "glorot_uniform", "bias": True,
                   "W_regularizer": None, "b_regularizer": None, "input_dim": None, "trainable": True}
            cfg.update(extra)
            return {"class_name": "Dense", "config": cfg}
        return json.dumps({"class_name": "Sequential", "keras_version": "1.2.2",
                           "config": [dense("dense_1", 8, "relu", batch_input_shape=[None, 4], input_dim=4),
                                      dense("dense_2", 3, "softmax")]})

    def numpy_forward(self, x):
        h = np.maximum(x @ self.w1 + self.b1, 0)
        z = h @ self.w2 + self.b2
        e = np.exp(z - z.max(1, keepdims=True))
        return e / e.sum(1, keepdims=True)


def test_wrapper_predicts_like_keras_and_trains():
    rs = np.random.RandomState(0)
    k = _KModel(rs)
    m = with_bigdl_backend(k)
    x = rs.randn(64, 4).astype(np.float32)
    assert np.allclose(m.predict(x), k.numpy_forward(x), atol=1e-5)       # definition + converted weights
    labels = (x[:, 0] + x[:, 1] > 0).astype(int) + (x[:, 2] > 1).astype(int)
    y = np.eye(3, dtype=np.float32)[labels]
    before = float(-(np.log(m.predict(x) + 1e-9) * y).sum(1).mean())
    m.fit(x, y, batch_size=16, nb_epoch=30)
    after = float(-(np.log(m.predict(x) + 1e-9) * y).sum(1).mean())
    assert after < 0.6 * before
    with pytest.raises(Exception):
        m.fit(x, y, validation_split=0.1)


def test_optim_converter_and_helpers():
    from bigdl_b200.nn import criterion as C
    from bigdl_b200.optim import optim_method as O
    assert isinstance(OptimConverter.to_bigdl_criterion("mse"), C.MSECriterion)
    assert isinstance(OptimConverter.to_bigdl_criterion("sparse_categorical_crossentropy"), C.ClassNLLCriterion)
    assert isinstance(OptimConverter.to_bigdl_criterion("kld"), C.KullbackLeiblerDivergenceCriterion)
    with pytest.raises(Exception):
        OptimConverter.to_bigdl_criterion("nope")

    class Adam:                                   # a Keras optimizer object: attributes, not a dict
        lr, beta_1, beta_2, epsilon, decay = 0.002, 0.8, 0.99, 1e-7, 0.0
    a = OptimConverter.to_bigdl_optim_method(Adam())
    assert isinstance(a, O.Adam) and a.learningRate == 0.002 and a.beta1 == 0.8
    s = OptimConverter.to_bigdl_optim_method({"name": "SGD", "lr": 0.1, "momentum": 0.9, "nesterov": True})
    assert isinstance(s, O.SGD) and s.momentum == 0.9 and s.nesterov
    assert isinstance(OptimConverter.to_bigdl_optim_method("rmsprop"), O.RMSprop)
    assert len(OptimConverter.to_bigdl_metrics(["accuracy"])) == 1
    assert to_bigdl_2d_ordering("tf") == "NHWC" and to_bigdl_3d_padding("same") == (-1, -1, -1)
    assert to_bigdl_2d_padding("valid") == (0, 0) and to_bigdl_2d_padding("same") == (-1, -1)
    assert to_bigdl_2d_padding("same", 28, 3, 1, 1, 28, 5, 1, 1) == (1, 2)
    assert type(to_bigdl_init("glorot_uniform")).__name__ == "Xavier"
    r = to_bigdl_reg({"l1": 0.0, "l2": 0.01})
    assert r.l2 == 0.01 and to_bigdl_reg(None) is None


def _klayer(cls_name, config, weights, input_shape):
    return type(cls_name, (), {"name": config.get("name", cls_name.lower()), "get_config": lambda self: dict(config),
                               "get_weights": lambda self: weights, "input_shape": input_shape})()


def test_per_layer_converters():
    import numpy as np
    import pytest
    import torch

    from bigdl.keras.converter import LayerConverter, WeightsConverter, unsupport_exp

    rs = np.random.RandomState(0)
    f = lambda *s: rs.randn(*s).astype(np.float32)  # noqa: E731
    dense = _klayer("Dense", {"output_dim": 3, "activation": "relu", "bias": True}, [f(4, 3), f(3)], (None, 4))
    w = WeightsConverter.convert_dense(dense, dense.get_weights())
    assert np.allclose(w[0], dense.get_weights()[0].T) and np.allclose(w[1], dense.get_weights()[1])
    layer = LayerConverter(dense).create()
    x = torch.rand(2, 4)
    layer.build([4])
    from bigdl_b200.interop.keras_converter import WeightLoader
    WeightLoader._apply(layer, dense.get_weights())
    want = torch.relu(x @ torch.from_numpy(dense.get_weights()[0]) + torch.from_numpy(dense.get_weights()[1]))
    assert torch.allclose(layer.forward(x), want, atol=1e-6)

    conv = _klayer("Convolution2D", {"nb_filter": 4, "nb_row": 3, "nb_col": 3, "dim_ordering": "tf"},
                   [f(3, 3, 2, 4), f(4)], (None, 8, 8, 2))
    cw = WeightsConverter.to_bigdl_weights(conv, conv.get_weights())
    assert cw[0].shape == (4, 2, 3, 3) and np.allclose(cw[0], conv.get_weights()[0].transpose(3, 2, 0, 1))
    lstm = _klayer("LSTM", {"output_dim": 2, "return_sequences": True}, [f(*s) for s in [(3, 2), (2, 2), (2,)] * 4],
                   (None, 5, 3))
    lw = WeightsConverter.get_bigdl_weights_from_klayer(lstm)
    ks = lstm.get_weights()
    assert [a.shape for a in lw] == [(8, 3), (8,), (8, 2)]
    assert np.allclose(lw[0][:2], ks[0].T) and np.allclose(lw[0][2:4], ks[3].T) and np.allclose(lw[2][4:6], ks[7].T)
    bn = _klayer("BatchNormalization", {"epsilon": 1e-3, "mode": 0, "axis": 1}, [f(3), f(3), f(3), np.abs(f(3))], (None, 3))
    bw = WeightsConverter.convert_batchnormalization(bn, bn.get_weights())
    assert len(bw) == 4 and np.allclose(bw[2], bn.get_weights()[2]) and np.allclose(bw[3], bn.get_weights()[3])
    model = type("M", (), {"layers": [dense, _klayer("Activation", {"activation": "tanh"}, [], (None, 3)), bn]})()
    assert len(WeightsConverter.get_weights_from_kmodel(model)) == 2 + 4
    with pytest.raises(ValueError):
        WeightsConverter.convert_lstm(dense, dense.get_weights())
    with pytest.raises(Exception, match="don't support Foo"):
        WeightsConverter.get_converter("Foo")
    with pytest.raises(Exception):
        unsupport_exp("x")

    lc = LayerConverter(lstm)
    assert type(lc.generate_lstm_cell()).__name__ == "LSTM" and lc.get_bdim_order() == "NCHW"
    assert LayerConverter(conv).get_bdim_order() == "NHWC" and LayerConverter(conv).get_bdim_order("3D") == "CHANNEL_LAST"
    with pytest.raises(ValueError):
        LayerConverter(dense).create_lstm()
    assert LayerConverter.get_value_from_init("one", (2,)).tolist() == [1.0, 1.0]
    fused = LayerConverter(dense).combo_parameter_layer(LayerConverter(dense).create(), {"name": "d1", "activation": "tanh"})
    assert len(fused.modules) == 2
    merged = LayerConverter({"class_name": "Merge", "config": {"mode": "concat", "concat_axis": 1, "name": "m"}}).create()
    assert type(merged).__name__ == "Merge"
    for t in ("Dense", "Dropout", "Flatten", "Convolution1D", "MaxPooling2D", "GRU", "Highway", "SReLU", "Cropping3D"):
        assert hasattr(LayerConverter, "create_" + t.lower())
