"""Keras-style API: shape inference against known Keras-1.2.2 shapes, functional Model, compile/fit/evaluate/predict
(mirrors T/keras/nn/*Spec.scala and pyspark/test/bigdl/keras/test_newapi.py)."""
import numpy as np
import pytest
import torch

import bigdl_b200.nn.keras as K


def _out(layer, shape):
    return tuple(layer.build(list(shape)))


def test_shape_inference_core_layers():
    assert _out(K.Dense(5, activation="relu"), (7,)) == (5,)
    assert _out(K.Dense(5), (3, 7)) == (3, 5)
    assert _out(K.Flatten(), (2, 3, 4)) == (24,)
    assert _out(K.Reshape((4, -1)), (2, 3, 4)) == (4, 6)
    assert _out(K.Permute((2, 1, 3)), (2, 3, 4)) == (3, 2, 4)
    assert _out(K.RepeatVector(4), (6,)) == (4, 6)
    assert _out(K.Embedding(100, 8), (12,)) == (12, 8)
    assert _out(K.BatchNormalization(), (3, 8, 8)) == (3, 8, 8)
    assert _out(K.Highway(), (6,)) == (6,)
    assert _out(K.MaxoutDense(3), (6,)) == (3,)


def test_shape_inference_conv_pool():
    assert _out(K.Convolution2D(8, 3, 3), (3, 16, 16)) == (8, 14, 14)
    assert _out(K.Convolution2D(8, 3, 3, border_mode="same", subsample=(2, 2)), (3, 16, 16)) == (8, 8, 8)
    assert _out(K.Convolution2D(8, 3, 3, dim_ordering="tf"), (16, 16, 3)) == (14, 14, 8)
    assert _out(K.Convolution1D(6, 3), (10, 4)) == (8, 6)
    assert _out(K.Convolution1D(6, 3, border_mode="same"), (10, 4)) == (10, 6)
    assert _out(K.Convolution3D(4, 2, 2, 2), (3, 6, 6, 6)) == (4, 5, 5, 5)
    assert _out(K.AtrousConvolution2D(4, 3, 3, atrous_rate=(2, 2)), (3, 12, 12)) == (4, 8, 8)
    assert _out(K.AtrousConvolution1D(4, 3, atrous_rate=2), (12, 5)) == (8, 4)
    assert _out(K.Deconvolution2D(3, 3, 3, subsample=(2, 2)), (4, 5, 5)) == (3, 11, 11)
    assert _out(K.SeparableConvolution2D(6, 3, 3), (3, 10, 10)) == (6, 8, 8)
    assert _out(K.MaxPooling2D(), (3, 8, 8)) == (3, 4, 4)
    assert _out(K.AveragePooling2D((3, 3), strides=(2, 2), border_mode="same"), (3, 9, 9)) == (3, 5, 5)
    assert _out(K.MaxPooling1D(2), (10, 4)) == (5, 4)
    assert _out(K.AveragePooling3D(), (2, 4, 4, 4)) == (2, 2, 2, 2)
    assert _out(K.GlobalAveragePooling2D(), (5, 7, 7)) == (5,)
    assert _out(K.GlobalMaxPooling1D(), (9, 4)) == (4,)
    assert _out(K.GlobalMaxPooling3D(), (3, 4, 4, 4)) == (3,)
    assert _out(K.ZeroPadding2D((1, 2)), (3, 8, 8)) == (3, 10, 12)
    assert _out(K.ZeroPadding1D(2), (5, 3)) == (9, 3)
    assert _out(K.Cropping2D(((1, 1), (2, 2))), (3, 8, 8)) == (3, 6, 4)
    assert _out(K.Cropping1D((1, 2)), (8, 3)) == (5, 3)
    assert _out(K.UpSampling2D((2, 3)), (3, 4, 4)) == (3, 8, 12)
    assert _out(K.UpSampling1D(3), (4, 5)) == (12, 5)
    assert _out(K.LocallyConnected2D(4, 3, 3), (2, 6, 6)) == (4, 4, 4)


def test_shape_inference_recurrent():
    assert _out(K.LSTM(6), (5, 4)) == (6,)
    assert _out(K.GRU(6, return_sequences=True), (5, 4)) == (5, 6)
    assert _out(K.SimpleRNN(3, go_backwards=True), (5, 4)) == (3,)
    assert _out(K.Bidirectional(K.LSTM(6, return_sequences=True)), (5, 4)) == (5, 12)
    assert _out(K.Bidirectional(K.GRU(6), merge_mode="sum"), (5, 4)) == (6,)
    assert _out(K.TimeDistributed(K.Dense(7)), (5, 4)) == (5, 7)
    assert _out(K.ConvLSTM2D(4, 3, return_sequences=True), (3, 2, 8, 8)) == (3, 4, 8, 8)


def test_sequential_requires_input_shape_and_tracks_shapes():
    m = K.Sequential()
    with pytest.raises(ValueError):
        m.add(K.Dense(3))
    m.add(K.Convolution2D(4, 3, 3, activation="relu", input_shape=(1, 12, 12)))
    m.add(K.MaxPooling2D())
    m.add(K.Flatten())
    m.add(K.Dense(10, activation="softmax"))
    assert m.getOutputShape() == (None, 10) and m.getInputShape() == (None, 1, 12, 12)
    y = m.forward(torch.randn(3, 1, 12, 12))
    assert tuple(y.shape) == (3, 10) and torch.allclose(y.sum(1), torch.ones(3), atol=1e-5)
    with pytest.raises(TypeError):
        m.add(__import__("bigdl_b200.nn.layer", fromlist=["Linear"]).Linear(10, 2))


def test_functional_model_and_merge():
    a = K.Input(shape=(8,))
    b = K.Input(shape=(8,))
    da = K.Dense(6, activation="tanh")(a)
    db = K.Dense(6, activation="tanh")(b)
    merged = K.merge([da, db], mode="concat")
    out = K.Dense(2)(merged)
    model = K.Model([a, b], out)
    assert model.getOutputShape() == (None, 2)
    y = model.forward([torch.randn(4, 8), torch.randn(4, 8)])
    assert tuple(y.shape) == (4, 2)


def test_compile_fit_evaluate_predict():
    from bigdl_b200.utils.random_generator import RNG

    RNG.setSeed(5)
    torch.manual_seed(5)
    rs = np.random.RandomState(0)
    x = rs.randn(128, 4).astype(np.float32)
    y = (x.sum(1) > 0).astype(np.float32) + 1          # 1-based classes
    model = K.Sequential()
    model.add(K.Dense(16, activation="relu", input_shape=(4,)))
    model.add(K.Dense(2, activation="softmax"))
    with pytest.raises(RuntimeError):
        model.fit(x, y)
    model.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    model.fit(x, y, batch_size=16, nb_epoch=40, validation_data=(x, y))
    loss, acc = model.evaluate(x, y, batch_size=32)
    assert acc > 0.9 and loss < 0.6
    pred = model.predict(x, batch_size=32)
    assert len(pred) == 128 and tuple(np.asarray(pred[0]).shape) == (2,)
    with pytest.raises(TypeError):
        model.compile(optimizer="nope", loss="mse")


def test_keras_predict_classes_tensorboard_and_checkpoint(tmp_path):
    import os
    import numpy as np
    import bigdl_b200.nn.keras as K
    from bigdl_b200.utils.random_generator import RNG

    RNG.setSeed(8)
    rs = np.random.RandomState(0)
    x = rs.randn(96, 4).astype(np.float32)
    y = (x[:, 0] > 0).astype(np.float32) + (x[:, 1] > 0).astype(np.float32)          # 3 classes: 0, 1, 2
    m = K.Sequential()
    m.add(K.Dense(16, activation="relu", input_shape=(4,)))
    m.add(K.Dense(3, activation="softmax"))
    m.compile(optimizer="adam", loss="sparse_categorical_crossentropy", metrics=["accuracy"])
    m.set_tensorboard(str(tmp_path / "tb"), "kapp").set_checkpoint(str(tmp_path / "ck"))
    os.makedirs(str(tmp_path / "ck"), exist_ok=True)
    m.fit(x, y + 1, batch_size=16, nb_epoch=100)             # class ids are 1-based for the loss, as in the reference
    zero = m.predict_classes(x, batch_size=32)
    one = m.predict_classes(x, batch_size=32, zero_based_label=False)
    assert len(zero) == 96 and [c + 1 for c in zero] == one and set(zero) <= {0, 1, 2}
    assert float(np.mean(np.asarray(zero) == y)) > 0.8
    assert len(m._train_summary.read_scalar("Loss")) > 0                              # fit() wrote TensorBoard scalars
    assert any(f.startswith("model") for _, _, fs in os.walk(str(tmp_path / "ck")) for f in fs)


def test_remaining_keras_layers_infer_and_produce_the_same_shape():
    """Keras-style layers not covered above: the inferred output shape must equal the shape actually produced."""
    import torch
    import bigdl_b200.nn.keras as K

    cases = [
        (lambda: K.Activation("tanh", input_shape=(5,)), (5,), (5,)),
        (lambda: K.Masking(0.0, input_shape=(4, 3)), (4, 3), (4, 3)),
        (lambda: K.AveragePooling1D(2, input_shape=(6, 3)), (6, 3), (3, 3)),
        (lambda: K.MaxPooling3D(input_shape=(2, 4, 4, 4)), (2, 4, 4, 4), (2, 2, 2, 2)),
        (lambda: K.GlobalAveragePooling1D(input_shape=(6, 3)), (6, 3), (3,)),
        (lambda: K.GlobalMaxPooling2D(input_shape=(3, 4, 4)), (3, 4, 4), (3,)),
        (lambda: K.GlobalAveragePooling3D(input_shape=(2, 4, 4, 4)), (2, 4, 4, 4), (2,)),
        (lambda: K.ZeroPadding3D((1, 1, 1), input_shape=(2, 3, 3, 3)), (2, 3, 3, 3), (2, 5, 5, 5)),
        (lambda: K.Cropping3D(((1, 1), (1, 1), (1, 1)), input_shape=(2, 4, 4, 4)), (2, 4, 4, 4), (2, 2, 2, 2)),
        (lambda: K.UpSampling3D((2, 2, 2), input_shape=(2, 3, 3, 3)), (2, 3, 3, 3), (2, 6, 6, 6)),
        (lambda: K.ThresholdedReLU(0.5, input_shape=(5,)), (5,), (5,)),
        (lambda: K.GaussianNoise(0.1, input_shape=(5,)), (5,), (5,)),
        (lambda: K.GaussianDropout(0.1, input_shape=(5,)), (5,), (5,)),
        (lambda: K.SpatialDropout1D(0.3, input_shape=(6, 3)), (6, 3), (6, 3)),
        (lambda: K.SpatialDropout2D(0.3, input_shape=(3, 4, 4)), (3, 4, 4), (3, 4, 4)),
        (lambda: K.SpatialDropout3D(0.3, input_shape=(2, 4, 4, 4)), (2, 4, 4, 4), (2, 4, 4, 4)),
    ]
    for make, in_shape, out_shape in cases:
        m = K.Sequential()
        m.add(make())
        y = m.forward(torch.randn(2, *in_shape))
        assert tuple(y.shape) == (2,) + out_shape, type(m.modules[0]).__name__
        assert tuple(m.getOutputShape())[1:] == out_shape
    thr = K.Sequential()
    thr.add(K.ThresholdedReLU(0.5, input_shape=(3,)))
    assert torch.equal(thr.forward(torch.tensor([[0.4, 0.6, -1.0]])), torch.tensor([[0.0, 0.6, 0.0]]))


def _hard_sigmoid(x):
    return np.clip(0.2 * x + 0.5, 0.0, 1.0)


def test_keras_recurrent_weight_conversion_matches_keras_math():
    """Keras-1.2.2 LSTM / GRU / SimpleRNN weight lists loaded through WeightLoader reproduce Keras' own recurrences
    (restated in numpy) — gate order and transposes of pyspark/bigdl/keras/converter.py convert_lstm/gru/simplernn."""
    import json as _json

    from bigdl_b200.interop.keras_converter import DefinitionLoader, WeightLoader

    rng = np.random.RandomState(0)
    T_, I, H = 5, 6, 4
    x = rng.randn(3, T_, I).astype(np.float32)

    def build(cls):
        spec = {"class_name": "Sequential", "config": [
            {"class_name": cls, "config": {"name": "r", "output_dim": H, "return_sequences": True,
                                           "batch_input_shape": [None, T_, I], "activation": "tanh",
                                           "inner_activation": "hard_sigmoid"}}]}
        return DefinitionLoader.from_json_str(_json.dumps(spec))

    # LSTM
    W = {g: rng.randn(I, H).astype(np.float32) * 0.5 for g in "icfo"}
    U = {g: rng.randn(H, H).astype(np.float32) * 0.5 for g in "icfo"}
    b = {g: rng.randn(H).astype(np.float32) * 0.1 for g in "icfo"}
    m = build("LSTM")
    WeightLoader.load_weights_from_arrays(m, {"r": [v for g in "icfo" for v in (W[g], U[g], b[g])]})
    h, c, ref = np.zeros((3, H), np.float32), np.zeros((3, H), np.float32), []
    for t in range(T_):
        xt = x[:, t]
        i = _hard_sigmoid(xt @ W["i"] + h @ U["i"] + b["i"])
        f = _hard_sigmoid(xt @ W["f"] + h @ U["f"] + b["f"])
        cc = np.tanh(xt @ W["c"] + h @ U["c"] + b["c"])
        o = _hard_sigmoid(xt @ W["o"] + h @ U["o"] + b["o"])
        c = f * c + i * cc
        h = o * np.tanh(c)
        ref.append(h)
    out = m.forward(torch.from_numpy(x)).detach().numpy()
    assert np.allclose(out, np.stack(ref, 1), atol=1e-5)

    # GRU
    W = {g: rng.randn(I, H).astype(np.float32) * 0.5 for g in "zrh"}
    U = {g: rng.randn(H, H).astype(np.float32) * 0.5 for g in "zrh"}
    b = {g: rng.randn(H).astype(np.float32) * 0.1 for g in "zrh"}
    m = build("GRU")
    WeightLoader.load_weights_from_arrays(m, {"r": [v for g in "zrh" for v in (W[g], U[g], b[g])]})
    h, ref = np.zeros((3, H), np.float32), []
    for t in range(T_):
        xt = x[:, t]
        z = _hard_sigmoid(xt @ W["z"] + h @ U["z"] + b["z"])
        r = _hard_sigmoid(xt @ W["r"] + h @ U["r"] + b["r"])
        hh = np.tanh(xt @ W["h"] + (r * h) @ U["h"] + b["h"])
        h = z * h + (1 - z) * hh
        ref.append(h)
    out = m.forward(torch.from_numpy(x)).detach().numpy()
    assert np.allclose(out, np.stack(ref, 1), atol=1e-5)

    # SimpleRNN
    Wr, Ur, br = rng.randn(I, H).astype(np.float32), rng.randn(H, H).astype(np.float32) * 0.3, rng.randn(H).astype(np.float32)
    m = build("SimpleRNN")
    WeightLoader.load_weights_from_arrays(m, {"r": [Wr, Ur, br]})
    h, ref = np.zeros((3, H), np.float32), []
    for t in range(T_):
        h = np.tanh(x[:, t] @ Wr + h @ Ur + br)
        ref.append(h)
    assert np.allclose(m.forward(torch.from_numpy(x)).detach().numpy(), np.stack(ref, 1), atol=1e-5)


def test_keras_converter_covers_the_reference_layer_list():
    import json as _json
    import os
    import re

    from bigdl_b200.interop.keras_converter import _BUILDERS, DefinitionLoader

    ref = "/root/reference/pyspark/bigdl/keras/converter.py"
    if os.path.exists(ref):
        names = set(re.findall(r"def create_([a-z0-9]+)\(", open(ref).read())) - {"model", "sequential"}
        have = {k.lower() for k in _BUILDERS} | {"timedistributed", "bidirectional", "inputlayer", "merge"}
        assert not (names - have), sorted(names - have)
    # a few of the newly covered ones build and run
    spec = {"class_name": "Sequential", "config": [
        {"class_name": "TimeDistributedDense", "config": {"name": "td", "output_dim": 3, "batch_input_shape": [None, 4, 5]}},
        {"class_name": "PReLU", "config": {"name": "p"}},
        {"class_name": "ParametricSoftplus", "config": {"name": "ps"}},
        {"class_name": "ActivityRegularization", "config": {"name": "ar", "l1": 0.01}}]}
    m = DefinitionLoader.from_json_str(_json.dumps(spec))
    y = m.forward(torch.randn(2, 4, 5))
    assert tuple(y.shape) == (2, 4, 3)
