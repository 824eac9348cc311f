"""Keras-1.2.2 model import (P/keras/converter.py:1-1759: ``DefinitionLoader`` builds a model from the JSON definition,
``WeightLoader`` copies HDF5 weights, layer by layer, converting Keras' weight layouts).

``load_keras(json_path, hdf5_path)`` mirrors ``Model.load_keras``. The definition side has no dependencies. HDF5 needs
``h5py``, which this image does not ship, so ``WeightLoader.load_weights_from_hdf5`` raises a clear error without it;
``load_weights_from_arrays`` takes {layer name: [arrays]} (e.g. from ``np.load`` of an exported ``.npz``) and applies
the same layout conversions."""
from __future__ import annotations

import json

import numpy as np
import torch

from ..nn import keras as K
from ..nn import layer as L


def _shape(cfg):
    bis = cfg.get("batch_input_shape")
    return tuple(bis[1:]) if bis else None


def _common(cfg):
    kw = {}
    s = _shape(cfg)
    if s is not None:
        kw["input_shape"] = s
    if cfg.get("name"):
        kw["name"] = cfg["name"]
    return kw


def _conv_common(cfg):
    return dict(init=cfg.get("init", "glorot_uniform"), activation=_act(cfg), border_mode=cfg.get("border_mode", "valid"),
                bias=cfg.get("bias", True))


def _act(cfg):
    a = cfg.get("activation")
    return None if a in (None, "linear") else a


_BUILDERS = {
    "Dense": lambda c: K.Dense(c["output_dim"], init=c.get("init", "glorot_uniform"), activation=_act(c),
                               bias=c.get("bias", True), **_common(c)),
    "Activation": lambda c: K.Activation(c["activation"], **_common(c)),
    "Dropout": lambda c: K.Dropout(c["p"], **_common(c)),
    "Flatten": lambda c: K.Flatten(**_common(c)),
    "Reshape": lambda c: K.Reshape(c["target_shape"], **_common(c)),
    "Permute": lambda c: K.Permute(c["dims"], **_common(c)),
    "RepeatVector": lambda c: K.RepeatVector(c["n"], **_common(c)),
    "Masking": lambda c: K.Masking(c.get("mask_value", 0.0), **_common(c)),
    "Highway": lambda c: K.Highway(activation=_act(c), bias=c.get("bias", True), **_common(c)),
    "MaxoutDense": lambda c: K.MaxoutDense(c["output_dim"], c.get("nb_feature", 4), bias=c.get("bias", True), **_common(c)),
    "Embedding": lambda c: K.Embedding(c["input_dim"], c["output_dim"], init=c.get("init", "uniform"),
                                       **({"input_shape": (c["input_length"],)} if c.get("input_length") else {}),
                                       **({"name": c["name"]} if c.get("name") else {})),
    "BatchNormalization": lambda c: K.BatchNormalization(c.get("epsilon", 1e-3), c.get("mode", 0), c.get("axis", 1),
                                                         c.get("momentum", 0.99), **_common(c)),
    "Convolution1D": lambda c: K.Convolution1D(c["nb_filter"], c["filter_length"], subsample_length=c.get("subsample_length", 1),
                                               **_conv_common(c), **_common(c)),
    "Convolution2D": lambda c: K.Convolution2D(c["nb_filter"], c["nb_row"], c["nb_col"], subsample=tuple(c.get("subsample", (1, 1))),
                                               dim_ordering=c.get("dim_ordering", "th"), **_conv_common(c), **_common(c)),
    "Convolution3D": lambda c: K.Convolution3D(c["nb_filter"], c["kernel_dim1"], c["kernel_dim2"], c["kernel_dim3"],
                                               subsample=tuple(c.get("subsample", (1, 1, 1))), **_conv_common(c), **_common(c)),
    "AtrousConvolution2D": lambda c: K.AtrousConvolution2D(c["nb_filter"], c["nb_row"], c["nb_col"], activation=_act(c),
                                                           subsample=tuple(c.get("subsample", (1, 1))),
                                                           atrous_rate=tuple(c.get("atrous_rate", (1, 1))), **_common(c)),
    "Deconvolution2D": lambda c: K.Deconvolution2D(c["nb_filter"], c["nb_row"], c["nb_col"], c.get("output_shape"),
                                                   activation=_act(c), subsample=tuple(c.get("subsample", (1, 1))),
                                                   bias=c.get("bias", True), **_common(c)),
    "SeparableConvolution2D": lambda c: K.SeparableConvolution2D(c["nb_filter"], c["nb_row"], c["nb_col"], activation=_act(c),
                                                                 border_mode=c.get("border_mode", "valid"),
                                                                 subsample=tuple(c.get("subsample", (1, 1))),
                                                                 depth_multiplier=c.get("depth_multiplier", 1),
                                                                 dim_ordering=c.get("dim_ordering", "th"), **_common(c)),
    "MaxPooling1D": lambda c: K.MaxPooling1D(c.get("pool_length", 2), c.get("stride"), c.get("border_mode", "valid"), **_common(c)),
    "AveragePooling1D": lambda c: K.AveragePooling1D(c.get("pool_length", 2), c.get("stride"), c.get("border_mode", "valid"), **_common(c)),
    "MaxPooling2D": lambda c: K.MaxPooling2D(tuple(c.get("pool_size", (2, 2))), c.get("strides"), c.get("border_mode", "valid"),
                                             c.get("dim_ordering", "th"), **_common(c)),
    "AveragePooling2D": lambda c: K.AveragePooling2D(tuple(c.get("pool_size", (2, 2))), c.get("strides"),
                                                     c.get("border_mode", "valid"), c.get("dim_ordering", "th"), **_common(c)),
    "MaxPooling3D": lambda c: K.MaxPooling3D(tuple(c.get("pool_size", (2, 2, 2))), c.get("strides"), **_common(c)),
    "AveragePooling3D": lambda c: K.AveragePooling3D(tuple(c.get("pool_size", (2, 2, 2))), c.get("strides"), **_common(c)),
    "GlobalMaxPooling1D": lambda c: K.GlobalMaxPooling1D(**_common(c)),
    "GlobalAveragePooling1D": lambda c: K.GlobalAveragePooling1D(**_common(c)),
    "GlobalMaxPooling2D": lambda c: K.GlobalMaxPooling2D(c.get("dim_ordering", "th"), **_common(c)),
    "GlobalAveragePooling2D": lambda c: K.GlobalAveragePooling2D(c.get("dim_ordering", "th"), **_common(c)),
    "ZeroPadding1D": lambda c: K.ZeroPadding1D(c.get("padding", 1), **_common(c)),
    "ZeroPadding2D": lambda c: K.ZeroPadding2D(tuple(c.get("padding", (1, 1))), c.get("dim_ordering", "th"), **_common(c)),
    "Cropping1D": lambda c: K.Cropping1D(tuple(c.get("cropping", (1, 1))), **_common(c)),
    "Cropping2D": lambda c: K.Cropping2D(c.get("cropping", ((0, 0), (0, 0))), c.get("dim_ordering", "th"), **_common(c)),
    "UpSampling1D": lambda c: K.UpSampling1D(c.get("length", 2), **_common(c)),
    "UpSampling2D": lambda c: K.UpSampling2D(tuple(c.get("size", (2, 2))), c.get("dim_ordering", "th"), **_common(c)),
    "UpSampling3D": lambda c: K.UpSampling3D(tuple(c.get("size", (2, 2, 2))), **_common(c)),
    "LeakyReLU": lambda c: K.LeakyReLU(c.get("alpha", 0.3), **_common(c)),
    "ELU": lambda c: K.ELU(c.get("alpha", 1.0), **_common(c)),
    "ThresholdedReLU": lambda c: K.ThresholdedReLU(c.get("theta", 1.0), **_common(c)),
    "GaussianNoise": lambda c: K.GaussianNoise(c["sigma"], **_common(c)),
    "GaussianDropout": lambda c: K.GaussianDropout(c["p"], **_common(c)),
    "SpatialDropout1D": lambda c: K.SpatialDropout1D(c.get("p", 0.5), **_common(c)),
    "SpatialDropout2D": lambda c: K.SpatialDropout2D(c.get("p", 0.5), c.get("dim_ordering", "th"), **_common(c)),
    "SimpleRNN": lambda c: K.SimpleRNN(c["output_dim"], c.get("activation", "tanh"), c.get("return_sequences", False),
                                       c.get("go_backwards", False), **_common(c)),
    "LSTM": lambda c: K.LSTM(c["output_dim"], c.get("activation", "tanh"), c.get("inner_activation", "hard_sigmoid"),
                             c.get("return_sequences", False), c.get("go_backwards", False), **_common(c)),
    "GRU": lambda c: K.GRU(c["output_dim"], c.get("activation", "tanh"), c.get("inner_activation", "hard_sigmoid"),
                           c.get("return_sequences", False), c.get("go_backwards", False), **_common(c)),
    # ---- the rest of pyspark/bigdl/keras/converter.py's create_* list
    "TimeDistributedDense": lambda c: K.TimeDistributed(K.Dense(c["output_dim"], init=c.get("init", "glorot_uniform"),
                                                                activation=_act(c), bias=c.get("bias", True)), **_common(c)),
    "AtrousConvolution1D": lambda c: K.AtrousConvolution1D(c["nb_filter"], c["filter_length"], activation=_act(c),
                                                           subsample_length=c.get("subsample_length", 1),
                                                           atrous_rate=c.get("atrous_rate", 1), bias=c.get("bias", True),
                                                           **_common(c)),
    "ConvLSTM2D": lambda c: K.ConvLSTM2D(c["nb_filter"], c["nb_row"], c.get("nb_col"), c.get("activation", "tanh"),
                                         c.get("inner_activation", "hard_sigmoid"), c.get("dim_ordering", "th"),
                                         tuple(c.get("subsample", (1, 1))), c.get("border_mode", "same"),
                                         return_sequences=c.get("return_sequences", False),
                                         go_backwards=c.get("go_backwards", False), **_common(c)),
    "GlobalMaxPooling3D": lambda c: K.GlobalMaxPooling3D(c.get("dim_ordering", "th"), **_common(c)),
    "GlobalAveragePooling3D": lambda c: K.GlobalAveragePooling3D(c.get("dim_ordering", "th"), **_common(c)),
    "ZeroPadding3D": lambda c: K.ZeroPadding3D(tuple(c.get("padding", (1, 1, 1))), c.get("dim_ordering", "th"), **_common(c)),
    "Cropping3D": lambda c: K.Cropping3D(c.get("cropping", ((1, 1), (1, 1), (1, 1))), c.get("dim_ordering", "th"), **_common(c)),
    "SpatialDropout3D": lambda c: K.SpatialDropout3D(c.get("p", 0.5), c.get("dim_ordering", "th"), **_common(c)),
    "SReLU": lambda c: K.SReLU(c.get("t_left_init", "zero"), c.get("a_left_init", "glorot_uniform"),
                               c.get("t_right_init", "glorot_uniform"), c.get("a_right_init", "one"),
                               c.get("shared_axes"), **_common(c)),
    "LocallyConnected1D": lambda c: K.LocallyConnected1D(c["nb_filter"], c["filter_length"], activation=_act(c),
                                                         border_mode=c.get("border_mode", "valid"),
                                                         subsample_length=c.get("subsample_length", 1),
                                                         bias=c.get("bias", True), **_common(c)),
    "LocallyConnected2D": lambda c: K.LocallyConnected2D(c["nb_filter"], c["nb_row"], c["nb_col"], activation=_act(c),
                                                         border_mode=c.get("border_mode", "valid"),
                                                         subsample=tuple(c.get("subsample", (1, 1))),
                                                         dim_ordering=c.get("dim_ordering", "th"), bias=c.get("bias", True),
                                                         **_common(c)),
    "PReLU": lambda c: _wrap_module(L.PReLU(0), c),
    "ParametricSoftplus": lambda c: _wrap_module(_ParametricSoftplus(c.get("alpha_init", 0.2), c.get("beta_init", 5.0)), c),
    "ActivityRegularization": lambda c: _wrap_module(L.ActivityRegularization(c.get("l1", 0.0), c.get("l2", 0.0)), c),
}


class _ParametricSoftplus(L.Module):
    """alpha * log(1 + exp(beta * x)) with learnable scalars (Keras 1.2.2 advanced activation; the reference only accepts
    it with the shared default parameters, pyspark/bigdl/keras/converter.py create_parametricsoftplus)."""

    def __init__(self, alpha=0.2, beta=5.0):
        super().__init__()
        self.register_parameter("alpha", torch.tensor([float(alpha)]))
        self.register_parameter("beta", torch.tensor([float(beta)]))

    def updateOutput(self, input):
        return self.alpha.to(input.dtype) * torch.nn.functional.softplus(self.beta.to(input.dtype) * input)


def _wrap_module(module, cfg):
    """A shape-preserving framework module as a Keras layer."""
    class _Wrapped(K.KerasLayer):
        def __init__(self, m, **kw):
            super().__init__(kw.pop("input_shape", None), **kw)
            self._m = m

        def doBuild(self, s):
            return self._m

        def computeOutputShape(self, s):
            return s

    return _Wrapped(module, **_common(cfg))


class DefinitionLoader:
    @staticmethod
    def _layer(spec):
        cls, cfg = spec["class_name"], spec["config"]
        if cls == "TimeDistributed":
            return K.TimeDistributed(DefinitionLoader._layer(cfg["layer"]), **_common(cfg))
        if cls == "Bidirectional":
            return K.Bidirectional(DefinitionLoader._layer(cfg["layer"]), cfg.get("merge_mode", "concat"), **_common(cfg))
        if cls == "InputLayer":
            return K.InputLayer(**_common(cfg))
        if cls in ("Sequential", "Model"):
            return DefinitionLoader.from_config(spec)
        if cls not in _BUILDERS:
            raise ValueError(f"Unsupported keras layer: {cls}")
        return _BUILDERS[cls](cfg)

    @staticmethod
    def from_config(model_spec):
        cls = model_spec["class_name"]
        if cls == "Sequential":
            m = K.Sequential()
            for spec in model_spec["config"]:
                m.add(DefinitionLoader._layer(spec))
            return m
        if cls == "Model":
            cfg = model_spec["config"]
            nodes = {}
            for spec in cfg["layers"]:
                name = spec["name"]
                if spec["class_name"] == "InputLayer":
                    nodes[name] = K.Input(shape=_shape(spec["config"]), name=name)
                    continue
                inbound = [nodes[i[0]] for i in spec["inbound_nodes"][0]]
                if spec["class_name"] == "Merge":
                    c = spec["config"]
                    layer = K.Merge(mode=c.get("mode", "sum"), concat_axis=c.get("concat_axis", -1), name=name)
                    nodes[name] = layer(inbound)
                else:
                    layer = DefinitionLoader._layer(spec)
                    nodes[name] = layer(inbound if len(inbound) > 1 else inbound[0])
            ins = [nodes[i[0]] for i in cfg["input_layers"]]
            outs = [nodes[o[0]] for o in cfg["output_layers"]]
            return K.Model(ins if len(ins) > 1 else ins[0], outs if len(outs) > 1 else outs[0])
        raise ValueError(f"Unsupported keras model type: {cls}")

    @staticmethod
    def from_json_str(s: str):
        return DefinitionLoader.from_config(json.loads(s))

    @staticmethod
    def from_kmodel(kmodel):
        """From a live Keras-1.2.2 model object (anything with ``to_json()``), as ``with_bigdl_backend`` does."""
        return DefinitionLoader.from_json_str(kmodel.to_json())

    @staticmethod
    def from_json_path(path: str):
        with open(path) as f:
            return DefinitionLoader.from_json_str(f.read())


def _find_layer(model, name):
    for m in model.all_modules():
        if m.hasName() and m.getName() == name:
            return m
    return None


class WeightLoader:
    """Keras layouts → framework layouts: Dense W [in,out] → [out,in]; Convolution2D th W [out,in,kh,kw] as is, tf W
    [kh,kw,in,out] → permuted; BatchNormalization [gamma, beta, mean, std-or-var]; Embedding [vocab, dim]."""

    @staticmethod
    def _apply(layer, arrays):
        params = [p for _, _, p in layer.leaf_parameters()]
        arrays = [np.asarray(a, dtype=np.float32) for a in arrays]
        cls = type(layer).__name__
        with torch.no_grad():
            if cls == "Dense":
                params[0].copy_(torch.from_numpy(arrays[0].T.copy()))
                if len(params) > 1 and len(arrays) > 1:
                    params[1].copy_(torch.from_numpy(arrays[1]))
            elif cls == "Convolution2D":
                w = arrays[0]
                if w.shape != tuple(params[0].shape):
                    w = w.transpose(3, 2, 0, 1)
                params[0].copy_(torch.from_numpy(np.ascontiguousarray(w)))
                if len(params) > 1 and len(arrays) > 1:
                    params[1].copy_(torch.from_numpy(arrays[1]))
            elif cls == "BatchNormalization":
                params[0].copy_(torch.from_numpy(arrays[0]))
                params[1].copy_(torch.from_numpy(arrays[1]))
                bn = [m for m in layer.all_modules() if "runningMean" in m._buffers][0]
                bn._buffers["runningMean"].copy_(torch.from_numpy(arrays[2]))
                bn._buffers["runningVar"].copy_(torch.from_numpy(arrays[3]))
            elif cls == "Embedding":
                params[0].copy_(torch.from_numpy(arrays[0]))
            elif cls in ("LSTM", "GRU", "SimpleRNN"):
                WeightLoader._recurrent(layer, cls, arrays)
            elif cls == "Bidirectional":
                half = len(arrays) // 2
                cells = [m for m in layer.all_modules() if "i2g" in m.__dict__ or "i2h" in m.__dict__]
                inner = type(layer.layer).__name__
                for cell, arrs in zip(cells, (arrays[:half], arrays[half:])):
                    WeightLoader._cell(cell, inner, arrs)
            elif cls == "TimeDistributed":
                WeightLoader._apply(layer.layer, arrays)
            elif cls in ("Convolution1D", "AtrousConvolution1D"):
                # keras [filter_length, 1, in, out] -> conv weight [out, in, 1|k, k|1] (the layer runs it as a 2-D conv)
                w = arrays[0]
                w = w.reshape(w.shape[0], w.shape[-2], w.shape[-1]).transpose(2, 1, 0)          # [out, in, k]
                params[0].copy_(torch.from_numpy(np.ascontiguousarray(w)).reshape(params[0].shape))
                if len(params) > 1 and len(arrays) > 1:
                    params[1].copy_(torch.from_numpy(arrays[1]))
            elif cls in ("Convolution3D",):
                w = arrays[0]
                if w.shape != tuple(params[0].shape):
                    w = w.transpose(4, 3, 0, 1, 2)                                               # tf: [d1,d2,d3,in,out]
                params[0].copy_(torch.from_numpy(np.ascontiguousarray(w)).reshape(params[0].shape))
                if len(params) > 1 and len(arrays) > 1:
                    params[1].copy_(torch.from_numpy(arrays[1]))
            elif cls in ("AtrousConvolution2D", "Deconvolution2D"):
                w = arrays[0]
                if w.size == params[0].numel() and w.shape != tuple(params[0].shape):
                    w = w.transpose(3, 2, 0, 1) if cls == "AtrousConvolution2D" else w
                    if cls == "Deconvolution2D" and w.shape != tuple(params[0].shape):
                        w = w.transpose(1, 0, 2, 3) if w.transpose(1, 0, 2, 3).shape == tuple(params[0].shape) else \
                            w.transpose(2, 3, 0, 1)
                params[0].copy_(torch.from_numpy(np.ascontiguousarray(w)).reshape(params[0].shape))
                if len(params) > 1 and len(arrays) > 1:
                    params[1].copy_(torch.from_numpy(arrays[1]))
            elif cls == "SeparableConvolution2D":
                # keras: depthwise [kh, kw, in, mult] (tf) or [in*mult... ] (th), pointwise [1,1,in*mult,out], bias
                dw, pw = arrays[0], arrays[1]
                if dw.ndim == 4 and dw.shape[:2] != tuple(params[0].shape[-2:]):
                    pass
                if dw.shape != tuple(params[0].shape):
                    dw = dw.transpose(2, 3, 0, 1).reshape(params[0].shape) if dw.size == params[0].numel() else dw
                if pw.shape != tuple(params[1].shape):
                    pw = pw.transpose(3, 2, 0, 1).reshape(params[1].shape) if pw.size == params[1].numel() else pw
                params[0].copy_(torch.from_numpy(np.ascontiguousarray(dw)).reshape(params[0].shape))
                params[1].copy_(torch.from_numpy(np.ascontiguousarray(pw)).reshape(params[1].shape))
                if len(params) > 2 and len(arrays) > 2:
                    params[2].copy_(torch.from_numpy(arrays[2]))
            elif cls == "Highway":
                # keras [W, W_carry, b, b_carry] (W [in, in]) -> Linear weights are [out, in]
                lin = [m for m in layer.all_modules() if type(m).__name__ == "Linear"]
                order = {0: (arrays[1], arrays[3] if len(arrays) > 3 else None),        # gate (carry) transform
                         1: (arrays[0], arrays[2] if len(arrays) > 2 else None)}        # candidate transform
                for i, m in enumerate(lin[:2]):
                    w, b = order[i]
                    m.weight.copy_(torch.from_numpy(w.T.copy()))
                    if b is not None and "bias" in m._parameters:
                        m.bias.copy_(torch.from_numpy(b))
            elif cls == "MaxoutDense":
                # keras W [nb_feature, in, out], b [nb_feature, out] -> one Linear(in, nb_feature * out)
                w, b = arrays[0], (arrays[1] if len(arrays) > 1 else None)
                k, i, o = w.shape
                params[0].copy_(torch.from_numpy(np.ascontiguousarray(w.transpose(0, 2, 1).reshape(k * o, i))))
                if b is not None and len(params) > 1:
                    params[1].copy_(torch.from_numpy(b.reshape(-1)))
            else:
                if len(arrays) != len(params):
                    raise ValueError(f"{cls} {layer.getName()}: expected {len(params)} weight arrays, got {len(arrays)}")
                for p, a in zip(params, arrays):
                    if a.size != p.numel():
                        raise ValueError(f"{cls} {layer.getName()}: weight size mismatch {a.shape} vs {tuple(p.shape)}")
                    p.copy_(torch.from_numpy(a.reshape(p.shape) if a.shape == tuple(p.shape) else a.T.reshape(p.shape)))
        for m in layer.all_modules():
            m._mark_weights_dirty()

    @staticmethod
    def _recurrent(layer, cls, arrays):
        cell = [m for m in layer.all_modules() if "i2g" in m.__dict__ or "i2h" in m.__dict__][0]
        WeightLoader._cell(cell, cls, arrays)

    @staticmethod
    def _cell(cell, cls, a):
        """Keras 1.2.2 recurrent weight lists -> the cells' fused gate matrices (reference:
        pyspark/bigdl/keras/converter.py convert_lstm / convert_gru / convert_simplernn).
        LSTM  [W_i,U_i,b_i, W_c,U_c,b_c, W_f,U_f,b_f, W_o,U_o,b_o]  -> gate chunks [input, candidate, forget, output]
        GRU   [W_z,U_z,b_z, W_r,U_r,b_r, W_h,U_h,b_h]               -> chunks [reset, update, candidate]
        SimpleRNN [W, U, b]."""
        t = lambda *xs: torch.from_numpy(np.ascontiguousarray(np.concatenate([x.T if x.ndim == 2 else x for x in xs])))  # noqa: E731
        if cls == "LSTM":
            cell.i2g.weight.copy_(t(a[0], a[3], a[6], a[9]))
            cell.h2g.weight.copy_(t(a[1], a[4], a[7], a[10]))
            cell.i2g.bias.copy_(t(a[2], a[5], a[8], a[11]))
        elif cls == "GRU":
            cell.i2g.weight.copy_(t(a[3], a[0], a[6]))
            cell.i2g.bias.copy_(t(a[5], a[2], a[8]))
            cell.h2g.weight.copy_(t(a[4], a[1]))
            cell.h2c.weight.copy_(t(a[7]))
        else:
            cell.i2h.weight.copy_(t(a[0]))
            cell.h2h.weight.copy_(t(a[1]))
            cell.i2h.bias.copy_(t(a[2]))
            if "bias" in cell.h2h._parameters:
                cell.h2h.bias.zero_()

    @staticmethod
    def load_weights_from_arrays(model, weights: dict, by_name: bool = False):
        for name, arrays in weights.items():
            layer = _find_layer(model, name)
            if layer is None:
                if by_name:
                    continue
                raise ValueError(f"layer {name} not found in the model")
            WeightLoader._apply(layer, arrays)
        return model

    @staticmethod
    def load_weights_from_kmodel(model, kmodel, by_name: bool = False):
        """Copy the weights of a live Keras model (``kmodel.layers[i].name`` / ``.get_weights()``)."""
        weights = {}
        for layer in kmodel.layers:
            arrays = layer.get_weights()
            if arrays:
                weights[layer.name] = [np.asarray(a) for a in arrays]
        return WeightLoader.load_weights_from_arrays(model, weights, by_name)

    @staticmethod
    def load_weights_from_hdf5(model, hdf5_path: str, by_name: bool = False):
        try:
            import h5py
        except ImportError as e:
            raise ImportError("reading Keras HDF5 weights needs h5py, which is not installed in this environment; "
                              "export the weights to .npz and use load_weights_from_arrays") from e
        weights = {}
        with h5py.File(hdf5_path, "r") as f:
            g = f["model_weights"] if "model_weights" in f else f
            for lname in [n.decode() if isinstance(n, bytes) else n for n in g.attrs["layer_names"]]:
                names = [n.decode() if isinstance(n, bytes) else n for n in g[lname].attrs["weight_names"]]
                if names:
                    weights[lname] = [np.asarray(g[lname][n]) for n in names]
        return WeightLoader.load_weights_from_arrays(model, weights, by_name)


def load_keras(json_path=None, hdf5_path=None, by_name=False):
    if json_path is None:
        raise ValueError("a model definition (json) is required: HDF5-only loading needs h5py")
    model = DefinitionLoader.from_json_path(json_path)
    if hdf5_path:
        if hdf5_path.endswith(".npz"):
            data = np.load(hdf5_path)
            grouped: dict = {}
            for k in data.files:
                lname, _, idx = k.rpartition("/")
                grouped.setdefault(lname, []).append((int(idx), data[k]))
            WeightLoader.load_weights_from_arrays(model, {k: [a for _, a in sorted(v)] for k, v in grouped.items()}, by_name)
        else:
            WeightLoader.load_weights_from_hdf5(model, hdf5_path, by_name)
    return model


# ------------------------------------------------------------------------------------------------ per-layer API
def _spec_of(klayer, kclayer=None):
    """{"class_name", "config"} of a Keras layer object (``get_config()``), of its JSON entry, or of both."""
    if isinstance(klayer, dict) and "class_name" in klayer:
        return klayer
    if kclayer is not None and isinstance(kclayer, dict) and "class_name" in kclayer:
        spec = {"class_name": kclayer["class_name"], "config": dict(kclayer["config"])}
    else:
        spec = {"class_name": type(klayer).__name__, "config": dict(klayer.get_config())}
    cfg = spec["config"]
    if "batch_input_shape" not in cfg and not isinstance(cfg, list):
        shp = getattr(klayer, "input_shape", None)
        if shp is not None and not (shp and isinstance(shp[0], (list, tuple))):
            cfg["batch_input_shape"] = list(shp)
    if not cfg.get("name") and getattr(klayer, "name", None):
        cfg["name"] = klayer.name
    return spec


class LayerConverter:
    """One Keras-1.2.2 layer → one Keras-style layer of this framework (P/keras/converter.py ``LayerConverter``):
    ``LayerConverter(klayer, kclayer, input_shape).create()``; the ``create_<type>()`` methods exist for every
    supported layer type and refuse a layer of another type."""

    def __init__(self, klayer, kclayer=None, input_shape=None):
        self.klayer, self.kclayer = klayer, kclayer
        self.spec = _spec_of(klayer, kclayer)
        self.config = self.spec["config"]
        if input_shape is not None and isinstance(self.config, dict):
            self.config["batch_input_shape"] = [None] + list(input_shape)[-len(input_shape):] \
                if input_shape and input_shape[0] is not None else list(input_shape)
        self.input_shape = _shape(self.config) if isinstance(self.config, dict) else None

    def create(self):
        fn = getattr(self, "create_" + self.spec["class_name"].lower(), None)
        if fn is None:
            unsupport_exp(self.spec["class_name"])
        return fn()

    def _make(self, expected: str):
        got = self.spec["class_name"]
        if got.lower() != expected.lower():
            raise ValueError(f"create_{expected.lower()} called for a {got} layer")
        return DefinitionLoader._layer(self.spec)

    # ---- recurrent pieces
    def _cell_of(self, expected):
        layer = self._make(expected)
        if self.input_shape is None:
            raise ValueError("the cell can only be generated once the input shape is known")
        layer.build(list(self.input_shape))
        return [m for m in layer.all_modules() if "i2g" in m.__dict__ or "i2h" in m.__dict__][0]

    def generate_simplernn_cell(self, *a, **k): return self._cell_of("SimpleRNN")
    def generate_lstm_cell(self, *a, **k): return self._cell_of("LSTM")
    def generate_gru_cell(self, *a, **k): return self._cell_of("GRU")
    def generate_convlstm2d_cell(self, *a, **k): return self._cell_of("ConvLSTM2D")

    # ---- small helpers the reference exposes
    def get_bdim_order(self, dim="2D"):
        order = self.config.get("dim_ordering", "th")
        if order not in ("tf", "th"):
            raise ValueError(f"Unsupport dim_ordering: {order}")
        return ("NHWC" if order == "tf" else "NCHW") if dim == "2D" else ("CHANNEL_LAST" if order == "tf" else "CHANNEL_FIRST")

    def combo_parameter_layer(self, blayer, config):
        """Attach the layer's name and, when it has one, its activation (the reference's post-processing of every
        created layer)."""
        if config.get("name"):
            blayer.setName(config["name"])
        act = _act(config)
        return LayerConverter.fuse(blayer, K.Activation(act)) if act else blayer

    @staticmethod
    def get_value_from_init(kinit_method, shape):
        if kinit_method == "zero":
            return np.zeros(shape, dtype=np.float32)
        if kinit_method == "one":
            return np.ones(shape, dtype=np.float32)
        raise ValueError(f"We don't support {kinit_method} for now")

    @staticmethod
    def fuse(src_blayer, activation, name=None):
        """layer → activation as one Sequential."""
        seq = K.Sequential()
        seq.add(src_blayer)
        seq.add(activation)
        if name:
            seq.setName(name)
        return seq

    def create_model(self): return self._make("Model")
    def create_sequential(self): return self._make("Sequential")
    def create_inputlayer(self): return self._make("InputLayer")
    def create_timedistributed(self): return self._make("TimeDistributed")
    def create_bidirectional(self): return self._make("Bidirectional")

    def create_merge(self):
        c = self.config
        return K.Merge(mode=c.get("mode", "sum"), concat_axis=c.get("concat_axis", -1), **_common(c))


def _add_creators():
    for cls_name in _BUILDERS:
        meth = "create_" + cls_name.lower()
        if not hasattr(LayerConverter, meth):
            setattr(LayerConverter, meth, (lambda n: lambda self: self._make(n))(cls_name))


_add_creators()


def unsupport_exp(name):
    raise Exception("We don't support %s for now" % name)


class WeightsConverter:
    """Keras weight lists → the parameter tensors of the converted layer, in that layer's parameter order
    (P/keras/converter.py ``WeightsConverter``). Each ``convert_<type>(klayer, weights)`` builds the converted layer
    from ``klayer``'s configuration, applies the same layout rules ``WeightLoader`` uses for whole models (transposes,
    gate re-ordering, th / tf kernel layouts) and returns the resulting arrays, so the per-layer and the whole-model
    paths cannot disagree. ``klayer`` needs ``get_config()`` and an input shape (``input_shape`` attribute or
    ``batch_input_shape`` in its config)."""

    @staticmethod
    def get_converter(class_name: str):
        fn = getattr(WeightsConverter, "convert_" + class_name.lower(), None)
        if fn is None:
            unsupport_exp(class_name)
        return fn

    @staticmethod
    def _converted(klayer, weights, expected=None):
        spec = _spec_of(klayer)
        if expected is not None and spec["class_name"].lower() != expected:
            raise ValueError(f"convert_{expected} called for a {spec['class_name']} layer")
        conv = LayerConverter(klayer)
        layer = conv.create()
        if conv.input_shape is None:
            raise ValueError(f"{spec['class_name']}: the layer's input shape is needed to lay out its weights")
        layer.build(list(conv.input_shape))
        WeightLoader._apply(layer, list(weights))
        out = [p.detach().cpu().numpy().copy() for _, _, p in layer.leaf_parameters()]
        if spec["class_name"] == "BatchNormalization":
            bn = [m for m in layer.all_modules() if "runningMean" in m._buffers][0]
            out += [bn._buffers["runningMean"].numpy().copy(), bn._buffers["runningVar"].numpy().copy()]
        return out

    @staticmethod
    def to_bigdl_weights(klayer, weights):
        return WeightsConverter.get_converter(_spec_of(klayer)["class_name"])(klayer, weights)

    @staticmethod
    def get_bigdl_weights_from_klayer(klayer):
        return WeightsConverter.to_bigdl_weights(klayer, klayer.get_weights())

    @staticmethod
    def get_weights_from_kmodel(kmodel):
        """Converted weights of every layer that has some, concatenated in layer order."""
        out = []
        for layer in kmodel.layers:
            w = layer.get_weights()
            if w:
                out.extend(WeightsConverter.get_bigdl_weights_from_klayer(layer))
        return out


def _add_weight_converters():
    names = ["timedistributed", "bidirectional", "dense", "timedistributeddense", "batchnormalization",
             "atrousconvolution2d", "atrousconvolution1d", "deconvolution2d", "convolution2d", "convolution1d",
             "convolution3d", "embedding", "simplernn", "lstm", "convlstm2d", "gru", "highway", "maxoutdense", "srelu",
             "separableconvolution2d", "locallyconnected1d", "locallyconnected2d"]
    for n in names:
        setattr(WeightsConverter, "convert_" + n,
                staticmethod((lambda e: lambda klayer, weights: WeightsConverter._converted(klayer, weights, e))(n)))


_add_weight_converters()


def _load_weights_from_json_hdf5(def_json, weights_hdf5, by_name=False):
    """Definition from the JSON file, weights from the HDF5 (or ``.npz``) file → model
    (``WeightLoader.load_weights_from_json_hdf5``)."""
    return load_keras(def_json, weights_hdf5, by_name)


WeightLoader.load_weights_from_json_hdf5 = staticmethod(_load_weights_from_json_hdf5)


def _from_hdf5_path(hdf5_path):
    """A model saved by ``keras_model.save(path)`` keeps its JSON configuration in the ``model_config`` attribute."""
    try:
        import h5py
    except ImportError as e:
        raise ImportError("reading a Keras HDF5 model needs h5py, which is not installed in this environment; "
                          "save the definition with model.to_json() and use from_json_path") from e
    with h5py.File(hdf5_path, "r") as f:
        cfg = f.attrs["model_config"]
        return DefinitionLoader.from_json_str(cfg.decode() if isinstance(cfg, bytes) else cfg)


DefinitionLoader.from_hdf5_path = staticmethod(_from_hdf5_path)
