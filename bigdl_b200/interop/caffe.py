"""Caffe model import / export.

Reference: S/utils/caffe/CaffeLoader.scala (``loadCaffe(defPath, modelPath)`` builds a Graph + optional criterion :283-
560; ``load(model, defPath, modelPath, matchAll)`` copies weights by layer name :57-280), V1LayerConverter.scala /
LayerConverter.scala (26 layer types, docs/docs/APIGuide/caffe_layer_list.md), CaffePersister.scala (export to prototxt
+ caffemodel). The protobuf schema is the runtime-built subset in ``caffe_pb``; prototxt parsing is protobuf's own
text_format. Caffe semantics kept: pooling output sizes round UP, InnerProduct flattens from ``axis``, BatchNorm blobs are
(mean, variance, scale_factor) and the usual BatchNorm+Scale pair folds into one affine batch-norm layer.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch
from google.protobuf import text_format

from ..serialization.module_io import packed_merge, packed_read
from . import caffe_pb as P


def _read_net(def_path: str | None, model_path: str | None):
    net_def, net_w = P.NetParameter(), P.NetParameter()
    if def_path:
        with open(def_path) as f:
            text_format.Parse(f.read(), net_def, allow_unknown_field=True)
    if model_path:
        with open(model_path, "rb") as f:
            net_w.ParseFromString(f.read())
    return net_def, net_w


def _blob(b) -> torch.Tensor:
    if len(b.data) >= 256 and sys.byteorder == "little":     # one raw record instead of a Python float per element
        data = packed_read(b, 5, np.float32)
    else:
        data = np.asarray(b.data if len(b.data) else b.double_data, dtype=np.float32)
    if b.HasField("shape") and len(b.shape.dim):
        shape = [int(d) for d in b.shape.dim]
    else:
        shape = [b.num, b.channels, b.height, b.width]
    if int(np.prod(shape)) != data.size:
        shape = [data.size]
    return torch.from_numpy(data.reshape(shape).copy())


def _layers(net):
    """Unified view over V2 ``layer`` and V1 ``layers``: (name, type string, bottoms, tops, message)."""
    out = []
    for l in net.layer:
        out.append((l.name, l.type, list(l.bottom), list(l.top), l))
    for l in net.layers:
        t = P.V1_TYPE_NAMES.get(l.type, "NONE")
        canon = {"CONVOLUTION": "Convolution", "INNER_PRODUCT": "InnerProduct", "RELU": "ReLU", "POOLING": "Pooling",
                 "LRN": "LRN", "DROPOUT": "Dropout", "SOFTMAX": "Softmax", "SOFTMAX_LOSS": "SoftmaxWithLoss",
                 "CONCAT": "Concat", "ELTWISE": "Eltwise", "TANH": "TanH", "SIGMOID": "Sigmoid", "FLATTEN": "Flatten",
                 "SPLIT": "Split", "POWER": "Power", "ABSVAL": "AbsVal", "BNLL": "BNLL", "DATA": "Data",
                 "SLICE": "Slice", "THRESHOLD": "Threshold", "EXP": "Exp", "DECONVOLUTION": "Deconvolution",
                 "ACCURACY": "Accuracy", "SILENCE": "Silence", "DUMMY_DATA": "DummyData", "MEMORY_DATA": "MemoryData",
                 "IMAGE_DATA": "ImageData", "HDF5_DATA": "HDF5Data", "EUCLIDEAN_LOSS": "EuclideanLoss"}.get(t, t)
        out.append((l.name, canon, list(l.bottom), list(l.top), l))
    return out


def _excluded_for_test(l) -> bool:
    if hasattr(l, "phase") and l.HasField("phase") and l.phase == 0:       # TRAIN-only layer
        return True
    inc = [r.phase for r in l.include if r.HasField("phase")]
    return bool(inc) and 1 not in inc


_DATA_TYPES = {"Data", "Input", "DummyData", "MemoryData", "ImageData", "HDF5Data", "AnnotatedData", "WindowData"}
_SKIP_TYPES = {"Accuracy", "Silence"}
_LOSS_TYPES = {"SoftmaxWithLoss", "EuclideanLoss", "SigmoidCrossEntropyLoss", "HingeLoss"}


def _conv_geometry(cp):
    def pick(rep, h, w, has_h, has_w, default):
        vals = list(rep)
        if has_h or has_w:
            return int(h), int(w)
        if not vals:
            return default, default
        return (int(vals[0]), int(vals[0])) if len(vals) == 1 else (int(vals[0]), int(vals[1]))
    kh, kw = pick(cp.kernel_size, cp.kernel_h, cp.kernel_w, cp.HasField("kernel_h"), cp.HasField("kernel_w"), 1)
    sh, sw = pick(cp.stride, cp.stride_h, cp.stride_w, cp.HasField("stride_h"), cp.HasField("stride_w"), 1)
    ph, pw = pick(cp.pad, cp.pad_h, cp.pad_w, cp.HasField("pad_h") and cp.pad_h > 0, cp.HasField("pad_w") and cp.pad_w > 0, 0)
    d = list(cp.dilation)
    dh, dw = (1, 1) if not d else ((int(d[0]), int(d[0])) if len(d) == 1 else (int(d[0]), int(d[1])))
    return kh, kw, sh, sw, ph, pw, dh, dw


class CaffeConversionException(ValueError):
    """Raised for a layer that cannot be converted (S/utils/caffe/CaffeConversionException.scala)."""

    def __init__(self, conversionMsg: str, error: BaseException | None = None):
        super().__init__(conversionMsg)
        self.conversionMsg, self.error = conversionMsg, error

    def __str__(self):
        return f"Caffe conversion error : {self.conversionMsg}" + (f"\n{self.error!r}" if self.error else "")


class Customizable:
    """Base of user converters handed to ``loadCaffe(customizedConverters={"TYPE": conv})`` (CaffeLoader.scala:37-47):
    implement ``convertor(layer)`` (optionally ``convertor(layer, blobs)``) returning the module for that layer; the
    loader registers the parsed net as ``contexts["netparam"]`` and the name → layer map as
    ``contexts["name2LayerV2"]`` before converting. A plain ``fn(layer, blobs)`` works too."""

    contexts: dict | None = None

    def convertor(self, layer, blobs=None):  # pragma: no cover - interface
        raise NotImplementedError

    def registerContext(self, name: str, context):
        if self.contexts is None:
            self.contexts = {}
        self.contexts[name] = context

    def __call__(self, layer, blobs=None):
        import inspect

        n = len(inspect.signature(self.convertor).parameters)
        return self.convertor(layer, blobs) if n >= 2 else self.convertor(layer)


class _Converter:
    """One caffe layer → one module (or a small Sequential). ``blobs`` are the trained tensors, possibly empty."""

    def __init__(self, customized=None):
        self.custom = dict(customized or {})

    def registerCutomizedConverter(self, layerType: str, converter):
        """(sic — the reference's spelling, Converter.scala:60) add a converter for a layer type."""
        self.custom[layerType] = converter
        return self

    registerCustomizedConverter = registerCutomizedConverter

    def convert(self, typ, l, blobs, in_channels=None):
        from ..nn import layer as L

        if typ in self.custom:
            return self.custom[typ](l, blobs)
        fn = getattr(self, "_" + typ.lower(), None)
        if fn is None:
            raise CaffeConversionException(f"unsupported caffe layer type {typ} (layer {l.name})")
        m = fn(L, l, blobs)
        if m is not None and l.name:
            m.setName(l.name)
        return m

    @staticmethod
    def _copy(m, names, blobs):
        with torch.no_grad():
            for n, b in zip(names, blobs):
                if n in m._parameters:
                    m._parameters[n].copy_(b.reshape(m._parameters[n].shape))
        m._mark_weights_dirty()
        return m

    def _convolution(self, L, l, blobs):
        cp = l.convolution_param
        kh, kw, sh, sw, ph, pw, dh, dw = _conv_geometry(cp)
        n_out, g = int(cp.num_output), int(cp.group)
        if blobs:
            n_in = blobs[0].shape[1] * g if blobs[0].dim() == 4 else blobs[0].numel() // (n_out * kh * kw) * g
        else:
            raise ValueError(f"convolution {l.name}: input channel count needs the caffemodel weights")
        if dh != 1 or dw != 1:
            if g != 1:
                raise ValueError("dilated grouped convolution is not supported")
            m = L.SpatialDilatedConvolution(n_in, n_out, kw, kh, sw, sh, pw, ph, dw, dh)
        else:
            m = L.SpatialConvolution(n_in, n_out, kw, kh, sw, sh, pw, ph, g, with_bias=bool(cp.bias_term))
        return self._copy(m, ["weight", "bias"], blobs)

    def _deconvolution(self, L, l, blobs):
        cp = l.convolution_param
        kh, kw, sh, sw, ph, pw, _, _ = _conv_geometry(cp)
        n_out, g = int(cp.num_output), int(cp.group)
        n_in = blobs[0].shape[0]
        m = L.SpatialFullConvolution(n_in, n_out, kw, kh, sw, sh, pw, ph, 0, 0, g, not bool(cp.bias_term))
        return self._copy(m, ["weight", "bias"], blobs)

    def _innerproduct(self, L, l, blobs):
        ip = l.inner_product_param
        n_out = int(ip.num_output)
        if not blobs:
            raise ValueError(f"inner product {l.name}: input size needs the caffemodel weights")
        w = blobs[0].reshape(n_out, -1) if not ip.transpose else blobs[0].reshape(-1, n_out).t()
        lin = L.Linear(w.shape[1], n_out, with_bias=bool(ip.bias_term))
        self._copy(lin, ["weight", "bias"], [w] + blobs[1:])
        lin.setName(l.name)
        seq = L.Sequential().add(L.InferReshape([-1, w.shape[1]], False)).add(lin)
        return seq

    def _relu(self, L, l, blobs):
        ns = float(l.relu_param.negative_slope)
        return L.ReLU() if ns == 0 else L.LeakyReLU(ns)

    def _tanh(self, L, l, blobs): return L.Tanh()
    def _sigmoid(self, L, l, blobs): return L.Sigmoid()
    def _absval(self, L, l, blobs): return L.Abs()
    def _bnll(self, L, l, blobs): return L.SoftPlus()
    def _split(self, L, l, blobs): return L.Identity()
    def _flatten(self, L, l, blobs): return L.InferReshape([0, -1], False)

    def _elu(self, L, l, blobs): return L.ELU(float(l.elu_param.alpha))
    def _threshold(self, L, l, blobs):
        # Caffe's Threshold emits 1 where x > threshold, else 0 (Converter.scala:205-212 maps it to BinaryThreshold; an
        # unset threshold is 1e-6 there)
        tp = l.threshold_param
        return L.BinaryThreshold(float(tp.threshold) if tp.HasField("threshold") else 1e-6)

    def _recurrent(self, L, l, blobs):
        return L.Recurrent()             # an empty time-loop container, as Converter.scala:200-203 (cells are added by the user)

    _rnn = _recurrent

    def _power(self, L, l, blobs):
        p = l.power_param
        return L.Power(float(p.power), float(p.scale), float(p.shift))

    def _exp(self, L, l, blobs):
        p = l.exp_param
        seq = L.Sequential()
        scale = float(p.scale) * (1.0 if p.base <= 0 else float(np.log(p.base)))
        shift = float(p.shift) * (1.0 if p.base <= 0 else float(np.log(p.base)))
        if scale != 1.0:
            seq.add(L.MulConstant(scale))
        if shift != 0.0:
            seq.add(L.AddConstant(shift))
        return seq.add(L.Exp())

    def _log(self, L, l, blobs):
        p = l.log_param
        seq = L.Sequential()
        if p.scale != 1.0:
            seq.add(L.MulConstant(float(p.scale)))
        if p.shift != 0.0:
            seq.add(L.AddConstant(float(p.shift)))
        seq.add(L.Log())
        if p.base > 0:
            seq.add(L.MulConstant(1.0 / float(np.log(p.base))))
        return seq

    def _prelu(self, L, l, blobs):
        n = 0 if l.prelu_param.channel_shared or not blobs else blobs[0].numel()
        m = L.PReLU(n)
        return self._copy(m, ["weight"], blobs)

    def _lrn(self, L, l, blobs):
        p = l.lrn_param
        if p.norm_region == 1:
            return L.SpatialWithinChannelLRN(int(p.local_size), float(p.alpha), float(p.beta))
        return L.SpatialCrossMapLRN(int(p.local_size), float(p.alpha), float(p.beta), float(p.k))

    def _pooling(self, L, l, blobs):
        p = l.pooling_param
        kh = int(p.kernel_h) if p.HasField("kernel_h") else int(p.kernel_size)
        kw = int(p.kernel_w) if p.HasField("kernel_w") else int(p.kernel_size)
        sh = int(p.stride_h) if p.HasField("stride_h") else int(p.stride)
        sw = int(p.stride_w) if p.HasField("stride_w") else int(p.stride)
        ph = int(p.pad_h) if p.pad_h else int(p.pad)
        pw = int(p.pad_w) if p.pad_w else int(p.pad)
        if p.pool == 0:
            if p.global_pooling:
                return L.Sequential().add(L.Max(4, 4)).add(L.Max(3, 4)).add(L.Unsqueeze(3)).add(L.Unsqueeze(4))
            return L.SpatialMaxPooling(kw, kh, sw, sh, pw, ph, to_ceil=True)
        if p.pool == 1:
            return L.SpatialAveragePooling(kw, kh, sw, sh, pw, ph, global_pooling=bool(p.global_pooling), ceil_mode=True)
        raise ValueError("stochastic pooling is not supported")

    def _dropout(self, L, l, blobs): return L.Dropout(float(l.dropout_param.dropout_ratio))
    def _softmax(self, L, l, blobs): return L.SoftMax()
    def _softmaxwithloss(self, L, l, blobs): return L.SoftMax()      # the criterion is returned separately

    def _concat(self, L, l, blobs):
        p = l.concat_param
        axis = int(p.axis) if p.HasField("axis") or not p.HasField("concat_dim") else int(p.concat_dim)
        return L.JoinTable(axis + 1)

    def _eltwise(self, L, l, blobs):
        p = l.eltwise_param
        if p.operation == 0:
            return L.CMulTable()
        if p.operation == 2:
            return L.CMaxTable()
        coeff = list(p.coeff)
        if coeff and any(c != 1.0 for c in coeff):
            if coeff == [1.0, -1.0]:
                return L.CSubTable()
            pt = L.ParallelTable()
            for c in coeff:
                pt.add(L.MulConstant(float(c)))
            return L.Sequential().add(pt).add(L.CAddTable())
        return L.CAddTable()

    def _reshape(self, L, l, blobs):
        dims = [int(d) for d in l.reshape_param.shape.dim]
        return L.InferReshape(dims, False)

    def _tile(self, L, l, blobs):
        return L.Tile(int(l.tile_param.axis) + 1, int(l.tile_param.tiles))

    def _batchnorm(self, L, l, blobs):
        if not blobs:
            raise ValueError(f"batch norm {l.name}: channel count needs the caffemodel weights")
        n = blobs[0].numel()
        m = L.SpatialBatchNormalization(n, float(l.batch_norm_param.eps), 1.0 - float(l.batch_norm_param.moving_average_fraction),
                                        affine=False)
        sf = float(blobs[2].reshape(-1)[0]) if len(blobs) > 2 else 1.0
        sf = 1.0 / sf if sf != 0 else 0.0
        m._buffers["runningMean"].copy_(blobs[0].reshape(-1) * sf)
        m._buffers["runningVar"].copy_(blobs[1].reshape(-1) * sf)
        return m

    def _scale(self, L, l, blobs):
        if not blobs:
            raise ValueError(f"scale {l.name}: needs the caffemodel weights")
        n = blobs[0].numel()
        m = L.Scale([1, n, 1, 1])
        with torch.no_grad():
            ps = [p for _, _, p in m.leaf_parameters()]
            ps[0].copy_(blobs[0].reshape(ps[0].shape))
            if len(blobs) > 1 and len(ps) > 1:
                ps[1].copy_(blobs[1].reshape(ps[1].shape))
        m._mark_weights_dirty()
        return m

    def _bias(self, L, l, blobs):
        n = blobs[0].numel()
        m = L.CAdd([1, n, 1, 1])
        return self._copy(m, ["bias"], blobs)

    def _slice(self, L, l, blobs):
        return None            # handled by the graph builder (needs the number of tops)


Converter = _Converter


class LayerConverter(_Converter):
    """Converter of the current ``LayerParameter`` message (LayerConverter.scala). Old-style ``V1LayerParameter``
    nets are normalised to the same (name, TYPE, bottoms, tops, message) tuples by ``_layers`` before conversion, so
    one converter serves both formats."""


class V1LayerConverter(_Converter):
    """Converter of the legacy ``V1LayerParameter`` message (V1LayerConverter.scala); see ``LayerConverter``."""


class CaffeLoader:
    @staticmethod
    def loadCaffe(defPath: str, modelPath: str, customizedConverters=None, outputNames=None):
        """→ (Graph, criterion or None)."""
        from ..nn import layer as L
        from ..nn.criterion import ClassNLLCriterion, MSECriterion
        from ..nn.module import Graph, Node

        net_def, net_w = _read_net(defPath, modelPath)
        weights = {name: [_blob(b) for b in msg.blobs] for name, _, _, _, msg in _layers(net_w)}
        conv = LayerConverter(customizedConverters)
        for c in (customizedConverters or {}).values():
            if isinstance(c, Customizable):
                c.registerContext("netparam", net_def)
                c.registerContext("name2LayerV2", {name: msg for name, _, _, _, msg in _layers(net_def)})
        tops: dict[str, Node] = {}
        inputs: list[Node] = []
        criterion = None

        def new_input(name):
            n = L.Input(name)
            tops[name] = n
            inputs.append(n)

        for name in net_def.input:
            new_input(name)
        consumed: set[str] = set()
        for name, typ, bottoms, top_names, msg in _layers(net_def):
            if _excluded_for_test(msg) and typ not in _DATA_TYPES:
                continue
            if typ in _DATA_TYPES:
                if top_names and top_names[0] not in tops:
                    new_input(top_names[0])
                continue
            if typ in _SKIP_TYPES:
                continue
            blobs = weights.get(name, [])
            if typ in _LOSS_TYPES:
                criterion = ClassNLLCriterion(logProbAsInput=False) if typ == "SoftmaxWithLoss" else MSECriterion()
                bottoms = bottoms[:1]
                if typ != "SoftmaxWithLoss":
                    continue
            prev = [tops[b] for b in bottoms if b in tops]
            consumed.update(b for b in bottoms if b in tops)
            if typ == "Slice":
                p = msg.slice_param
                axis = (int(p.axis) if p.HasField("axis") or not p.HasField("slice_dim") else int(p.slice_dim)) + 1
                pts = [0] + [int(v) for v in p.slice_point]
                for i, t in enumerate(top_names):
                    if i < len(pts) - 1 or (i == len(pts) - 1 and p.slice_point):
                        length = (pts[i + 1] - pts[i]) if i + 1 < len(pts) else -1
                        if length == -1:
                            mod = L.Narrow(axis, pts[i] + 1, -1)
                        else:
                            mod = L.Narrow(axis, pts[i] + 1, length)
                    else:
                        raise ValueError("Slice without slice_point needs static shapes; give slice_point")
                    mod.setName(f"{name}_{i}")
                    tops[t] = Node(mod, prev)
                continue
            module = conv.convert(typ, msg, blobs)
            node = Node(module, prev)
            for t in top_names:
                tops[t] = node
        if outputNames:
            outs = [tops[n] for n in outputNames]
        else:
            seen, outs = set(), []
            for t, n in tops.items():
                if t not in consumed and id(n) not in seen and n not in inputs:
                    seen.add(id(n))
                    outs.append(n)
        g = Graph(inputs, outs)
        g.setName(net_def.name or "caffe_model")
        return g, criterion

    @staticmethod
    def load(model, defPath: str | None, modelPath: str, matchAll: bool = True):
        """Copy trained blobs into ``model`` by layer name (CaffeLoader.load :57-280). With ``matchAll`` every
        parameterised layer of the model must be found in the caffemodel."""
        _, net_w = _read_net(None, modelPath)
        weights = {name: [_blob(b) for b in msg.blobs] for name, _, _, _, msg in _layers(net_w) if len(msg.blobs)}
        for m in model.all_modules():
            own = [n for n in ("weight", "bias") if n in m._parameters]
            if not own:
                continue
            blobs = weights.get(m.getName())
            if blobs is None:
                if matchAll:
                    raise ValueError(f"module {m.getName()} cannot map a layer in caffe model")
                continue
            with torch.no_grad():
                for n, b in zip(own, blobs):
                    p = m._parameters[n]
                    if b.numel() != p.numel():
                        raise ValueError(f"{n} element number is not equal between caffe layer and bigdl module "
                                         f"{m.getName()}, data shape in caffe is {list(b.shape)}, while data shape in "
                                         f"bigdl is {list(p.shape)}")
                    p.copy_(b.reshape(p.shape))
            m._mark_weights_dirty()
        return model


# ------------------------------------------------------------------------------------------------ export
class CaffePersister:
    """Module tree → prototxt (definition) + caffemodel (definition + blobs) (S/utils/caffe/CaffePersister.scala)."""

    @staticmethod
    def persist(prototxtPath: str, modelPath: str, module, useV2: bool = True, overwrite: bool = False):
        from ..nn import layer as L
        from ..nn.module import Graph, Sequential

        for p in (prototxtPath, modelPath):
            if os.path.exists(p) and not overwrite:
                raise FileExistsError(p)
        net = P.NetParameter()
        net.name = module.getName()
        net.input.append("data")
        counter = [0]

        def blob(layer, t):
            b = layer.blobs.add()
            b.shape.dim.extend(list(t.shape))
            flat = t.detach().float().reshape(-1)
            if flat.numel() >= 256 and sys.byteorder == "little":
                packed_merge(b, 5, flat.cpu().numpy())
            else:
                b.data.extend(flat.tolist())

        def emit(m, bottoms):
            """Returns the top name produced by ``m``."""
            if isinstance(m, Sequential):
                cur = bottoms
                for c in m.modules:
                    cur = [emit(c, cur)]
                return cur[0]
            if isinstance(m, (L.Identity,)) or (type(m).__name__ == "Input"):
                return bottoms[0]
            if isinstance(m, L.ConcatTable):
                raise ValueError("ConcatTable must be followed by a table layer; export a Graph instead")
            layer = net.layer.add()
            counter[0] += 1
            layer.name = m.getName() if m.hasName() else f"{type(m).__name__}{counter[0]}"
            layer.bottom.extend(bottoms)
            layer.top.append(layer.name)
            if isinstance(m, L.SpatialConvolution):
                layer.type = "Convolution"
                cp = layer.convolution_param
                cp.num_output, cp.group, cp.bias_term = m.n_output_plane, m.n_group, m.with_bias
                cp.kernel_h, cp.kernel_w, cp.stride_h, cp.stride_w = m.kernel_h, m.kernel_w, m.stride_h, m.stride_w
                cp.pad_h, cp.pad_w = max(m.pad_h, 0), max(m.pad_w, 0)
                if m.dilation_h != 1 or m.dilation_w != 1:
                    cp.dilation.extend([m.dilation_h, m.dilation_w])
                blob(layer, m.weight.contiguous())
                if m.with_bias:
                    blob(layer, m.bias)
            elif isinstance(m, L.Linear):
                layer.type = "InnerProduct"
                layer.inner_product_param.num_output, layer.inner_product_param.bias_term = m.output_size, m.with_bias
                blob(layer, m.weight)
                if m.with_bias:
                    blob(layer, m.bias)
            elif type(m) is L.ReLU:
                layer.type = "ReLU"
            elif isinstance(m, L.LeakyReLU):
                layer.type = "ReLU"
                layer.relu_param.negative_slope = m.negval
            elif isinstance(m, L.Tanh):
                layer.type = "TanH"
            elif isinstance(m, L.Sigmoid):
                layer.type = "Sigmoid"
            elif isinstance(m, L.SpatialMaxPooling) or isinstance(m, L.SpatialAveragePooling):
                layer.type = "Pooling"
                pp = layer.pooling_param
                pp.pool = 0 if isinstance(m, L.SpatialMaxPooling) else 1
                pp.kernel_h, pp.kernel_w, pp.stride_h, pp.stride_w = m.kh, m.kw, m.dh, m.dw
                pp.pad_h, pp.pad_w = max(m.pad_h, 0), max(m.pad_w, 0)
                if getattr(m, "global_pooling", False):
                    pp.global_pooling = True
            elif isinstance(m, L.SpatialCrossMapLRN):
                layer.type = "LRN"
                lp = layer.lrn_param
                lp.local_size, lp.alpha, lp.beta, lp.k = m.size, m.alpha, m.beta, m.k
            elif isinstance(m, L.Dropout):
                layer.type = "Dropout"
                layer.dropout_param.dropout_ratio = m.p
            elif isinstance(m, (L.SoftMax, L.LogSoftMax)):
                layer.type = "Softmax"
            elif isinstance(m, L.BatchNormalization):
                layer.type = "BatchNorm"
                layer.batch_norm_param.eps = m.eps
                layer.batch_norm_param.use_global_stats = True
                blob(layer, m._buffers["runningMean"])
                blob(layer, m._buffers["runningVar"])
                blob(layer, torch.ones(1))
                if m.affine:
                    sc = net.layer.add()
                    sc.name, sc.type = layer.name + "_scale", "Scale"
                    sc.bottom.append(layer.name)
                    sc.top.append(sc.name)
                    sc.scale_param.bias_term = True
                    blob(sc, m.weight)
                    blob(sc, m.bias)
                    return sc.name
            elif isinstance(m, L.CAddTable):
                layer.type = "Eltwise"
                layer.eltwise_param.operation = 1
            elif isinstance(m, L.CMulTable):
                layer.type = "Eltwise"
                layer.eltwise_param.operation = 0
            elif isinstance(m, L.CMaxTable):
                layer.type = "Eltwise"
                layer.eltwise_param.operation = 2
            elif isinstance(m, L.JoinTable):
                layer.type = "Concat"
                layer.concat_param.axis = m.dimension - 1
            elif isinstance(m, (L.View, L.Reshape, L.InferReshape)):
                layer.type = "Flatten"
            # the rest of Converter.scala:297-329's toCaffe table
            elif isinstance(m, L.SpatialFullConvolution):
                layer.type = "Deconvolution"
                cp = layer.convolution_param
                has_bias = getattr(m, "bias", None) is not None
                cp.num_output, cp.group, cp.bias_term = m.nout, m.n_group, has_bias
                cp.kernel_h, cp.kernel_w, cp.stride_h, cp.stride_w = m.kh, m.kw, m.dh, m.dw
                cp.pad_h, cp.pad_w = m.pad_h, m.pad_w
                blob(layer, m.weight.reshape(m.nin, m.nout // m.n_group, m.kh, m.kw))
                if has_bias:
                    blob(layer, m.bias)
            elif isinstance(m, L.SpatialWithinChannelLRN):
                layer.type = "LRN"
                lp = layer.lrn_param
                lp.local_size, lp.alpha, lp.beta, lp.norm_region = m.size, m.alpha, m.beta, 1
            elif isinstance(m, (L.BinaryThreshold, L.Threshold)):
                # Caffe's Threshold is a 0/1 step (= BinaryThreshold); the reference also writes its value-replacing
                # Threshold under this type (Converter.scala toCaffeThreshold), which reads back as BinaryThreshold
                layer.type = "Threshold"
                layer.threshold_param.threshold = m.th
            elif isinstance(m, L.ELU):
                layer.type = "ELU"
                layer.elu_param.alpha = m.alpha
            elif isinstance(m, L.PReLU):
                layer.type = "PReLU"
                layer.prelu_param.channel_shared = m.n_output_plane == 0
                blob(layer, m.weight)
            elif isinstance(m, L.Power):
                layer.type = "Power"
                layer.power_param.power, layer.power_param.scale, layer.power_param.shift = m.power, m.scale, m.shift
            elif isinstance(m, L.Abs):
                layer.type = "AbsVal"
            elif isinstance(m, L.Exp):
                layer.type = "Exp"
            elif isinstance(m, L.Log):
                layer.type = "Log"
            elif isinstance(m, L.Scale):
                layer.type = "Scale"
                layer.scale_param.bias_term = True
                ps = [p for _, _, p in m.leaf_parameters()]
                blob(layer, ps[0].reshape(-1))
                blob(layer, ps[1].reshape(-1))
            elif isinstance(m, (L.Add, L.CAdd)):
                layer.type = "Bias"
                blob(layer, m.bias.reshape(-1))
            elif isinstance(m, L.Replicate):
                layer.type = "Tile"
                layer.tile_param.axis, layer.tile_param.tiles = m.dim - 1, m.n_features
            elif isinstance(m, L.CSubTable):
                layer.type = "Eltwise"
                layer.eltwise_param.operation = 1
                layer.eltwise_param.coeff.extend([1.0, -1.0])
            else:
                raise ValueError(f"saveCaffe: unsupported module {type(m).__name__}")
            return layer.name

        if isinstance(module, Graph):
            names: dict[int, str] = {}
            for n in module.exec_order:
                if not n.prev:
                    names[id(n)] = "data"
                    continue
                names[id(n)] = emit(n.element, [names[id(p)] for p in n.prev])
        else:
            emit(module, ["data"])
        with open(modelPath, "wb") as f:
            f.write(net.SerializeToString())
        proto = P.NetParameter()
        proto.CopyFrom(net)
        for l in proto.layer:
            del l.blobs[:]
        with open(prototxtPath, "w") as f:
            f.write(text_format.MessageToString(proto))
