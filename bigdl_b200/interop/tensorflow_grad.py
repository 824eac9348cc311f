"""TensorFlow op loaders for the backward half of a training graph and the 3-D / morphological families.

The reference keeps one loader per op under S/utils/tf/loaders/ (161 files); the forward ops live in
``interop/tensorflow.py``. This module adds the ``*Grad`` / ``*Backprop*`` ops a *training* GraphDef contains (the graphs
``tf.train.Optimizer.minimize`` emits: S/utils/tf/loaders/{ReluGrad,MaxPoolGrad,Conv2DBackpropFilter,BiasAddGrad,
FusedBatchNormGrad,LRNGrad,ResizeBilinearGrad,...}.scala), ``Conv3D`` and ``Dilation2D`` with their backprops,
``SoftmaxCrossEntropyWithLogits`` and a few stragglers. Gradients of composite ops are evaluated as vector-Jacobian
products of the forward op under torch autograd, which is exactly what the TF kernels compute.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from ..utils.table import Table
from .tensorflow import _Fn, _attr, _hw, _ints, _nchw, _same_pad, op


def _vjp(fwd, primal: torch.Tensor, cotangent: torch.Tensor) -> torch.Tensor:
    with torch.enable_grad():
        x = primal.detach().float().requires_grad_(True)
        y = fwd(x)
        (g,) = torch.autograd.grad(y, x, cotangent.to(y.dtype).reshape(y.shape))
    return g.to(primal.dtype)


def _to_nchw(x, nchw):
    return x if nchw else x.permute(0, 3, 1, 2)


def _from_nchw(x, nchw):
    return x if nchw else x.permute(0, 2, 3, 1)


# ---------------------------------------------------------------------------------------------- element-wise gradients
@op("ReluGrad")
def _relu_grad(n): return _Fn(lambda g, f: g * (f > 0).to(g.dtype), 2)


@op("Relu6Grad")
def _relu6_grad(n): return _Fn(lambda g, f: g * ((f > 0) & (f < 6)).to(g.dtype), 2)


@op("EluGrad")
def _elu_grad(n): return _Fn(lambda g, y: g * torch.where(y > 0, torch.ones_like(y), y + 1), 2)


@op("SigmoidGrad")
def _sigmoid_grad(n): return _Fn(lambda y, dy: dy * y * (1 - y), 2)


@op("TanhGrad")
def _tanh_grad(n): return _Fn(lambda y, dy: dy * (1 - y * y), 2)


@op("SoftplusGrad")
def _softplus_grad(n): return _Fn(lambda g, f: g * torch.sigmoid(f), 2)


@op("SoftsignGrad")
def _softsign_grad(n): return _Fn(lambda g, f: g / (1 + f.abs()) ** 2, 2)


@op("SqrtGrad")
def _sqrt_grad(n): return _Fn(lambda y, dy: dy * 0.5 / y, 2)


@op("RsqrtGrad")
def _rsqrt_grad(n): return _Fn(lambda y, dy: -0.5 * dy * y ** 3, 2)


@op("ReciprocalGrad", "InvGrad")
def _recip_grad(n): return _Fn(lambda y, dy: -dy * y * y, 2)


@op("ApproximateEqual")
def _approx_equal(n):
    tol = float(_attr(n, "tolerance", 1e-5))
    return _Fn(lambda a, b: (a - b).abs() < tol, 2)


@op("TruncateMod")
def _truncate_mod(n): return _Fn(lambda a, b: torch.fmod(a, b), 2)


@op("BiasAddV1")
def _bias_add_v1(n): return _Fn(lambda x, b: x + b, 2)


@op("BiasAddGrad")
def _bias_add_grad(n):
    nchw = _nchw(n)

    def fn(g):
        dims = [d for d in range(g.dim()) if d != (1 if nchw else g.dim() - 1)]
        return g.sum(dims)
    return _Fn(fn, 1)


@op("BroadcastGradientArgs")
def _broadcast_gradient_args(n):
    def fn(s0, s1):
        a, b = _ints(s0), _ints(s1)
        r = max(len(a), len(b))
        a2, b2 = [1] * (r - len(a)) + a, [1] * (r - len(b)) + b
        r0 = [i for i in range(r) if a2[i] == 1 and b2[i] != 1 or i < r - len(a)]
        r1 = [i for i in range(r) if b2[i] == 1 and a2[i] != 1 or i < r - len(b)]
        return Table(torch.tensor(r0, dtype=torch.int32), torch.tensor(r1, dtype=torch.int32))
    return _Fn(fn, 2)


@op("SoftmaxCrossEntropyWithLogits")
def _softmax_xent(n):
    def fn(logits, labels):
        lp = F.log_softmax(logits.float(), -1)
        loss = -(labels.float() * lp).sum(-1)
        return Table(loss.to(logits.dtype), (lp.exp() - labels.float()).to(logits.dtype))
    return _Fn(fn, 2)


# ---------------------------------------------------------------------------------------------- pooling / LRN / resize
def _pool_geometry(n):
    nchw = _nchw(n)
    kh, kw = _hw(list(_attr(n, "ksize")), nchw)
    sh, sw = _hw(list(_attr(n, "strides")), nchw)
    pad = _attr(n, "padding", b"VALID")
    pad = pad.decode() if isinstance(pad, bytes) else pad
    return nchw, kh, kw, sh, sw, pad


def _pool_fwd(kind, x, kh, kw, sh, sw, pad):
    if pad == "SAME":
        fill = float("-inf") if kind == "max" else 0.0
        H, W = x.shape[2:]
        oh, ow = -(-H // sh), -(-W // sw)
        ph, pw = max((oh - 1) * sh + kh - H, 0), max((ow - 1) * sw + kw - W, 0)
        if kind == "avg":
            ones = F.pad(torch.ones_like(x[:1, :1]), (pw // 2, pw - pw // 2, ph // 2, ph - ph // 2))
            cnt = F.avg_pool2d(ones, (kh, kw), (sh, sw)) * (kh * kw)
            xs = F.pad(x, (pw // 2, pw - pw // 2, ph // 2, ph - ph // 2))
            return F.avg_pool2d(xs, (kh, kw), (sh, sw)) * (kh * kw) / cnt
        x = F.pad(x, (pw // 2, pw - pw // 2, ph // 2, ph - ph // 2), value=fill)
    return F.max_pool2d(x, (kh, kw), (sh, sw)) if kind == "max" else F.avg_pool2d(x, (kh, kw), (sh, sw))


@op("MaxPoolGrad")
def _max_pool_grad(n):
    nchw, kh, kw, sh, sw, pad = _pool_geometry(n)

    def fn(x, y, g):
        gx = _vjp(lambda t: _pool_fwd("max", t, kh, kw, sh, sw, pad), _to_nchw(x, nchw), _to_nchw(g, nchw))
        return _from_nchw(gx, nchw)
    return _Fn(fn, 3)


@op("AvgPoolGrad")
def _avg_pool_grad(n):
    nchw, kh, kw, sh, sw, pad = _pool_geometry(n)

    def fn(shape, g):
        shp = _ints(shape)
        x = torch.zeros(shp, dtype=g.dtype, device=g.device)
        gx = _vjp(lambda t: _pool_fwd("avg", t, kh, kw, sh, sw, pad), _to_nchw(x, nchw), _to_nchw(g, nchw))
        return _from_nchw(gx, nchw)
    return _Fn(fn, 2)


@op("LRNGrad")
def _lrn_grad(n):
    r = int(_attr(n, "depth_radius", 5))
    bias, alpha, beta = float(_attr(n, "bias", 1.0)), float(_attr(n, "alpha", 1.0)), float(_attr(n, "beta", 0.5))

    def lrn(x):                                   # NHWC in TF; torch's LRN divides alpha by the window size
        size = 2 * r + 1
        return F.local_response_norm(x.permute(0, 3, 1, 2), size, alpha * size, beta, bias).permute(0, 2, 3, 1)
    return _Fn(lambda g, x, y: _vjp(lrn, x, g), 3)


@op("ResizeBilinearGrad")
def _resize_bilinear_grad(n):
    align = bool(_attr(n, "align_corners", False))

    def fn(g, orig):
        size = tuple(g.shape[1:3])
        return _vjp(lambda t: F.interpolate(t.permute(0, 3, 1, 2), size=size, mode="bilinear", align_corners=align)
                    .permute(0, 2, 3, 1), orig, g)
    return _Fn(fn, 2)


# ---------------------------------------------------------------------------------------------- convolutions
def _conv_attrs(n, dims=2):
    nchw = _nchw(n)
    strides = list(_attr(n, "strides"))
    dil = list(_attr(n, "dilations", [1] * (dims + 2)) or [1] * (dims + 2))
    sp = slice(2, None) if nchw else slice(1, -1)
    pad = _attr(n, "padding", b"VALID")
    return nchw, tuple(strides[sp]), tuple(dil[sp]), (pad.decode() if isinstance(pad, bytes) else pad)


def _conv_nd(x, w, strides, dil, pad, groups=1):
    """x channels-first, w in torch layout [O, I/g, *k]; TF SAME padding may be asymmetric."""
    nd = x.dim() - 2
    k = w.shape[2:]
    if pad == "SAME":
        pads = []
        for i in reversed(range(nd)):
            n_in = x.shape[2 + i]
            out = -(-n_in // strides[i])
            total = max((out - 1) * strides[i] + (k[i] - 1) * dil[i] + 1 - n_in, 0)
            pads += [total // 2, total - total // 2]
        x = F.pad(x, pads)
    fn = F.conv2d if nd == 2 else F.conv3d
    return fn(x, w, None, strides, 0, dil, groups)


def _cf(x, nchw):          # channels-last -> channels-first for 4-D / 5-D
    if nchw:
        return x
    return x.permute(0, x.dim() - 1, *range(1, x.dim() - 1))


def _cl(x, nchw):
    if nchw:
        return x
    return x.permute(0, *range(2, x.dim()), 1)


def _w_torch(w):           # TF [*k, I, O] -> torch [O, I, *k]
    nd = w.dim() - 2
    return w.permute(nd + 1, nd, *range(nd))


def _w_tf(w):
    nd = w.dim() - 2
    return w.permute(*range(2, 2 + nd), 1, 0)


@op("Conv2DBackpropFilter")
def _conv2d_backprop_filter(n):
    nchw, strides, dil, pad = _conv_attrs(n)

    def fn(x, sizes, g):
        w0 = torch.zeros(_ints(sizes), dtype=torch.float32, device=x.device)
        xc, gc = _cf(x, nchw).float(), _cf(g, nchw)
        return _vjp(lambda w: _conv_nd(xc, _w_torch(w), strides, dil, pad), w0, gc).to(x.dtype)
    return _Fn(fn, 3)


@op("Conv3D")
def _conv3d(n):
    nchw, strides, dil, pad = _conv_attrs(n, 3)
    return _Fn(lambda x, w: _cl(_conv_nd(_cf(x, nchw), _w_torch(w), strides, dil, pad), nchw), 2)


@op("Conv3DBackpropInput", "Conv3DBackpropInputV2")
def _conv3d_backprop_input(n):
    nchw, strides, dil, pad = _conv_attrs(n, 3)
    v2 = n.op.endswith("V2")

    def fn(a, w, g):
        shape = _ints(a) if v2 else list(a.shape)
        x0 = torch.zeros(shape, dtype=torch.float32, device=g.device)
        wt = _w_torch(w).float()
        return _cl(_vjp(lambda x: _conv_nd(x, wt, strides, dil, pad), _cf(x0, nchw), _cf(g, nchw)), nchw).to(g.dtype)
    return _Fn(fn, 3)


@op("Conv3DBackpropFilter", "Conv3DBackpropFilterV2")
def _conv3d_backprop_filter(n):
    nchw, strides, dil, pad = _conv_attrs(n, 3)
    v2 = n.op.endswith("V2")

    def fn(x, b, g):
        shape = _ints(b) if v2 else list(b.shape)
        w0 = torch.zeros(shape, dtype=torch.float32, device=x.device)
        xc = _cf(x, nchw).float()
        return _vjp(lambda w: _conv_nd(xc, _w_torch(w), strides, dil, pad), w0, _cf(g, nchw)).to(x.dtype)
    return _Fn(fn, 3)


def _dw_torch(w):          # TF depthwise filter [kh, kw, I, M] -> torch [I*M, 1, kh, kw]
    kh, kw, i, m = w.shape
    return w.permute(2, 3, 0, 1).reshape(i * m, 1, kh, kw)


@op("DepthwiseConv2dNativeBackpropInput")
def _dw_backprop_input(n):
    nchw, strides, dil, pad = _conv_attrs(n)

    def fn(sizes, w, g):
        x0 = torch.zeros(_ints(sizes), dtype=torch.float32, device=g.device)
        wt = _dw_torch(w).float()
        groups = w.shape[2]
        return _cl(_vjp(lambda x: _conv_nd(x, wt, strides, dil, pad, groups), _cf(x0, nchw), _cf(g, nchw)), nchw).to(g.dtype)
    return _Fn(fn, 3)


@op("DepthwiseConv2dNativeBackpropFilter")
def _dw_backprop_filter(n):
    nchw, strides, dil, pad = _conv_attrs(n)

    def fn(x, sizes, g):
        w0 = torch.zeros(_ints(sizes), dtype=torch.float32, device=x.device)
        xc = _cf(x, nchw).float()
        groups = w0.shape[2]
        return _vjp(lambda w: _conv_nd(xc, _dw_torch(w), strides, dil, pad, groups), w0, _cf(g, nchw)).to(x.dtype)
    return _Fn(fn, 3)


# ---------------------------------------------------------------------------------------------- morphological dilation
def _dilation2d(x, w, strides, rates, pad):
    """Grayscale dilation (TF Dilation2D): out[b,y,x,c] = max_{dy,dx} in[b, y*s+dy*r, x*s+dx*r, c] + filter[dy,dx,c].
    NHWC input, filter [kh, kw, C]."""
    kh, kw, c = w.shape
    sh, sw = strides
    rh, rw = rates
    xc = x.permute(0, 3, 1, 2)
    if pad == "SAME":
        H, W = xc.shape[2:]
        oh, ow = -(-H // sh), -(-W // sw)
        ph = max((oh - 1) * sh + (kh - 1) * rh + 1 - H, 0)
        pw = max((ow - 1) * sw + (kw - 1) * rw + 1 - W, 0)
        xc = F.pad(xc, (pw // 2, pw - pw // 2, ph // 2, ph - ph // 2), value=float("-inf"))
    cols = F.unfold(xc, (kh, kw), dilation=(rh, rw), stride=(sh, sw))            # [B, C*kh*kw, L]
    B, _, L = cols.shape
    cols = cols.view(B, c, kh * kw, L) + w.permute(2, 0, 1).reshape(1, c, kh * kw, 1)
    oh = (xc.shape[2] - (kh - 1) * rh - 1) // sh + 1
    return cols.amax(2).view(B, c, oh, -1).permute(0, 2, 3, 1)


def _dil_attrs(n):
    s, r = list(_attr(n, "strides")), list(_attr(n, "rates"))
    pad = _attr(n, "padding", b"VALID")
    return (s[1], s[2]), (r[1], r[2]), (pad.decode() if isinstance(pad, bytes) else pad)


@op("Dilation2D")
def _dilation2d_loader(n):
    st, rt, pad = _dil_attrs(n)
    return _Fn(lambda x, w: _dilation2d(x.float(), w.float(), st, rt, pad).to(x.dtype), 2)


@op("Dilation2DBackpropInput")
def _dilation_backprop_input(n):
    st, rt, pad = _dil_attrs(n)
    return _Fn(lambda x, w, g: _vjp(lambda t: _dilation2d(t, w.float(), st, rt, pad), x, g), 3)


@op("Dilation2DBackpropFilter")
def _dilation_backprop_filter(n):
    st, rt, pad = _dil_attrs(n)
    return _Fn(lambda x, w, g: _vjp(lambda t: _dilation2d(x.float(), t, st, rt, pad), w, g), 3)


# ---------------------------------------------------------------------------------------------- batch norm
@op("FusedBatchNormGrad", "FusedBatchNormGradV2", "FusedBatchNormGradV3")
def _fused_bn_grad(n):
    nchw = _nchw(n)
    eps = float(_attr(n, "epsilon", 1e-3))
    training = bool(_attr(n, "is_training", True))

    def fn(dy, x, scale, mean, var, *rest):
        xc, gc = _cf(x, nchw).float(), _cf(dy, nchw).float()
        dims = [0] + list(range(2, xc.dim()))
        shp = [1, -1] + [1] * (xc.dim() - 2)
        with torch.enable_grad():
            xr = xc.detach().requires_grad_(True)
            sr = scale.detach().float().requires_grad_(True)
            if training:
                mu, v = xr.mean(dims), xr.var(dims, unbiased=False)
            else:
                mu, v = mean.float(), var.float()
            y = (xr - mu.view(shp)) * torch.rsqrt(v.view(shp) + eps) * sr.view(shp)
            dx, dscale = torch.autograd.grad(y, (xr, sr), gc)
        doffset = gc.sum(dims)
        z = torch.zeros(0, dtype=dy.dtype, device=dy.device)
        return Table(_cl(dx, nchw).to(dy.dtype), dscale.to(scale.dtype), doffset.to(scale.dtype), z, z)
    return _Fn(fn)


# ---------------------------------------------------------------------------------------------- records
_TORCH_OF_DT = {1: torch.float32, 2: torch.float64, 3: torch.int32, 4: torch.uint8, 5: torch.int16, 6: torch.int8,
                9: torch.int64}


@op("DecodeRaw")
def _decode_raw(n):
    from ..nn import tf as TFO

    return TFO.DecodeRaw(_TORCH_OF_DT.get(int(_attr(n, "out_type", 4)), torch.uint8), bool(_attr(n, "little_endian", True)))


def _dense_spec(n):
    types = list(n.attr["Tdense"].list.type) if "Tdense" in n.attr else []
    names = {1: "float", 2: "float", 7: "string"}
    shapes = [[int(d.size) for d in sh.dim] for sh in n.attr["dense_shapes"].list.shape] if "dense_shapes" in n.attr else []
    while len(shapes) < len(types):
        shapes.append([])
    return [names.get(int(t), "int64") for t in types], shapes


@op("ParseExample")
def _parse_example(n):
    from ..nn import tf as TFO

    types, shapes = _dense_spec(n)
    return TFO.ParseExample(int(_attr(n, "Ndense", len(types)) or len(types)), types, shapes)


@op("ParseSingleExample")
def _parse_single_example(n):
    """Keys come from the ``dense_keys`` attribute; without ``Tdense`` the types follow the record itself."""
    from ..nn import tf as TFO

    types, shapes = _dense_spec(n)
    keys = list(n.attr["dense_keys"].list.s) if "dense_keys" in n.attr else []
    if len(types) < len(keys):
        types = types + [None] * (len(keys) - len(types))
    return TFO.ParseSingleExample(types, keys, shapes)


@op("RandomShuffle")
def _random_shuffle(n):
    def fn(x):
        perm = torch.randperm(x.shape[0], device=x.device if isinstance(x, torch.Tensor) else None)
        return x[perm]
    return _Fn(fn, 1)
