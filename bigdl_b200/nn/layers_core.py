"""The hot-path layers: Linear, SpatialConvolution family, pooling, LRN, BatchNorm, ReLU, Dropout, (Log)SoftMax.

On CUDA tensors these dispatch to the sm_100a kernels (``bigdl_b200.ops.functions``); on CPU tensors they run plain
fp32 PyTorch ops, which double as the numerics oracle for the kernels (the role Torch7 ``TH`` plays in the reference's
tests, T/torch/TH.scala).

Reference files: S/nn/Linear.scala, SpatialConvolution.scala, SpatialShareConvolution.scala,
SpatialDilatedConvolution.scala, SpatialMaxPooling.scala, SpatialAveragePooling.scala, SpatialCrossMapLRN.scala,
BatchNormalization.scala, SpatialBatchNormalization.scala, ReLU.scala, Dropout.scala, SoftMax.scala, LogSoftMax.scala,
and the mkldnn twins under S/nn/mkldnn/.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from ..utils import config
from ..utils.random_generator import RNG
from .initialization_method import (InitializationMethod, Ones, RandomUniform, VariableFormat, Zeros)
from .module import Module


def native(x: torch.Tensor) -> bool:
    """True when the hand-written kernel path should run for this tensor."""
    return x.is_cuda and config.get("bigdl.engineType") == "b200"


def _compute_dtype(x: torch.Tensor) -> torch.dtype:
    return torch.bfloat16 if x.is_cuda else torch.float32


def conv_out_size(in_size: int, k: int, stride: int, pad: int, dilation: int = 1, ceil_mode: bool = False) -> int:
    """Torch output-size rule used by the reference (S/nn/Utils.scala:340-390)."""
    dk = dilation * (k - 1) + 1
    num = in_size - dk + 2 * pad
    o = (math.ceil(num / stride) if ceil_mode else math.floor(num / stride)) + 1
    if pad != 0 and (o - 1) * stride >= in_size + pad:
        o -= 1
    return int(o)


def same_padding(in_size: int, k: int, stride: int) -> tuple[int, int, int]:
    """TF-style SAME: (pad_before, pad_after, out) (S/nn/Utils.scala:307-332)."""
    o = math.ceil(in_size / stride)
    total = max(0, (o - 1) * stride + k - in_size)
    return total // 2, total - total // 2, o


class Linear(Module):
    """y = x W^T + b, weight [outputSize, inputSize] (S/nn/Linear.scala:78-167)."""

    def __init__(self, input_size: int, output_size: int, with_bias: bool = True, wRegularizer=None,
                 bRegularizer=None, init_weight=None, init_bias=None, init_grad_weight=None, init_grad_bias=None):
        super().__init__()
        self.input_size, self.output_size, self.with_bias = input_size, output_size, with_bias
        self.wRegularizer, self.bRegularizer = wRegularizer, bRegularizer
        self.register_parameter("weight", torch.zeros(output_size, input_size))
        if with_bias:
            self.register_parameter("bias", torch.zeros(output_size))
        self._fuse_relu = False
        self.reset()
        with torch.no_grad():
            if init_weight is not None:
                self.weight.copy_(torch.as_tensor(init_weight, dtype=torch.float32).reshape(self.weight.shape))
            if init_bias is not None and with_bias:
                self.bias.copy_(torch.as_tensor(init_bias, dtype=torch.float32).reshape(self.bias.shape))

    def reset(self):
        wi = self._init_methods.get("weight", RandomUniform())
        bi = self._init_methods.get("bias", RandomUniform())
        wi(self.weight.data, VariableFormat.OUT_IN)
        if self.with_bias:
            if isinstance(bi, RandomUniform) and bi.lower is None:
                stdv = 1.0 / math.sqrt(self.input_size)
                RNG.fill_uniform(self.bias.data, -stdv, stdv)
            else:
                bi(self.bias.data, VariableFormat.ONE_D)
        self.zeroGradParameters()
        self._mark_weights_dirty()
        return self

    def updateOutput(self, input):
        bias = self._parameters.get("bias")
        if native(input) and self.input_size % 8 == 0:
            from ..ops.functions import LinearFn

            w16 = self.weight_for_compute("weight")
            if w16.dtype != torch.bfloat16:
                w16 = w16.detach().to(torch.bfloat16)
            # sharded parameter plane: a callable that names the weight rows to gather from their owners during this
            # training step (parallel/bucketed.py), absent otherwise
            gp = self.__dict__.get("_gather_plan")
            return LinearFn.apply(input, self.weight, bias, w16, self._fuse_relu, self.scaleW, self.scaleB,
                                  gp() if gp is not None else None)
        w = self.weight
        if input.dtype != w.dtype:
            w = w.to(input.dtype)
            bias = bias.to(input.dtype) if bias is not None else None
        y = F.linear(input, w, bias)
        return F.relu(y) if self._fuse_relu else y


class SpatialConvolution(Module):
    """2-D convolution (S/nn/SpatialConvolution.scala). Weight is exposed as [nOutputPlane, nInputPlane/nGroup, kH, kW]
    (the reference's 5-D (nGroup, out/g, in/g, kH, kW) view with the group axis folded into the output axis) and
    stored channels-last so the GEMM sees [out, kH*kW*in]. ``pad_w = pad_h = -1`` selects TF-style SAME padding."""

    def __init__(self, n_input_plane: int, n_output_plane: int, kernel_w: int, kernel_h: int, stride_w: int = 1,
                 stride_h: int = 1, pad_w: int = 0, pad_h: int = 0, n_group: int = 1, propagate_back: bool = True,
                 wRegularizer=None, bRegularizer=None, init_weight=None, init_bias=None, init_grad_weight=None,
                 init_grad_bias=None, with_bias: bool = True, data_format: str = "NCHW", dilation_w: int = 1,
                 dilation_h: int = 1):
        super().__init__()
        if n_input_plane % n_group or n_output_plane % n_group:
            raise ValueError("Number of input/output channels should be multiples of group.")
        self.n_input_plane, self.n_output_plane = n_input_plane, n_output_plane
        self.kernel_w, self.kernel_h, self.stride_w, self.stride_h = kernel_w, kernel_h, stride_w, stride_h
        self.pad_w, self.pad_h, self.n_group = pad_w, pad_h, n_group
        self.dilation_w, self.dilation_h = dilation_w, dilation_h
        self.propagate_back, self.with_bias, self.data_format = propagate_back, with_bias, data_format.upper()
        self.wRegularizer, self.bRegularizer = wRegularizer, bRegularizer
        w = torch.zeros(n_output_plane, n_input_plane // n_group, kernel_h, kernel_w).contiguous(
            memory_format=torch.channels_last)
        self.register_parameter("weight", w)
        if with_bias:
            self.register_parameter("bias", torch.zeros(n_output_plane))
        self._fuse_relu = False
        self.reset()
        with torch.no_grad():
            if init_weight is not None:
                self.weight.copy_(torch.as_tensor(init_weight, dtype=torch.float32).reshape(self.weight.shape))
            if init_bias is not None and with_bias:
                self.bias.copy_(torch.as_tensor(init_bias, dtype=torch.float32).reshape(self.bias.shape))

    def register_parameter(self, name, value):
        p = super().register_parameter(name, value)
        if value.dim() == 4:   # keep the OHWI physical layout through clone()
            q = value.detach().clone(memory_format=torch.channels_last)
            q.requires_grad_(True)
            q.grad = torch.zeros_like(q, memory_format=torch.channels_last)
            self._parameters[name] = q
            return q
        return p

    def reset(self):
        wi = self._init_methods.get("weight", RandomUniform())
        bi = self._init_methods.get("bias", RandomUniform())
        tmp = torch.zeros(self.weight.shape)
        wi(tmp, VariableFormat.OUT_IN_KW_KH)
        with torch.no_grad():
            self.weight.copy_(tmp)
        if self.with_bias:
            if isinstance(bi, RandomUniform) and bi.lower is None:
                fan_in = (self.n_input_plane // self.n_group) * self.kernel_h * self.kernel_w
                stdv = 1.0 / math.sqrt(fan_in)
                RNG.fill_uniform(self.bias.data, -stdv, stdv)
            else:
                bi(self.bias.data, VariableFormat.ONE_D)
        self.zeroGradParameters()
        self._mark_weights_dirty()
        return self

    def _padding(self, H: int, W: int):
        if self.pad_w == -1 or self.pad_h == -1:
            dkh = self.dilation_h * (self.kernel_h - 1) + 1
            dkw = self.dilation_w * (self.kernel_w - 1) + 1
            pt, pb, _ = same_padding(H, dkh, self.stride_h)
            pl, pr, _ = same_padding(W, dkw, self.stride_w)
            return pt, pl, pb, pr
        return self.pad_h, self.pad_w, self.pad_h, self.pad_w

    def updateOutput(self, input):
        x = input
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        if self.data_format == "NHWC":
            x = x.permute(0, 3, 1, 2)
        pt, pl, pb, pr = self._padding(x.shape[2], x.shape[3])
        bias = self._parameters.get("bias")
        if native(x):
            from ..ops.functions import ConvFn

            w16 = self.weight_for_compute("weight")
            if w16.dtype != torch.bfloat16:
                w16 = w16.detach().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
            geom = (self.kernel_h, self.kernel_w, self.stride_h, self.stride_w, pt, pl, pb, pr, self.dilation_h,
                    self.dilation_w, self.n_group)
            xin = x if self.propagate_back else x.detach()
            res = self.__dict__.pop("_next_residual", None)     # conv + sum fusion (set by a fused ConcatTable)
            if res is not None:
                from ..ops.functions import set_residual_hint

                set_residual_hint(res)
            hint = self.__dict__.pop("_next_out_hint", None)
            if hint is not None:
                from ..ops.functions import set_out_hint

                set_out_hint(hint)
            stats = None
            if self.__dict__.get("_bn_stats") and self.train and not self._fuse_relu:
                # the next layer is a BatchNorm (nn/fusion.py): let the GEMM epilogue sum the output columns
                from ..ops.functions import zeros_f32

                stats = zeros_f32(2 * self.n_output_plane, x.device)
            y = ConvFn.apply(xin, self.weight, bias, w16, geom, self._fuse_relu, self.scaleW, self.scaleB,
                             self.__dict__.pop("_skip_token", None),      # residual-block input pairing (SkipToken)
                             stats,
                             # our input is the fused-ReLU output of the convolution right before us and we are its
                             # only consumer (nn/fusion.py): its ReLU backward + bias sums run in our dgrad epilogue
                             bool(self.__dict__.get("_mask_input")) and self.train)
            if stats is not None and getattr(stats, "_bdl_done", False):
                y._bdl_bn_stats = stats

        else:
            w = self.weight
            if x.dtype != w.dtype:
                w = w.to(x.dtype)
                bias = bias.to(x.dtype) if bias is not None else None
            xin = x if self.propagate_back else x.detach()
            if (pt, pl) != (pb, pr):
                xin = F.pad(xin, (pl, pr, pt, pb))
                pad = (0, 0)
            else:
                pad = (pt, pl)
            y = F.conv2d(xin, w, bias, (self.stride_h, self.stride_w), pad, (self.dilation_h, self.dilation_w),
                         self.n_group)
            if self._fuse_relu:
                y = F.relu(y)
        if self.data_format == "NHWC":
            y = y.permute(0, 2, 3, 1)
        return y.squeeze(0) if squeeze else y


class SpatialShareConvolution(SpatialConvolution):
    """Same math as SpatialConvolution; the reference variant only shares im2col buffers between replicas
    (S/nn/SpatialShareConvolution.scala) — here every conv already shares one device workspace."""


class SpatialDilatedConvolution(SpatialConvolution):
    """Dilated ("atrous") convolution (S/nn/SpatialDilatedConvolution.scala)."""

    def __init__(self, n_input_plane, n_output_plane, kw, kh, dw=1, dh=1, pad_w=0, pad_h=0, dilation_w=1,
                 dilation_h=1, wRegularizer=None, bRegularizer=None):
        super().__init__(n_input_plane, n_output_plane, kw, kh, dw, dh, pad_w, pad_h, 1, True, wRegularizer,
                         bRegularizer, dilation_w=dilation_w, dilation_h=dilation_h)


def _pool_geometry(mod, H, W):
    if mod.pad_h == -1 or mod.pad_w == -1:
        pt, _, Ho = same_padding(H, mod.kh, mod.dh)
        pl, _, Wo = same_padding(W, mod.kw, mod.dw)
        return pt, pl, Ho, Wo
    Ho = conv_out_size(H, mod.kh, mod.dh, mod.pad_h, 1, mod.ceil_mode)
    Wo = conv_out_size(W, mod.kw, mod.dw, mod.pad_w, 1, mod.ceil_mode)
    return mod.pad_h, mod.pad_w, Ho, Wo


class SpatialMaxPooling(Module):
    """Max pooling; ``.ceil()`` switches to ceil-mode output sizes (S/nn/SpatialMaxPooling.scala)."""

    def __init__(self, kw: int, kh: int, dw: int | None = None, dh: int | None = None, pad_w: int = 0, pad_h: int = 0,
                 to_ceil: bool = False, format: str = "NCHW"):
        super().__init__()
        self.kw, self.kh = kw, kh
        self.dw, self.dh = (dw or kw), (dh or kh)
        self.pad_w, self.pad_h = pad_w, pad_h
        self.ceil_mode = to_ceil
        self.data_format = format.upper()

    def ceil(self):
        self.ceil_mode = True
        return self

    def floor(self):
        self.ceil_mode = False
        return self

    def updateOutput(self, input):
        x = input
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        if self.data_format == "NHWC":
            x = x.permute(0, 3, 1, 2)
        H, W = x.shape[2], x.shape[3]
        ph, pw, Ho, Wo = _pool_geometry(self, H, W)
        if native(x) and x.shape[1] % 8 == 0 and self.kh * self.kw <= 127:
            from ..ops.functions import MaxPoolFn

            y = MaxPoolFn.apply(x, self.kh, self.kw, self.dh, self.dw, ph, pw, Ho, Wo)
        else:
            need_h = (Ho - 1) * self.dh + self.kh - H - ph
            need_w = (Wo - 1) * self.dw + self.kw - W - pw
            xp = F.pad(x, (pw, max(need_w, 0), ph, max(need_h, 0)), value=float("-inf"))
            y = F.max_pool2d(xp, (self.kh, self.kw), (self.dh, self.dw))[:, :, :Ho, :Wo]
        if self.data_format == "NHWC":
            y = y.permute(0, 2, 3, 1)
        return y.squeeze(0) if squeeze else y


class SpatialAveragePooling(Module):
    """Average pooling with ceil mode, countIncludePad and divide flags; ``global_pooling`` pools the whole map
    (S/nn/SpatialAveragePooling.scala)."""

    def __init__(self, kw: int, kh: int, dw: int = 1, dh: int = 1, pad_w: int = 0, pad_h: int = 0,
                 global_pooling: bool = False, ceil_mode: bool = False, count_include_pad: bool = True,
                 divide: bool = True, format: str = "NCHW"):
        super().__init__()
        self.kw, self.kh, self.dw, self.dh = kw, kh, dw, dh
        self.pad_w, self.pad_h = pad_w, pad_h
        self.global_pooling, self.ceil_mode = global_pooling, ceil_mode
        self.count_include_pad, self.divide = count_include_pad, divide
        self.data_format = format.upper()

    def ceil(self):
        self.ceil_mode = True
        return self

    def floor(self):
        self.ceil_mode = False
        return self

    def updateOutput(self, input):
        x = input
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        if self.data_format == "NHWC":
            x = x.permute(0, 3, 1, 2)
        H, W = x.shape[2], x.shape[3]
        if self.global_pooling:
            kh, kw, dh, dw = H, W, 1, 1
            ph = pw = 0
            Ho = Wo = 1
        else:
            kh, kw, dh, dw = self.kh, self.kw, self.dh, self.dw
            ph, pw, Ho, Wo = _pool_geometry(self, H, W)
        if native(x) and x.shape[1] % 8 == 0:
            from ..ops.functions import AvgPoolFn

            y = AvgPoolFn.apply(x, kh, kw, dh, dw, ph, pw, Ho, Wo, self.count_include_pad, self.divide)
        else:
            y = _avg_pool_reference(x, kh, kw, dh, dw, ph, pw, Ho, Wo, self.count_include_pad, self.divide)
        if self.data_format == "NHWC":
            y = y.permute(0, 2, 3, 1)
        return y.squeeze(0) if squeeze else y


def _avg_pool_reference(x, kh, kw, dh, dw, ph, pw, Ho, Wo, count_include_pad, divide):
    """Window sums via unfold-free conv with a ones kernel, divisor computed analytically (handles ceil mode)."""
    N, C, H, W = x.shape
    need_h = max((Ho - 1) * dh + kh - H - ph, 0)
    need_w = max((Wo - 1) * dw + kw - W - pw, 0)
    xp = F.pad(x, (pw, need_w, ph, need_h))
    s = F.avg_pool2d(xp, (kh, kw), (dh, dw), divisor_override=1)[:, :, :Ho, :Wo]
    if not divide:
        return s
    hs = torch.arange(Ho, device=x.device) * dh - ph
    ws = torch.arange(Wo, device=x.device) * dw - pw
    he = torch.clamp(hs + kh, max=H + ph)
    we = torch.clamp(ws + kw, max=W + pw)
    if count_include_pad:
        div = (he - hs)[:, None] * (we - ws)[None, :]
    else:
        div = (torch.clamp(he, max=H) - torch.clamp(hs, min=0))[:, None] * \
              (torch.clamp(we, max=W) - torch.clamp(ws, min=0))[None, :]
        div = torch.clamp(div, min=1)
    return s / div.to(s.dtype)


class SpatialCrossMapLRN(Module):
    """y = x * (k + alpha/size * sum_{window over channels} x^2)^-beta (S/nn/SpatialCrossMapLRN.scala:204-237)."""

    def __init__(self, size: int = 5, alpha: float = 1.0, beta: float = 0.75, k: float = 1.0, data_format="NCHW"):
        super().__init__()
        self.size, self.alpha, self.beta, self.k = size, alpha, beta, k
        self.data_format = data_format.upper()

    def updateOutput(self, input):
        x = input
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        if self.data_format == "NHWC":
            x = x.permute(0, 3, 1, 2)
        if native(x) and x.shape[1] % 8 == 0:
            from ..ops.functions import LRNFn

            y = LRNFn.apply(x, self.size, float(self.alpha), float(self.beta), float(self.k))
        else:
            pre = (self.size - 1) // 2
            post = self.size - 1 - pre
            sq = F.pad((x * x).unsqueeze(1), (0, 0, 0, 0, pre, post))
            ssum = F.avg_pool3d(sq, (self.size, 1, 1), stride=1, divisor_override=1).squeeze(1)
            y = x * (self.k + self.alpha / self.size * ssum).pow(-self.beta)
        if self.data_format == "NHWC":
            y = y.permute(0, 2, 3, 1)
        return y.squeeze(0) if squeeze else y


class ReLU(Module):
    def __init__(self, ip: bool = False):
        super().__init__()
        self.inplace = ip

    def updateOutput(self, input):
        if getattr(self, "_fused", False):      # folded into the producer's epilogue (nn/fusion.py)
            return input
        if native(input) and input.dtype == torch.bfloat16:
            from ..ops.functions import ReLUFn

            return ReLUFn.apply(input)
        return F.relu(input)


class Dropout(Module):
    """Drop with probability ``init_p``; kept activations are scaled by 1/(1-p) when ``scale`` (S/nn/Dropout.scala:40-75)."""

    def __init__(self, init_p: float = 0.5, inplace: bool = False, scale: bool = True):
        super().__init__()
        self.p, self.inplace, self.scale = init_p, inplace, scale

    def setP(self, p: float):
        self.p = p
        return self

    def updateOutput(self, input):
        if not self.train or self.p <= 0:
            return input
        if self.p >= 1:
            return input * 0
        if native(input) and input.dtype == torch.bfloat16 and self.scale:
            from ..ops.functions import DropoutFn

            return DropoutFn.apply(input, float(self.p), RNG.random() | (RNG.random() << 32))
        keep = torch.empty(input.shape, dtype=torch.float32)
        RNG.fill_bernoulli(keep, 1 - self.p)
        keep = keep.to(device=input.device, dtype=input.dtype)
        return input * keep / (1 - self.p) if self.scale else input * keep


class BatchNormalization(Module):
    """Batch norm over dim 1 of a 2-D (or N-D, channel-first) input: eps 1e-5, momentum 0.1, gamma ~ U(0,1), beta 0,
    running variance fed with the *unbiased* batch variance (S/nn/BatchNormalization.scala:40-110,
    SpatialBatchNormalization.scala:485-494)."""

    n_dim = 2

    def __init__(self, n_output: int, eps: float = 1e-5, momentum: float = 0.1, affine: bool = True, init_weight=None,
                 init_bias=None, init_grad_weight=None, init_grad_bias=None, data_format: str = "NCHW"):
        super().__init__()
        self.n_output, self.eps, self.momentum, self.affine = n_output, eps, momentum, affine
        self.data_format = data_format.upper()
        if affine:
            self.register_parameter("weight", torch.zeros(n_output))
            self.register_parameter("bias", torch.zeros(n_output))
        self.register_buffer("runningMean", torch.zeros(n_output))
        self.register_buffer("runningVar", torch.ones(n_output))
        self._fuse_relu = False
        self._sync = None   # set by setParallism / sync-BN
        self.reset()
        with torch.no_grad():
            if init_weight is not None and affine:
                self.weight.copy_(torch.as_tensor(init_weight, dtype=torch.float32).reshape(-1))
            if init_bias is not None and affine:
                self.bias.copy_(torch.as_tensor(init_bias, dtype=torch.float32).reshape(-1))

    def reset(self):
        if self.affine:
            wi = self._init_methods.get("weight", RandomUniform(0.0, 1.0))
            bi = self._init_methods.get("bias", Zeros())
            wi(self.weight.data, VariableFormat.ONE_D)
            bi(self.bias.data, VariableFormat.ONE_D)
            self.zeroGradParameters()
        self._buffers["runningMean"].zero_()
        self._buffers["runningVar"].fill_(1.0)
        self._mark_weights_dirty()
        return self

    def setParallism(self, n: int):
        """Reference API for intra-node sync-BN (S/nn/BatchNormalization.scala:65-76). With one replica per GPU the
        cross-replica statistics come from ``bigdl_b200.parallel.sync_bn`` instead."""
        self._sync_parallelism = n
        return self

    def updateOutput(self, input):
        if getattr(self, "_folded", False):     # folded into the preceding convolution (nn/fusion.py fold_bn)
            return F.relu(input) if self._fuse_relu else input
        x = input
        if self.data_format == "NHWC" and x.dim() == 4:
            x = x.permute(0, 3, 1, 2)
        w = self._parameters.get("weight")
        b = self._parameters.get("bias")
        rm, rv = self._buffers["runningMean"], self._buffers["runningVar"]
        if native(x) and x.dim() == 4 and x.shape[1] % 8 == 0 and x.dtype == torch.bfloat16:
            from ..ops.batchnorm import BatchNormFn

            if self.__dict__.pop("_defer_now", False) and self.data_format != "NHWC" and not self._fuse_relu:
                # last layer of a residual branch: the CAddTable that follows runs bn + add (+ ReLU) as ONE kernel
                # (nn/fusion.py marks the pattern, ConcatTable arms this flag for exactly one call)
                out = x.view_as(x)
                out._bdl_bn = self
                st = getattr(x, "_bdl_bn_stats", None)
                if st is not None:
                    out._bdl_bn_stats = st
                return out
            y = BatchNormFn.apply(x, w, b, rm, rv, self.train, self.momentum, self.eps, self._fuse_relu, self._sync,
                                  None, None, getattr(x, "_bdl_bn_stats", None) if self.train else None)
        else:
            xf = x.float() if x.dtype != torch.float32 else x
            y = F.batch_norm(xf, rm, rv, w, b, self.train, self.momentum, self.eps)
            if self._fuse_relu:
                y = F.relu(y)
            if y.dtype != x.dtype:
                y = y.to(x.dtype)
        if self.data_format == "NHWC" and y.dim() == 4:
            y = y.permute(0, 2, 3, 1)
        return y


class SpatialBatchNormalization(BatchNormalization):
    n_dim = 4


def bn_add_relu(bn: BatchNormalization, pre: torch.Tensor, shortcut: torch.Tensor, relu: bool,
                skip_token=None) -> torch.Tensor:
    """relu?(bn(pre) + shortcut) in one pass for a deferred BN (see ``BatchNormalization.updateOutput``).
    ``skip_token``: the shortcut IS the block input (identity shortcut) and the branch's first convolution holds the
    same token, so the two input gradients are summed in that convolution's dgrad epilogue."""
    from ..ops.batchnorm import BatchNormFn

    return BatchNormFn.apply(pre, bn._parameters.get("weight"), bn._parameters.get("bias"), bn._buffers["runningMean"],
                             bn._buffers["runningVar"], bn.train, bn.momentum, bn.eps, relu, bn._sync, shortcut,
                             skip_token, getattr(pre, "_bdl_bn_stats", None) if bn.train else None)


class SoftMax(Module):
    def __init__(self, pos: int = 1):
        super().__init__()
        self.pos = pos

    def updateOutput(self, input):
        dim = input.dim() - self.pos if input.dim() > 1 else 0
        dim = 1 if input.dim() in (2, 4) else (0 if input.dim() in (1, 3) else dim)
        if native(input) and dim == input.dim() - 1:
            from ..ops.functions import SoftmaxFn

            return SoftmaxFn.apply(input, False, input.dtype if input.dtype in (torch.bfloat16, torch.float32) else torch.float32)
        return F.softmax(input.float(), dim=dim).to(input.dtype)


class SoftMin(Module):
    def updateOutput(self, input):
        dim = 1 if input.dim() in (2, 4) else 0
        return F.softmin(input.float(), dim=dim).to(input.dtype)


class LogSoftMax(Module):
    """log(softmax(x)) over the class dimension (S/nn/LogSoftMax.scala). During training with a ClassNLLCriterion the
    optimizer marks the model's final LogSoftMax ``_defer_to_criterion``: it then hands the logits through, tagged, and
    the criterion runs the fused log-softmax + NLL kernels on them (csrc/loss_softmax.cu) — one pass over the logits
    instead of a materialised log-probability tensor plus gather/sum kernels."""

    def updateOutput(self, input):
        dim = 1 if input.dim() >= 2 else 0
        if (self.__dict__.get("_defer_to_criterion") and self.isTraining() and torch.is_grad_enabled() and native(input)
                and input.dim() == 2):
            out = input.view_as(input)
            out._bdl_logits = True
            return out
        if native(input) and dim == input.dim() - 1:
            from ..ops.functions import SoftmaxFn

            return SoftmaxFn.apply(input, True, torch.float32)
        return F.log_softmax(input.float(), dim=dim)
