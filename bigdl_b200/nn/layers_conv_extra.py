"""Convolution variants, temporal/volumetric pooling, resizing and ROI layers.

Reference files (S/nn/): SpatialFullConvolution, SpatialSeparableConvolution, SpatialConvolutionMap,
TemporalConvolution, VolumetricConvolution, VolumetricFullConvolution, LocallyConnected1D, LocallyConnected2D,
TemporalMaxPooling, VolumetricMaxPooling, VolumetricAveragePooling, ResizeBilinear, UpSampling1D/2D/3D, RoiPooling,
RoiAlign, Pooler.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from ..utils.random_generator import RNG
from ..utils.table import Table, to_table
from .initialization_method import BilinearFiller, RandomUniform, VariableFormat, Zeros
from .layers_core import conv_out_size, same_padding
from .module import Container, Module


def _init_conv(mod, weight, bias, fan_in, fmt):
    wi = mod._init_methods.get("weight", RandomUniform())
    bi = mod._init_methods.get("bias", RandomUniform())
    if isinstance(wi, RandomUniform) and wi.lower is None:
        RNG.fill_uniform(weight.data, -1.0 / math.sqrt(fan_in), 1.0 / math.sqrt(fan_in))
    else:
        wi(weight.data, fmt)
    if bias is not None:
        if isinstance(bi, RandomUniform) and bi.lower is None:
            RNG.fill_uniform(bias.data, -1.0 / math.sqrt(fan_in), 1.0 / math.sqrt(fan_in))
        else:
            bi(bias.data, VariableFormat.ONE_D)
    mod.zeroGradParameters()
    mod._mark_weights_dirty()


class SpatialFullConvolution(Module):
    """Transposed convolution ("deconvolution") with ``adj`` output padding; accepts a table (input, reference) to
    derive adj from the reference tensor's size (S/nn/SpatialFullConvolution.scala). Weight [nIn, nOut/g, kH, kW]."""

    def __init__(self, n_input_plane, n_output_plane, kw, kh, dw=1, dh=1, pad_w=0, pad_h=0, adj_w=0, adj_h=0,
                 n_group=1, no_bias=False, wRegularizer=None, bRegularizer=None):
        super().__init__()
        if adj_w > dw - 1 or adj_h > dh - 1:
            raise ValueError("adjW and adjH must be smaller than dW - 1 and dH - 1 respectively")
        self.nin, self.nout, self.kw, self.kh, self.dw, self.dh = n_input_plane, n_output_plane, kw, kh, dw, dh
        self.pad_w, self.pad_h, self.adj_w, self.adj_h, self.n_group = pad_w, pad_h, adj_w, adj_h, n_group
        self.wRegularizer, self.bRegularizer = wRegularizer, bRegularizer
        self.register_parameter("weight", torch.zeros(n_input_plane, n_output_plane // n_group, kh, kw))
        if not no_bias:
            self.register_parameter("bias", torch.zeros(n_output_plane))
        self.reset()

    def reset(self):
        _init_conv(self, self.weight, self._parameters.get("bias"), (self.nout // self.n_group) * self.kh * self.kw,
                   VariableFormat.IN_OUT_KW_KH)
        return self

    def updateOutput(self, input):
        adj_h, adj_w = self.adj_h, self.adj_w
        x = input
        if not isinstance(input, torch.Tensor):
            x, ref = to_table(input).to_list()
            th, tw = ref.shape[-2], ref.shape[-1]
            adj_h = th - ((x.shape[-2] - 1) * self.dh - 2 * self.pad_h + self.kh)
            adj_w = tw - ((x.shape[-1] - 1) * self.dw - 2 * self.pad_w + self.kw)
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        y = F.conv_transpose2d(x, self.weight.to(x.dtype), None if "bias" not in self._parameters else
                               self.bias.to(x.dtype), (self.dh, self.dw), (self.pad_h, self.pad_w), (adj_h, adj_w),
                               self.n_group)
        return y.squeeze(0) if squeeze else y


class SpatialSeparableConvolution(Module):
    """Depthwise conv (depth multiplier) followed by 1x1 pointwise conv (S/nn/SpatialSeparableConvolution.scala)."""

    def __init__(self, n_input_channel, n_output_channel, depth_multiplier, kw, kh, sw=1, sh=1, pw=0, ph=0,
                 with_bias=True, data_format="NCHW", w_regularizer=None, b_regularizer=None, p_regularizer=None):
        super().__init__()
        self.nin, self.nout, self.mult = n_input_channel, n_output_channel, depth_multiplier
        self.kw, self.kh, self.sw, self.sh, self.pw, self.ph = kw, kh, sw, sh, pw, ph
        self.data_format = data_format.upper()
        self.register_parameter("depthWeight", torch.zeros(n_input_channel * depth_multiplier, 1, kh, kw))
        self.register_parameter("pointWeight", torch.zeros(n_output_channel, n_input_channel * depth_multiplier, 1, 1))
        if with_bias:
            self.register_parameter("bias", torch.zeros(n_output_channel))
        self.reset()

    def reset(self):
        RNG.fill_uniform(self.depthWeight.data, -1.0 / math.sqrt(self.kh * self.kw), 1.0 / math.sqrt(self.kh * self.kw))
        RNG.fill_uniform(self.pointWeight.data, -1.0 / math.sqrt(self.nin * self.mult), 1.0 / math.sqrt(self.nin * self.mult))
        if "bias" in self._parameters:
            self.bias.data.zero_()
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        x = input.permute(0, 3, 1, 2) if self.data_format == "NHWC" else input
        if self.pw == -1 or self.ph == -1:
            pt, pb, _ = same_padding(x.shape[2], self.kh, self.sh)
            pl, pr, _ = same_padding(x.shape[3], self.kw, self.sw)
            x = F.pad(x, (pl, pr, pt, pb))
            pad = (0, 0)
        else:
            pad = (self.ph, self.pw)
        y = F.conv2d(x, self.depthWeight.to(x.dtype), None, (self.sh, self.sw), pad, 1, self.nin)
        y = F.conv2d(y, self.pointWeight.to(x.dtype), self._parameters.get("bias"))
        return y.permute(0, 2, 3, 1) if self.data_format == "NHWC" else y


class SpatialConvolutionMap(Module):
    """Convolution with an explicit (input plane → output plane) connection table (S/nn/SpatialConvolutionMap.scala;
    tables built by ``full``, ``oneToOne``, ``random``). One 2-D kernel per connection."""

    def __init__(self, conn_table, kw, kh, dw=1, dh=1, pad_w=0, pad_h=0, wRegularizer=None, bRegularizer=None):
        super().__init__()
        ct = torch.as_tensor(conn_table).long().reshape(-1, 2)
        self.conn_table = ct
        self.nin, self.nout = int(ct[:, 0].max()), int(ct[:, 1].max())
        self.kw, self.kh, self.dw, self.dh, self.pad_w, self.pad_h = kw, kh, dw, dh, pad_w, pad_h
        self.register_parameter("weight", torch.zeros(ct.shape[0], kh, kw))
        self.register_parameter("bias", torch.zeros(self.nout))
        self.reset()

    @staticmethod
    def full(nin, nout):
        return torch.tensor([[i, o] for o in range(1, nout + 1) for i in range(1, nin + 1)])

    @staticmethod
    def oneToOne(nfeat):
        return torch.tensor([[i, i] for i in range(1, nfeat + 1)])

    @staticmethod
    def random(nin, nout, nto):
        rows = []
        for o in range(1, nout + 1):
            perm = RNG.randperm(nin)[:nto]
            rows.extend([[int(i) + 1, o] for i in perm])
        return torch.tensor(rows)

    def reset(self):
        ninp = torch.zeros(self.nout)
        for _, o in self.conn_table.tolist():
            ninp[o - 1] += 1
        for k, (_, o) in enumerate(self.conn_table.tolist()):
            stdv = 1.0 / math.sqrt(self.kw * self.kh * float(ninp[o - 1]))
            RNG.fill_uniform(self.weight.data[k], -stdv, stdv)
        for o in range(self.nout):
            stdv = 1.0 / math.sqrt(self.kw * self.kh * max(float(ninp[o]), 1.0))
            RNG.fill_uniform(self.bias.data[o:o + 1], -stdv, stdv)
        self.zeroGradParameters()
        return self

    def updateOutput(self, input):
        x = input if input.dim() == 4 else input.unsqueeze(0)
        dense = torch.zeros(self.nout, self.nin, self.kh, self.kw, dtype=x.dtype, device=x.device)
        i = self.conn_table[:, 0].to(x.device) - 1
        o = self.conn_table[:, 1].to(x.device) - 1
        dense = dense.index_put((o, i), self.weight.to(x.dtype), accumulate=True)
        y = F.conv2d(x, dense, self.bias.to(x.dtype), (self.dh, self.dw), (self.pad_h, self.pad_w))
        return y if input.dim() == 4 else y.squeeze(0)


class TemporalConvolution(Module):
    """1-D convolution over [B, T, inputFrameSize] (S/nn/TemporalConvolution.scala); weight [out, kW*in]."""

    def __init__(self, input_frame_size, output_frame_size, kernel_w, stride_w=1, propagate_back=True,
                 weight_regularizer=None, bias_regularizer=None, init_weight=None, init_bias=None,
                 init_grad_weight=None, init_grad_bias=None):
        super().__init__()
        self.inp, self.out, self.kw, self.dw = input_frame_size, output_frame_size, kernel_w, stride_w
        self.propagate_back = propagate_back
        self.wRegularizer, self.bRegularizer = weight_regularizer, bias_regularizer
        self.register_parameter("weight", torch.zeros(output_frame_size, kernel_w * input_frame_size))
        self.register_parameter("bias", torch.zeros(output_frame_size))
        self.reset()

    def reset(self):
        _init_conv(self, self.weight, self.bias, self.kw * self.inp, VariableFormat.OUT_IN)
        return self

    def updateOutput(self, input):
        x = input if input.dim() == 3 else input.unsqueeze(0)
        if not self.propagate_back:
            x = x.detach()
        w = self.weight.reshape(self.out, self.kw, self.inp).permute(0, 2, 1).to(x.dtype)
        y = F.conv1d(x.transpose(1, 2), w, self.bias.to(x.dtype), self.dw).transpose(1, 2)
        return y if input.dim() == 3 else y.squeeze(0)


class VolumetricConvolution(Module):
    def __init__(self, n_input_plane, n_output_plane, k_t, k_w, k_h, d_t=1, d_w=1, d_h=1, pad_t=0, pad_w=0, pad_h=0,
                 with_bias=True, wRegularizer=None, bRegularizer=None):
        super().__init__()
        self.nin, self.nout = n_input_plane, n_output_plane
        self.k, self.d, self.p = (k_t, k_h, k_w), (d_t, d_h, d_w), (pad_t, pad_h, pad_w)
        self.register_parameter("weight", torch.zeros(n_output_plane, n_input_plane, k_t, k_h, k_w))
        if with_bias:
            self.register_parameter("bias", torch.zeros(n_output_plane))
        self.reset()

    def reset(self):
        _init_conv(self, self.weight, self._parameters.get("bias"), self.nin * self.k[0] * self.k[1] * self.k[2],
                   VariableFormat.OUT_IN_KT_KH_KW)
        return self

    def updateOutput(self, input):
        x = input if input.dim() == 5 else input.unsqueeze(0)
        if -1 in self.p:
            pads = []
            for size, k, s in zip(x.shape[2:], self.k, self.d):
                a, b, _ = same_padding(size, k, s)
                pads.append((a, b))
            x = F.pad(x, (pads[2][0], pads[2][1], pads[1][0], pads[1][1], pads[0][0], pads[0][1]))
            pad = 0
        else:
            pad = self.p
        b = self._parameters.get("bias")
        y = F.conv3d(x, self.weight.to(x.dtype), None if b is None else b.to(x.dtype), self.d, pad)
        return y if input.dim() == 5 else y.squeeze(0)


class VolumetricFullConvolution(Module):
    def __init__(self, n_input_plane, n_output_plane, kt, kw, kh, dt=1, dw=1, dh=1, pad_t=0, pad_w=0, pad_h=0,
                 adj_t=0, adj_w=0, adj_h=0, n_group=1, no_bias=False, wRegularizer=None, bRegularizer=None):
        super().__init__()
        self.nin, self.nout, self.n_group = n_input_plane, n_output_plane, n_group
        self.k, self.d, self.p, self.adj = (kt, kh, kw), (dt, dh, dw), (pad_t, pad_h, pad_w), (adj_t, adj_h, adj_w)
        self.register_parameter("weight", torch.zeros(n_input_plane, n_output_plane // n_group, kt, kh, kw))
        if not no_bias:
            self.register_parameter("bias", torch.zeros(n_output_plane))
        self.reset()

    def reset(self):
        _init_conv(self, self.weight, self._parameters.get("bias"),
                   (self.nout // self.n_group) * self.k[0] * self.k[1] * self.k[2], VariableFormat.DEFAULT)
        return self

    def updateOutput(self, input):
        x = input if input.dim() == 5 else input.unsqueeze(0)
        b = self._parameters.get("bias")
        y = F.conv_transpose3d(x, self.weight.to(x.dtype), None if b is None else b.to(x.dtype), self.d, self.p,
                               self.adj, self.n_group)
        return y if input.dim() == 5 else y.squeeze(0)


class LocallyConnected2D(Module):
    """Convolution with untied weights: one filter bank per output position (S/nn/LocallyConnected2D.scala)."""

    def __init__(self, n_input_plane, input_width, input_height, n_output_plane, kernel_w, kernel_h, stride_w=1,
                 stride_h=1, pad_w=0, pad_h=0, propagate_back=True, wRegularizer=None, bRegularizer=None,
                 init_weight=None, init_bias=None, init_grad_weight=None, init_grad_bias=None, with_bias=True,
                 data_format="NCHW"):
        super().__init__()
        self.nin, self.nout, self.kw, self.kh = n_input_plane, n_output_plane, kernel_w, kernel_h
        self.sw, self.sh, self.pw, self.ph = stride_w, stride_h, pad_w, pad_h
        self.data_format = data_format.upper()
        self.oh = conv_out_size(input_height, kernel_h, stride_h, max(pad_h, 0))
        self.ow = conv_out_size(input_width, kernel_w, stride_w, max(pad_w, 0))
        self.register_parameter("weight", torch.zeros(self.oh * self.ow, n_output_plane, n_input_plane * kernel_h * kernel_w))
        if with_bias:
            self.register_parameter("bias", torch.zeros(n_output_plane, self.oh, self.ow))
        self.reset()

    def reset(self):
        _init_conv(self, self.weight, self._parameters.get("bias"), self.nin * self.kh * self.kw, VariableFormat.DEFAULT)
        return self

    def updateOutput(self, input):
        x = input.permute(0, 3, 1, 2) if self.data_format == "NHWC" else input
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        cols = F.unfold(x, (self.kh, self.kw), 1, (self.ph, self.pw), (self.sh, self.sw))   # [B, C*kh*kw, L]
        y = torch.einsum("lok,bkl->bol", self.weight.to(x.dtype), cols).reshape(x.shape[0], self.nout, self.oh, self.ow)
        if "bias" in self._parameters:
            y = y + self.bias.to(x.dtype)
        if squeeze:
            y = y.squeeze(0)
        return y.permute(0, 2, 3, 1) if self.data_format == "NHWC" else y


class LocallyConnected1D(Module):
    """Untied-weight temporal convolution over [B, T, in] (S/nn/LocallyConnected1D.scala)."""

    def __init__(self, n_input_frame, input_frame_size, output_frame_size, kernel_w, stride_w=1, propagate_back=True,
                 weight_regularizer=None, bias_regularizer=None, init_weight=None, init_bias=None,
                 init_grad_weight=None, init_grad_bias=None):
        super().__init__()
        self.T, self.inp, self.out, self.kw, self.dw = n_input_frame, input_frame_size, output_frame_size, kernel_w, stride_w
        self.n_out_frame = (n_input_frame - kernel_w) // stride_w + 1
        self.register_parameter("weight", torch.zeros(self.n_out_frame, output_frame_size, kernel_w * input_frame_size))
        self.register_parameter("bias", torch.zeros(self.n_out_frame, output_frame_size))
        self.reset()

    def reset(self):
        _init_conv(self, self.weight, self.bias, self.kw * self.inp, VariableFormat.DEFAULT)
        return self

    def updateOutput(self, input):
        x = input if input.dim() == 3 else input.unsqueeze(0)
        win = x.unfold(1, self.kw, self.dw)                        # [B, L, in, kw]
        win = win.permute(0, 1, 3, 2).reshape(x.shape[0], self.n_out_frame, self.kw * self.inp)
        y = torch.einsum("lok,blk->blo", self.weight.to(x.dtype), win) + self.bias.to(x.dtype)
        return y if input.dim() == 3 else y.squeeze(0)


# ---------------------------------------------------------------- pooling (1-D / 3-D)
class TemporalMaxPooling(Module):
    def __init__(self, k_w, d_w=None):
        super().__init__()
        self.kw, self.dw = k_w, d_w or k_w

    def updateOutput(self, input):
        x = input if input.dim() == 3 else input.unsqueeze(0)
        y = F.max_pool1d(x.transpose(1, 2), self.kw, self.dw).transpose(1, 2)
        return y if input.dim() == 3 else y.squeeze(0)


class VolumetricMaxPooling(Module):
    def __init__(self, k_t, k_w, k_h, d_t=None, d_w=None, d_h=None, pad_t=0, pad_w=0, pad_h=0):
        super().__init__()
        self.k = (k_t, k_h, k_w)
        self.d = (d_t or k_t, d_h or k_h, d_w or k_w)
        self.p = (pad_t, pad_h, pad_w)
        self.ceil_mode = False

    def ceil(self):
        self.ceil_mode = True
        return self

    def floor(self):
        self.ceil_mode = False
        return self

    def updateOutput(self, input):
        x = input if input.dim() == 5 else input.unsqueeze(0)
        y = F.max_pool3d(x, self.k, self.d, self.p, ceil_mode=self.ceil_mode)
        return y if input.dim() == 5 else y.squeeze(0)


class VolumetricAveragePooling(Module):
    def __init__(self, k_t, k_w, k_h, d_t=None, d_w=None, d_h=None, pad_t=0, pad_w=0, pad_h=0,
                 count_include_pad=True, ceil_mode=False):
        super().__init__()
        self.k = (k_t, k_h, k_w)
        self.d = (d_t or k_t, d_h or k_h, d_w or k_w)
        self.p = (pad_t, pad_h, pad_w)
        self.count_include_pad, self.ceil_mode = count_include_pad, ceil_mode

    def ceil(self):
        self.ceil_mode = True
        return self

    def updateOutput(self, input):
        x = input if input.dim() == 5 else input.unsqueeze(0)
        y = F.avg_pool3d(x, self.k, self.d, self.p, self.ceil_mode, self.count_include_pad)
        return y if input.dim() == 5 else y.squeeze(0)


# ---------------------------------------------------------------- resizing
class ResizeBilinear(Module):
    def __init__(self, output_height, output_width, align_corner=False, data_format="NCHW"):
        super().__init__()
        self.oh, self.ow, self.align_corner, self.data_format = output_height, output_width, align_corner, data_format.upper()

    def updateOutput(self, input):
        x = input.permute(0, 3, 1, 2) if self.data_format == "NHWC" else input
        y = F.interpolate(x.float(), (self.oh, self.ow), mode="bilinear", align_corners=self.align_corner).to(x.dtype)
        return y.permute(0, 2, 3, 1) if self.data_format == "NHWC" else y


class UpSampling1D(Module):
    def __init__(self, length: int):
        super().__init__()
        self.length = length

    def updateOutput(self, input):
        return input.repeat_interleave(self.length, dim=1)


class UpSampling2D(Module):
    def __init__(self, size, data_format="NCHW"):
        super().__init__()
        self.size, self.data_format = list(size), data_format.upper()

    def updateOutput(self, input):
        hd, wd = (2, 3) if self.data_format == "NCHW" else (1, 2)
        return input.repeat_interleave(self.size[0], dim=hd).repeat_interleave(self.size[1], dim=wd)


class UpSampling3D(Module):
    def __init__(self, size):
        super().__init__()
        self.size = list(size)

    def updateOutput(self, input):
        y = input
        for i, s in enumerate(self.size):
            y = y.repeat_interleave(s, dim=2 + i)
        return y


# ---------------------------------------------------------------- ROI ops
class RoiPooling(Module):
    """Max-pool each ROI to pooled_h x pooled_w. input = table(features [N,C,H,W], rois [R,5] (batch idx (0-based as in
    Caffe), x1, y1, x2, y2)) (S/nn/RoiPooling.scala)."""

    def __init__(self, pooled_w, pooled_h, spatial_scale):
        super().__init__()
        self.pw, self.ph, self.scale = pooled_w, pooled_h, spatial_scale

    def updateOutput(self, input):
        feat, rois = to_table(input).to_list()
        R = rois.shape[0]
        C, H, W = feat.shape[1:]
        out = feat.new_zeros(R, C, self.ph, self.pw)
        roi_rows = rois.detach().cpu().tolist()          # box coordinates are indices here, not differentiable inputs
        for r in range(R):
            b = int(roi_rows[r][0])
            x1, y1, x2, y2 = [int(round(float(v) * self.scale)) for v in roi_rows[r][1:5]]
            rh, rw = max(y2 - y1 + 1, 1), max(x2 - x1 + 1, 1)
            bh, bw = rh / self.ph, rw / self.pw
            for i in range(self.ph):
                hs = min(max(int(math.floor(i * bh)) + y1, 0), H)
                he = min(max(int(math.ceil((i + 1) * bh)) + y1, 0), H)
                for j in range(self.pw):
                    ws = min(max(int(math.floor(j * bw)) + x1, 0), W)
                    we = min(max(int(math.ceil((j + 1) * bw)) + x1, 0), W)
                    if he > hs and we > ws:
                        out[r, :, i, j] = feat[b, :, hs:he, ws:we].amax(dim=(1, 2))
        return out


def _bilinear_gather(feat, y, x):
    """feat [C,H,W]; y, x [...] float coordinates → [C, ...] bilinear samples (zero outside [-1, size])."""
    C, H, W = feat.shape
    valid = (y >= -1.0) & (y <= H) & (x >= -1.0) & (x <= W)
    y = y.clamp(min=0.0)
    x = x.clamp(min=0.0)
    y0 = y.floor().clamp(max=H - 1)
    x0 = x.floor().clamp(max=W - 1)
    y1 = (y0 + 1).clamp(max=H - 1)
    x1 = (x0 + 1).clamp(max=W - 1)
    y = torch.where(y0 >= H - 1, y0, y)
    x = torch.where(x0 >= W - 1, x0, x)
    ly, lx = y - y0, x - x0
    hy, hx = 1 - ly, 1 - lx
    y0l, x0l, y1l, x1l = y0.long(), x0.long(), y1.long(), x1.long()
    v = (feat[:, y0l, x0l] * (hy * hx) + feat[:, y0l, x1l] * (hy * lx) +
         feat[:, y1l, x0l] * (ly * hx) + feat[:, y1l, x1l] * (ly * lx))
    return v * valid.to(v.dtype)


class RoiAlign(Module):
    """Average (or max) of bilinear samples per ROI bin (Mask R-CNN), rois [R,4] (x1,y1,x2,y2) for a single image or
    [R,5] with batch index (S/nn/RoiAlign.scala)."""

    def __init__(self, spatial_scale, sampling_ratio, pooled_h, pooled_w, mode="avg", aligned=True):
        super().__init__()
        self.scale, self.ratio, self.ph, self.pw, self.mode, self.aligned = (spatial_scale, sampling_ratio, pooled_h,
                                                                          pooled_w, mode, aligned)

    def updateOutput(self, input):
        feat, rois = to_table(input).to_list()
        if feat.dim() == 3:
            feat = feat.unsqueeze(0)
        R = rois.shape[0]
        C = feat.shape[1]
        out = []
        off = 0.5 if self.aligned else 0.0
        for r in range(R):
            if rois.shape[1] == 5:
                b, box = int(rois[r, 0]), rois[r, 1:5]
            else:
                b, box = 0, rois[r, :4]
            x1, y1, x2, y2 = [box[i].float() * self.scale - off for i in range(4)]
            rw, rh = x2 - x1, y2 - y1
            if not self.aligned:
                rw, rh = rw.clamp(min=1.0), rh.clamp(min=1.0)
            bh, bw = rh / self.ph, rw / self.pw
            gh = self.ratio if self.ratio > 0 else int(math.ceil(float(rh) / self.ph))
            gw = self.ratio if self.ratio > 0 else int(math.ceil(float(rw) / self.pw))
            gh, gw = max(gh, 1), max(gw, 1)
            iy = (torch.arange(self.ph, device=feat.device).float().repeat_interleave(gh) * bh + y1 +
                  (torch.arange(gh, device=feat.device).float().repeat(self.ph) + 0.5) * bh / gh)
            ix = (torch.arange(self.pw, device=feat.device).float().repeat_interleave(gw) * bw + x1 +
                  (torch.arange(gw, device=feat.device).float().repeat(self.pw) + 0.5) * bw / gw)
            yy, xx = torch.meshgrid(iy, ix, indexing="ij")
            v = _bilinear_gather(feat[b].float(), yy, xx).reshape(C, self.ph, gh, self.pw, gw)
            out.append(v.amax(dim=(2, 4)) if self.mode == "max" else v.mean(dim=(2, 4)))
        return torch.stack(out).to(feat.dtype) if out else feat.new_zeros(0, C, self.ph, self.pw)


class Pooler(Module):
    """FPN ROI pooler: route each ROI to a pyramid level by its scale, RoiAlign there, and merge in ROI order
    (S/nn/Pooler.scala). input = table(table(features per level), rois [R,4])."""

    def __init__(self, resolution, scales, sampling_ratio):
        super().__init__()
        self.resolution, self.scales, self.sampling_ratio = resolution, list(scales), sampling_ratio
        self.poolers = [RoiAlign(s, sampling_ratio, resolution, resolution) for s in self.scales]
        self.lvl_min = -int(round(math.log2(self.scales[0])))
        self.lvl_max = -int(round(math.log2(self.scales[-1])))

    def updateOutput(self, input):
        feats, rois = to_table(input).to_list()
        feats = to_table(feats).to_list()
        if isinstance(rois, Table):
            rois = torch.cat(rois.to_list(), dim=0)
        area = ((rois[:, 2] - rois[:, 0] + 1) * (rois[:, 3] - rois[:, 1] + 1)).float().clamp_min(1e-6)
        lvl = torch.floor(4 + torch.log2(area.sqrt() / 224.0 + 1e-6)).clamp(self.lvl_min, self.lvl_max).long() - self.lvl_min
        C = feats[0].shape[1]
        out = feats[0].new_zeros(rois.shape[0], C, self.resolution, self.resolution)
        for l, (f, p) in enumerate(zip(feats, self.poolers)):
            idx = torch.nonzero(lvl == l).reshape(-1)
            if idx.numel():
                out[idx] = p._call(Table(f, rois[idx]))
        return out
