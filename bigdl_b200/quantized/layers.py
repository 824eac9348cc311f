"""Quantised inference layers. The reference quantises Linear / SpatialConvolution / SpatialDilatedConvolution to int8
for BigQuant's mixed-precision GEMM (S/nn/quantized/{Linear,SpatialConvolution,SpatialDilatedConvolution}.scala) and, on
the MKL-DNN backend, runs int8 convolutions with per-tensor input scales and per-channel weight scales
(S/nn/mkldnn/SpatialConvolution.scala:237-247, S/nn/MklInt8Convertible.scala). On B200 the narrow tensor-core type is
fp8 e4m3 (INT8 MMA is a legacy path), so weights are stored as e4m3 with one abs-max per output channel and activations
are quantised per tensor — dynamically (abs-max of the batch) or statically after ``Quantizer.calibrate``.

GPU path: quantise kernel → tcgen05 ``kind::f8f6f4`` GEMM / TMA-im2col implicit GEMM with the dequantisation
(``act_scale * w_scale[n]``), bias and ReLU in the epilogue; a producer may requantise its output in the epilogue so a
chain of quantised layers never materialises bf16. CPU path: identical arithmetic emulated with torch.float8_e4m3fn.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from ..nn.layers_core import conv_out_size, native, same_padding
from ..nn.module import Module
from ..ops import quant as Q


class QuantizedModule(Module):
    """Inference-only module (S/nn/quantized/QuantizedModule.scala): backward is an error."""

    def backward(self, input, gradOutput):
        raise ValueError("Doesn't updateGradInput for quantized model")

    def updateGradInput(self, input, gradOutput):
        raise ValueError("Doesn't updateGradInput for quantized model")

    def accGradParameters(self, input, gradOutput):
        raise ValueError("Doesn't accGradParameters for quantized model")

    # ---- activation scale bookkeeping (MklInt8Convertible.calcScales / getInputScales / setInputScales)
    def _init_quant_state(self):
        self.register_buffer("input_absmax", torch.zeros(1))
        self.static_scale = False        # use input_absmax instead of measuring each batch
        self.calibrating = False
        self._colscale = None
        self._emit_q = None              # consumer layer whose static scale the epilogue requantises for

    def _act_absmax(self, x: torch.Tensor) -> torch.Tensor:
        from ..ops.elementwise import absmax

        if self.calibrating:
            absmax(x.contiguous(), self._buffers["input_absmax"])
            self._colscale = None
        if self.static_scale and not self.calibrating:
            return self._buffers["input_absmax"]
        return absmax(x.contiguous())

    def _col_scale(self, am: torch.Tensor) -> torch.Tensor:
        if self.static_scale and not self.calibrating:
            if self._colscale is None or self._colscale.device != am.device:
                self._colscale = Q.colscale(am, self._buffers["weight_absmax"])
            return self._colscale
        return Q.colscale(am, self._buffers["weight_absmax"])

    def _out_qscale(self):
        c = self._emit_q
        if c is None or not c.static_scale or c.calibrating or getattr(c, "block_scale", False):
            return None
        qs = c.__dict__.get("_in_qscale")
        if qs is None or qs.device != c._buffers["input_absmax"].device:
            am = c._buffers["input_absmax"]
            qs = torch.where(am > 0, Q.E4M3_MAX / am, torch.zeros_like(am))
            c.__dict__["_in_qscale"] = qs
        return qs

    def getInputScales(self):
        return [[float(self._buffers["input_absmax"])]]

    def setInputScales(self, scales):
        self._buffers["input_absmax"].fill_(float(scales[0][0]))
        self.static_scale, self._colscale = True, None
        self.__dict__.pop("_in_qscale", None)
        return self

    def getWeightScales(self):
        return [self._buffers["weight_absmax"].tolist()]

    def _on_device_change(self, device):
        self._colscale = None
        self.__dict__.pop("_in_qscale", None)


def _quantize_weight(w2d: torch.Tensor):
    am = Q.row_absmax(w2d.contiguous())
    return Q.quantize_e4m3(w2d.contiguous(), am, per_row=True, pad_to=16), am


class Linear(QuantizedModule):
    """fp8 Linear (S/nn/quantized/Linear.scala:30-140)."""

    def __init__(self, input_size: int, output_size: int, with_bias: bool = True):
        super().__init__()
        self.input_size, self.output_size, self.with_bias = input_size, output_size, with_bias
        kp = (input_size + 15) // 16 * 16
        self.register_buffer("weight_q", torch.zeros(output_size, kp, dtype=torch.uint8))
        self.register_buffer("weight_absmax", torch.zeros(output_size))
        self.register_buffer("bias", torch.zeros(output_size))
        self._fuse_relu = False
        self._init_quant_state()

    def enable_block_scale(self, w: torch.Tensor | None = None):
        """Switch to MX block scaling: e4m3 with one power-of-two scale per 32 consecutive inputs, for the weights (stored)
        and the activations (quantised per call by ONE kernel: no abs-max pass, no calibration), multiplied by
        ``tcgen05.mma.kind::mxf8f6f4.block_scale``. The scale follows the data block by block like BigQuant's per-window
        ("local") thresholds do (core/bigquant/native/ops/shuffle/shuffle_im2col.h)."""
        if w is None:
            w = self.__dict__.pop("_w_float", None)      # the float weights, when the Quantizer is still converting
        if w is None:
            w = self.dequantized_weight()
        wq, sf = Q.quantize_mxfp8(w.float().contiguous())
        self.static_scale = False
        self._buffers["weight_q"], self._buffers["weight_sf"] = wq, sf
        self.block_scale = True
        return self

    def _block_scale_forward(self, x):
        lead = x.shape[:-1]
        K = self.input_size
        x2 = x.detach().reshape(-1, K)
        if native(x):
            if x2.dtype != torch.bfloat16:
                x2 = x2.to(torch.bfloat16)
            xq, sfa = Q.quantize_mxfp8(x2.contiguous())
            y = Q.gemm_mxfp8(xq, sfa, self._buffers["weight_q"], self._buffers["weight_sf"], K,
                             bias=self._buffers["bias"], relu=self._fuse_relu)
        else:
            xq, sfa = Q.quantize_mxfp8(x2.float())
            xd = Q.dequantize_mxfp8(xq, sfa, K)
            wd = Q.dequantize_mxfp8(self._buffers["weight_q"].cpu(), self._buffers["weight_sf"].cpu(), K)
            y = F.linear(xd, wd, self._buffers["bias"].cpu())
            if self._fuse_relu:
                y = F.relu(y)
        return y.reshape(*lead, self.output_size)

    @classmethod
    def from_float(cls, m) -> "Linear":
        q = cls(m.input_size, m.output_size, m.with_bias)
        w = m.weight.detach().float()
        wq, am = _quantize_weight(w)
        q._buffers["weight_q"], q._buffers["weight_absmax"] = wq, am
        q.__dict__["_w_float"] = w               # dropped by Quantizer.quantize once the scaling mode is settled
        if m.with_bias:
            q._buffers["bias"] = m.bias.detach().float().clone()
        else:
            q._buffers["bias"] = torch.zeros(m.output_size, device=w.device)
        q._fuse_relu = getattr(m, "_fuse_relu", False)
        if m.hasName():
            q.setName(m.getName())
        return q

    def dequantized_weight(self) -> torch.Tensor:
        if getattr(self, "block_scale", False):
            return Q.dequantize_mxfp8(self._buffers["weight_q"], self._buffers["weight_sf"], self.input_size)
        return Q.dequantize_e4m3(self._buffers["weight_q"], self._buffers["weight_absmax"], per_row=True)[
            :, :self.input_size]

    def updateOutput(self, input):
        x = input
        if getattr(self, "block_scale", False) and x.dtype != torch.uint8:
            with torch.no_grad():
                return self._block_scale_forward(x)
        lead = x.shape[:-1]
        K = self.input_size
        if native(x):
            with torch.no_grad():
                if x.dtype == torch.uint8:            # producer already requantised for us
                    xq, am = x.reshape(-1, x.shape[-1]), self._buffers["input_absmax"]
                else:
                    x2 = x.detach().reshape(-1, K)
                    if x2.dtype != torch.bfloat16:
                        x2 = x2.to(torch.bfloat16)
                    x2 = x2.contiguous()
                    am = self._act_absmax(x2)
                    xq = Q.quantize_e4m3(x2, am)
                qs = self._out_qscale()
                y = Q.gemm_fp8(xq, self._buffers["weight_q"], self._col_scale(am), k=K, bias=self._buffers["bias"],
                               relu=self._fuse_relu, out_dtype=torch.uint8 if qs is not None else torch.bfloat16,
                               out_qscale=qs)
            return y.reshape(*lead, self.output_size)
        with torch.no_grad():
            x2 = x.detach().float().reshape(-1, K)
            am = self._act_absmax(x2)
            xd = Q.dequantize_e4m3(Q.quantize_e4m3(x2, am, pad_to=1), am)
            y = F.linear(xd, self.dequantized_weight().to(xd.device), self._buffers["bias"])
            if self._fuse_relu:
                y = F.relu(y)
        return y.reshape(*lead, self.output_size)


class SpatialConvolution(QuantizedModule):
    """fp8 convolution (S/nn/quantized/SpatialConvolution.scala:35-260; dilation from SpatialDilatedConvolution.scala).
    Weights are [O, kH*kW*Cp] e4m3 (tap-major, channels-last, C padded to 16) with one abs-max per output channel."""

    def __init__(self, n_input_plane: int, n_output_plane: int, kernel_w: int, kernel_h: int, stride_w: int = 1,
                 stride_h: int = 1, pad_w: int = 0, pad_h: int = 0, n_group: int = 1, with_bias: bool = True,
                 data_format: str = "NCHW", dilation_w: int = 1, dilation_h: int = 1):
        super().__init__()
        self.n_input_plane, self.n_output_plane = n_input_plane, n_output_plane
        self.kernel_w, self.kernel_h, self.stride_w, self.stride_h = kernel_w, kernel_h, stride_w, stride_h
        self.pad_w, self.pad_h, self.n_group, self.with_bias = pad_w, pad_h, n_group, with_bias
        self.dilation_w, self.dilation_h, self.data_format = dilation_w, dilation_h, data_format.upper()
        cg = n_input_plane // n_group
        self.cp = (cg + 15) // 16 * 16
        self.register_buffer("weight_q", torch.zeros(n_output_plane, kernel_h * kernel_w * self.cp, dtype=torch.uint8))
        self.register_buffer("weight_absmax", torch.zeros(n_output_plane))
        self.register_buffer("bias", torch.zeros(n_output_plane))
        self._fuse_relu = False
        self._init_quant_state()

    @classmethod
    def from_float(cls, m) -> "SpatialConvolution":
        q = cls(m.n_input_plane, m.n_output_plane, m.kernel_w, m.kernel_h, m.stride_w, m.stride_h, m.pad_w, m.pad_h,
                m.n_group, m.with_bias, m.data_format, m.dilation_w, m.dilation_h)
        w = m.weight.detach().float()                           # [O, C/g, kh, kw]
        O, cg, kh, kw = w.shape
        w_ohwi = torch.zeros(O, kh, kw, q.cp, device=w.device)
        w_ohwi[..., :cg] = w.permute(0, 2, 3, 1)
        am = Q.row_absmax(w_ohwi.reshape(O, -1))
        q._buffers["weight_q"] = Q.quantize_e4m3(w_ohwi.reshape(O, -1), am, per_row=True, pad_to=16)
        q._buffers["weight_absmax"] = am
        q._buffers["bias"] = m.bias.detach().float().clone() if m.with_bias else torch.zeros(O, device=w.device)
        q._fuse_relu = getattr(m, "_fuse_relu", False)
        if cg < 16 and m.n_group == 1:
            # An RGB stem would waste 7/8 of every 128-byte fp8 k-block on channel padding, and first layers are the
            # usual high-precision exception in int8/fp8 deployments: keep a bf16 copy and run the bf16 kernel on GPU.
            q.register_buffer("weight_hp", w.to(torch.bfloat16).contiguous(memory_format=torch.channels_last))
        if m.hasName():
            q.setName(m.getName())
        return q

    def dequantized_weight(self) -> torch.Tensor:
        """[O, C/g, kH, kW] fp32."""
        O, cg = self.n_output_plane, self.n_input_plane // self.n_group
        w = Q.dequantize_e4m3(self._buffers["weight_q"], self._buffers["weight_absmax"], per_row=True)
        return w.reshape(O, self.kernel_h, self.kernel_w, self.cp)[..., :cg].permute(0, 3, 1, 2).contiguous()

    def _padding(self, H, W):
        if self.pad_w == -1 or self.pad_h == -1:
            pt, pb, _ = same_padding(H, self.dilation_h * (self.kernel_h - 1) + 1, self.stride_h)
            pl, pr, _ = same_padding(W, self.dilation_w * (self.kernel_w - 1) + 1, self.stride_w)
            return pt, pl, pb, pr
        return self.pad_h, self.pad_w, self.pad_h, self.pad_w

    def _out_hw(self, H, W, pads):
        pt, pl, pb, pr = pads
        Ho = (H + pt + pb - (self.dilation_h * (self.kernel_h - 1) + 1)) // self.stride_h + 1
        Wo = (W + pl + pr - (self.dilation_w * (self.kernel_w - 1) + 1)) // self.stride_w + 1
        return Ho, Wo

    def updateOutput(self, input):
        x = input
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        if self.data_format == "NHWC":
            x = x.permute(0, 3, 1, 2)
        N, C, H, W = x.shape
        pads = self._padding(H, W)
        Ho, Wo = self._out_hw(H, W, pads)
        sym = pads[0] == pads[2] and pads[1] == pads[3]
        with torch.no_grad():
            if native(x) and "weight_hp" in self._buffers and x.dtype != torch.uint8:
                from ..ops.functions import ConvFn

                if self.calibrating:
                    self._act_absmax(x.detach())
                geom = (self.kernel_h, self.kernel_w, self.stride_h, self.stride_w, *pads, self.dilation_h,
                        self.dilation_w, 1)
                w16 = self._buffers["weight_hp"]
                y = ConvFn.apply(x.detach(), w16, self._buffers["bias"], w16, geom, self._fuse_relu, 1.0, 1.0)
            elif native(x) and self.n_group == 1 and sym:
                y = self._forward_fp8(x, N, H, W, Ho, Wo, pads)
            else:
                y = self._forward_emulated(x, pads)
        if self.data_format == "NHWC":
            y = y.permute(0, 2, 3, 1)
        return y.squeeze(0) if squeeze else y

    def _forward_fp8(self, x, N, H, W, Ho, Wo, pads):
        cp = self.cp
        if x.dtype == torch.uint8:                # [N, Cp, H, W] logical, NHWC physical, requantised by the producer
            xq = x.permute(0, 2, 3, 1)
            am = self._buffers["input_absmax"]
        else:
            from ..ops.functions import nhwc_pitched, rows2d

            static = self.static_scale and not self.calibrating
            am = self._buffers["input_absmax"] if static else None
            hit = getattr(x, "_bdl_q", None) if static else None
            if hit is not None and hit[0] == am.data_ptr():
                xq = hit[1]                   # a sibling branch sharing our scale already quantised this tensor
            else:
                xp, ld = nhwc_pitched(x.detach())
                x2 = rows2d(xp, ld)
                if not static:
                    am = self._act_absmax(x2 if ld == x2.shape[1] else x2.contiguous())
                xq = Q.quantize_e4m3(x2, am, pad_to=16).view(N, H, W, cp)
                if static:
                    x._bdl_q = (am.data_ptr(), xq)
        geom = (N, H, W, cp, Ho, Wo, self.kernel_h, self.kernel_w, self.stride_h, self.stride_w, pads[0], pads[1],
                self.dilation_h, self.dilation_w)
        qs = self._out_qscale()
        O = self.n_output_plane
        emit = qs is not None and O % 16 == 0
        pointwise = self.kernel_h == 1 and self.kernel_w == 1 and self.stride_h == 1 and self.stride_w == 1 and \
            pads[0] == 0 and pads[1] == 0
        y2 = Q.gemm_fp8(xq.reshape(N * H * W, cp) if pointwise else xq, self._buffers["weight_q"],
                        self._col_scale(am), k=cp, bias=self._buffers["bias"], relu=self._fuse_relu,
                        out_dtype=torch.uint8 if emit else torch.bfloat16, out_qscale=qs if emit else None,
                        conv_geom=None if pointwise else geom)
        return y2.view(N, Ho, Wo, O).permute(0, 3, 1, 2)

    def _forward_emulated(self, x, pads):
        xf = x.detach().float()
        am = self._act_absmax(xf)
        xd = Q.dequantize_e4m3(Q.quantize_e4m3(xf.reshape(1, -1), am, pad_to=1), am).reshape(xf.shape)
        pt, pl, pb, pr = pads
        if (pt, pl) != (pb, pr):
            xd = F.pad(xd, (pl, pr, pt, pb))
            pt = pl = 0
        y = F.conv2d(xd, self.dequantized_weight().to(xd.device), self._buffers["bias"].to(xd.device),
                     (self.stride_h, self.stride_w), (pt, pl), (self.dilation_h, self.dilation_w), self.n_group)
        if self._fuse_relu:
            y = F.relu(y)
        return y.to(torch.bfloat16) if x.is_cuda else y


class SpatialDilatedConvolution(SpatialConvolution):
    """S/nn/quantized/SpatialDilatedConvolution.scala — the same kernel with dilation > 1."""
