"""Fused LSTM gate kernels (csrc/recurrent.cu) as an autograd Function; the cell state is carried in fp32."""
from __future__ import annotations

import torch
from torch.autograd import Function

from . import _lib


class _LstmGates(Function):
    @staticmethod
    def forward(ctx, g, c_prev):
        lib = _lib.cuda_lib()
        B, H4 = g.shape
        H = H4 // 4
        g = g if g.stride(1) == 1 else g.contiguous()
        cp = c_prev.float().contiguous()
        h = torch.empty((B, H), dtype=torch.bfloat16, device=g.device)
        c = torch.empty((B, H), dtype=torch.float32, device=g.device)
        gates = torch.empty((B, H4), dtype=torch.bfloat16, device=g.device)
        _lib.check(lib.bdl_lstm_gates_fwd(g.data_ptr(), cp.data_ptr(), h.data_ptr(), c.data_ptr(), gates.data_ptr(), B,
                                          H, g.stride(0), _lib.stream_ptr()), "bdl_lstm_gates_fwd")
        _lib.count_launch()
        ctx.save_for_backward(gates, c, cp)
        ctx.c_dtype = c_prev.dtype
        return h, c

    @staticmethod
    def backward(ctx, dh, dc):
        gates, c, cp = ctx.saved_tensors
        lib = _lib.cuda_lib()
        B, H = c.shape
        dg = torch.empty_like(gates)
        dcp = torch.empty_like(c)
        dhp = None if dh is None else dh.to(torch.bfloat16).contiguous()
        dcf = None if dc is None else dc.float().contiguous()
        _lib.check(lib.bdl_lstm_gates_bwd(_lib.ptr(dhp), _lib.ptr(dcf), gates.data_ptr(), c.data_ptr(), cp.data_ptr(),
                                          dg.data_ptr(), dcp.data_ptr(), B, H, _lib.stream_ptr()), "bdl_lstm_gates_bwd")
        _lib.count_launch()
        return dg, dcp.to(ctx.c_dtype)


def lstm_gates(g: torch.Tensor, c_prev: torch.Tensor):
    """(h [B,H] bf16, c [B,H] fp32) from pre-activations g [B,4H] (order: input, candidate, forget, output)."""
    return _LstmGates.apply(g, c_prev)


def _rows(t: torch.Tensor) -> torch.Tensor:
    """bf16 2-D tensor with unit inner stride (column slices of a wider matrix are used in place via their pitch)."""
    t = t if t.dtype == torch.bfloat16 else t.to(torch.bfloat16)
    return t if t.stride(1) == 1 else t.contiguous()


class _GruRZ(Function):
    """(z, h*r) from x_rz [B,2H], h.U_rz [B,2H] and h [B,H] — stage 1 of a GRU step (S/nn/GRU.scala:96-130)."""

    @staticmethod
    def forward(ctx, xrz, hg, h):
        lib = _lib.cuda_lib()
        xrz, hg, h = _rows(xrz), _rows(hg).contiguous(), _rows(h)
        B, H = h.shape
        rz = torch.empty((B, 2 * H), dtype=torch.bfloat16, device=h.device)
        hr = torch.empty((B, H), dtype=torch.bfloat16, device=h.device)
        _lib.check(lib.bdl_gru_rz_fwd(xrz.data_ptr(), hg.data_ptr(), h.data_ptr(), rz.data_ptr(), hr.data_ptr(), B, H,
                                      xrz.stride(0), h.stride(0), _lib.stream_ptr()), "bdl_gru_rz_fwd")
        _lib.count_launch()
        ctx.save_for_backward(rz, h)
        return rz[:, H:], hr

    @staticmethod
    def backward(ctx, d_z, d_hr):
        rz, h = ctx.saved_tensors
        lib = _lib.cuda_lib()
        B, H = h.shape
        d_pre = torch.empty_like(rz)
        d_h = torch.empty((B, H), dtype=torch.bfloat16, device=h.device)
        dz = None if d_z is None else d_z.to(torch.bfloat16).contiguous()
        dhr = None if d_hr is None else d_hr.to(torch.bfloat16).contiguous()
        _lib.check(lib.bdl_gru_rz_bwd(_lib.ptr(dhr), _lib.ptr(dz), rz.data_ptr(), h.data_ptr(), d_pre.data_ptr(),
                                      d_h.data_ptr(), B, H, h.stride(0), _lib.stream_ptr()), "bdl_gru_rz_bwd")
        _lib.count_launch()
        return d_pre, d_pre, d_h


class _GruOut(Function):
    """h' = (1 - z) * tanh(x_c + hr.U_c) + z * h — stage 2 of a GRU step."""

    @staticmethod
    def forward(ctx, xc, hc, z, h):
        lib = _lib.cuda_lib()
        xc, hc, z, h = _rows(xc), _rows(hc).contiguous(), _rows(z), _rows(h)
        B, H = h.shape
        hh = torch.empty((B, H), dtype=torch.bfloat16, device=h.device)
        out = torch.empty((B, H), dtype=torch.bfloat16, device=h.device)
        _lib.check(lib.bdl_gru_out_fwd(xc.data_ptr(), hc.data_ptr(), z.data_ptr(), h.data_ptr(), hh.data_ptr(),
                                       out.data_ptr(), B, H, xc.stride(0), z.stride(0), h.stride(0),
                                       _lib.stream_ptr()), "bdl_gru_out_fwd")
        _lib.count_launch()
        ctx.save_for_backward(hh, z, h)
        return out

    @staticmethod
    def backward(ctx, d_out):
        hh, z, h = ctx.saved_tensors
        lib = _lib.cuda_lib()
        B, H = h.shape
        d_pre = torch.empty((B, H), dtype=torch.bfloat16, device=h.device)
        d_z = torch.empty_like(d_pre)
        d_h = torch.empty_like(d_pre)
        g = d_out.to(torch.bfloat16).contiguous()
        _lib.check(lib.bdl_gru_out_bwd(g.data_ptr(), hh.data_ptr(), z.data_ptr(), h.data_ptr(), d_pre.data_ptr(),
                                       d_z.data_ptr(), d_h.data_ptr(), B, H, z.stride(0), h.stride(0),
                                       _lib.stream_ptr()), "bdl_gru_out_bwd")
        _lib.count_launch()
        return d_pre, d_pre, d_z, d_h


def gru_rz(xrz, hg, h):
    return _GruRZ.apply(xrz, hg, h)


def gru_out(xc, hc, z, h):
    return _GruOut.apply(xc, hc, z, h)
