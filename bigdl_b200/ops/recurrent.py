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
