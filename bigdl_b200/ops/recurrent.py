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


# ---------------------------------------------------------------------------------------------- whole-sequence LSTM
class LstmSeqFn(Function):
    """All T steps of an LSTM layer in one persistent kernel (csrc/lstm_seq.cu), forward and BPTT.

    ``xw`` [B,T,4H] bf16 is the hoisted input projection (x W_ih^T + b, already an autograd tensor: its gradient flows
    back into the i2g Linear, which turns it into dW_ih / db / dx with single GEMMs); ``whh`` is the fp32 recurrent
    weight parameter (its gradient is accumulated in place like every layer's), ``whh16`` its bf16 compute copy.
    Returns hs [B,T,H] bf16 and the final (h, c)."""

    @staticmethod
    def forward(ctx, xw, whh, whh16, h0, c0, reverse):
        lib = _lib.cuda_lib()
        B, T, H4 = xw.shape
        H = H4 // 4
        xw = xw if xw.stride(2) == 1 else xw.contiguous()
        dev = xw.device
        # rows of hs start 16-byte aligned (pitch padded to 8 elements, pad columns stay zero): every CTA re-reads all
        # of h_{t-1} each step, and PTB's 650 / 1500 would otherwise force 4- / 8-byte loads
        Hpad = (H + 7) // 8 * 8
        hs = (torch.zeros if Hpad != H else torch.empty)((B, T, Hpad), dtype=torch.bfloat16, device=dev)[:, :, :H]
        cs = torch.empty((T, B, H), dtype=torch.float32, device=dev)
        gates = torch.empty((T, B, H4), dtype=torch.bfloat16, device=dev)
        c_last = torch.empty((B, H), dtype=torch.float32, device=dev)
        counter = torch.zeros(1, dtype=torch.int32, device=dev)
        h0b = None if h0 is None else h0.to(torch.bfloat16).contiguous()
        c0f = None if c0 is None else c0.float().contiguous()
        w16 = whh16 if whh16.is_contiguous() else whh16.contiguous()
        _lib.check(lib.bdl_lstm_seq_fwd(xw.data_ptr(), xw.stride(0), xw.stride(1), w16.data_ptr(), _lib.ptr(h0b),
                                        _lib.ptr(c0f), hs.data_ptr(), hs.stride(0), hs.stride(1), Hpad, cs.data_ptr(),
                                        gates.data_ptr(), c_last.data_ptr(), B, T, H, int(reverse), counter.data_ptr(),
                                        _lib.stream_ptr()), "bdl_lstm_seq_fwd")
        _lib.count_launch(2)
        ctx.save_for_backward(hs, cs, gates, w16, h0b, c0f)
        ctx.cfg = (whh, bool(reverse), B, T, H)
        h_last = hs[:, 0 if reverse else T - 1]
        ctx.mark_non_differentiable(c_last)        # the carried cell state is exposed for getHiddenState only
        return hs, h_last.clone(), c_last

    @staticmethod
    def backward(ctx, dhs, dh_last, _dc_last):
        from .gemm import gemm_bf16

        hs, cs, gates, w16, h0b, c0f = ctx.saved_tensors
        whh, reverse, B, T, H = ctx.cfg
        lib = _lib.cuda_lib()
        dev = hs.device
        if dhs is None:
            dhs = torch.zeros_like(hs)
        dhs = dhs if dhs.dtype == torch.bfloat16 else dhs.to(torch.bfloat16)
        if dh_last is not None:
            dhs = dhs.clone()
            dhs[:, 0 if reverse else T - 1] += dh_last.to(torch.bfloat16)
        dhs = dhs if dhs.stride(2) == 1 else dhs.contiguous()
        Hg = (H + 7) // 8 * 8           # gate blocks of a dG row start 16-byte aligned; pad columns stay zero
        dGp = (torch.zeros if Hg != H else torch.empty)((T, B, 4, Hg), dtype=torch.bfloat16, device=dev)
        dh_rec = torch.zeros((2, B, H), dtype=torch.float32, device=dev)
        dc = torch.zeros((B, H), dtype=torch.float32, device=dev)
        counter = torch.zeros(1, dtype=torch.int32, device=dev)
        _lib.check(lib.bdl_lstm_seq_bwd(dhs.data_ptr(), dhs.stride(0), dhs.stride(1), w16.data_ptr(), gates.data_ptr(),
                                        cs.data_ptr(), _lib.ptr(c0f), dGp.data_ptr(), Hg, dh_rec.data_ptr(), dc.data_ptr(),
                                        B, T, H, int(reverse), counter.data_ptr(), _lib.stream_ptr()), "bdl_lstm_seq_bwd")
        _lib.count_launch(4)
        dG = dGp.reshape(T, B, 4 * H) if Hg == H else dGp[..., :H].reshape(T, B, 4 * H)
        # dW_hh += sum_t dG_t^T h_{prev(t)}: one GEMM over all (t, b) rows
        gw = whh.grad if (whh is not None and whh.requires_grad) else None
        if gw is not None:
            Hpad = (H + 7) // 8 * 8             # operand row pitch of the tensor-core GEMM: a multiple of 8 elements
            hp = torch.empty((T, B, Hpad), dtype=torch.bfloat16, device=dev)[:, :, :H]
            first = T - 1 if reverse else 0
            if reverse:
                hp[:T - 1] = hs[:, 1:].transpose(0, 1)
            else:
                hp[1:] = hs[:, :T - 1].transpose(0, 1)
            if h0b is None:
                hp[first].zero_()
            else:
                hp[first] = h0b
            hp2 = hp.as_strided((T * B, H), (hp.stride(1), 1))
            gemm_bf16(dG.reshape(T * B, 4 * H), hp2, a_mn=True, b_mn=True, out=gw, accumulate=True)
        dxw = dG.permute(1, 0, 2)
        dh0 = dh_rec[T & 1].to(torch.bfloat16) if (h0b is not None and ctx.needs_input_grad[3]) else None
        dc0 = dc if (c0f is not None and ctx.needs_input_grad[4]) else None
        return dxw, None, None, dh0, dc0, None


def lstm_seq_supported(H: int, device) -> bool:
    """Shapes the persistent kernel takes: the W_hh slice of one CTA (4 * ceil(H / #SM) rows, fwd; ceil(H / (#SM / 4))
    columns, bwd) must fit shared memory and 64 rows."""
    sms = torch.cuda.get_device_properties(device).multi_processor_count
    hp = (H + 63) // 64 * 64 + 8
    hj = max(4, -(-H // sms))
    hc = max(8, -(-H // (sms // 4)))
    if 4 * hj > 64 or hc > 64:
        return False
    r1, r2 = (4 * hj + 15) // 16 * 16, (hc + 15) // 16 * 16
    fwd = r1 * hp * 2 + 16 * hp * 2 + hj * 32 * 4 + (0 if 2 * hp >= 32 * r1 else 8 * r1 * 16 * 4)
    bwd = r2 * hp * 2 + 16 * hp * 2 + (0 if 2 * hp >= 32 * r2 else 8 * r2 * 16 * 4)   # (chunk 16; 24 is taken when it fits)
    return max(fwd, bwd) <= 227 * 1024
