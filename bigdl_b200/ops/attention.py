"""Fused multi-head attention (csrc/attention.cu) as an autograd Function.

``attention(q, k, v, bias, heads, scale, causal)`` takes the *projected* activations in the layout the Q/K/V ``Linear``
layers produce — ``[B, L, heads * D]`` — and returns the combined-heads context ``[B, Lq, heads * D]``: the split /
transpose / combine reshapes of S/nn/Attention.scala:67-114 become tensor-map coordinates. The forward kernel keeps the
logits on chip and saves only the row log-sum-exp; backward recomputes the probabilities from it.
"""
from __future__ import annotations

import ctypes

import torch
from torch.autograd import Function

from . import _lib

_P, _I, _L, _F = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
_lib.declare("bdl_attention_fwd", [_P, _L, _L, _P, _L, _L, _P, _L, _L, _P, _L, _L, _L, _P, _L, _L, _P, _I, _I, _I, _I,
                                   _I, _F, _I, _P])
_lib.declare("bdl_attention_bwd", [_P, _L, _L, _P, _L, _L, _P, _L, _L, _P, _L, _L, _P, _L, _L, _P, _P, _P, _L, _L, _L, _P, _L,
                                   _L, _P, _L, _L, _P, _L, _L, _I, _I, _I, _I, _I, _F, _I, _P])
_lib.declare("bdl_attention_set_two_mode", [_I])


def set_two_tile_mode(mode: int) -> int:
    """-1 heuristic / 0 one query tile per CTA / 1 two tiles per CTA wherever Lq > 128; returns the previous mode."""
    return int(_lib.cuda_lib().bdl_attention_set_two_mode(int(mode)))


def supported(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int) -> bool:
    """Shapes / layouts the kernel takes: CUDA bf16, head size 64 or 128, unit inner stride, 16-byte aligned rows."""
    if not (q.is_cuda and q.dtype == torch.bfloat16 and k.dtype == torch.bfloat16 and v.dtype == torch.bfloat16):
        return False
    if _lib.cuda_lib() is None or q.dim() != 3 or q.shape[2] % heads:
        return False
    if q.shape[2] // heads not in (64, 128):
        return False
    return True


def _rows(t: torch.Tensor) -> torch.Tensor:
    ok = t.stride(2) == 1 and t.stride(1) % 8 == 0 and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0
    return t if ok else t.contiguous()


def _bias_view(bias, B, heads, Lq, Lk):
    """fp32 bias broadcastable to [B, heads, Lq, Lk] -> (tensor, strides over batch / head / query) with key stride 1."""
    if bias is None:
        return None, (0, 0, 0)
    b = bias.float() if bias.dtype != torch.float32 else bias
    while b.dim() < 4:
        b = b.unsqueeze(0)
    if b.shape[3] != Lk:
        b = b.expand(b.shape[0], b.shape[1], b.shape[2], Lk)
    if b.stride(3) != 1:
        b = b.contiguous()
    st = []
    for dim, full in ((0, B), (1, heads), (2, Lq)):
        if b.shape[dim] == 1:
            st.append(0)
        elif b.shape[dim] == full:
            st.append(b.stride(dim))
        else:
            raise ValueError(f"attention bias of shape {tuple(bias.shape)} does not broadcast to {(B, heads, Lq, Lk)}")
    return b, tuple(st)


_FORCE_TORCH_BACKWARD = False        # tests flip this to compare the native backward with the torch composition


class AttentionFn(Function):
    @staticmethod
    def forward(ctx, q, k, v, bias, heads, scale, causal):
        lib = _lib.cuda_lib()
        q, k, v = _rows(q), _rows(k), _rows(v)
        B, Lq, W = q.shape
        Lk = k.shape[1]
        D = W // heads
        out = torch.empty((B, Lq, W), dtype=torch.bfloat16, device=q.device)
        lse = torch.empty((B, heads, Lq), dtype=torch.float32, device=q.device)
        bt, (sb, sh, sq) = _bias_view(bias, B, heads, Lq, Lk)
        _lib.check(lib.bdl_attention_fwd(q.data_ptr(), q.stride(0), q.stride(1), k.data_ptr(), k.stride(0), k.stride(1),
                                         v.data_ptr(), v.stride(0), v.stride(1), _lib.ptr(bt), sb, sh, sq,
                                         out.data_ptr(), out.stride(0), out.stride(1), lse.data_ptr(), B, heads, Lq, Lk,
                                         D, float(scale), int(bool(causal)), _lib.stream_ptr()), "bdl_attention_fwd")
        _lib.count_launch()
        ctx.save_for_backward(q, k, v, out, lse, bt if bt is not None else torch.empty(0, device=q.device))
        ctx.cfg = (heads, float(scale), bool(causal), bias is not None)
        return out

    @staticmethod
    def backward(ctx, d_out):
        q, k, v, out, lse, bias = ctx.saved_tensors
        heads, scale, causal, has_bias = ctx.cfg
        B, Lq, W = q.shape
        Lk, D = k.shape[1], W // heads
        lib = _lib.cuda_lib()
        if lib is not None and not _FORCE_TORCH_BACKWARD:
            # two tcgen05 kernels (csrc/attention_bwd.cu): dQ (+ Delta), then dK / dV; P is recomputed from lse
            go = _rows(d_out if d_out.dtype == torch.bfloat16 else d_out.to(torch.bfloat16))
            qkv_grad = torch.empty((3, B, max(Lq, Lk), W), dtype=torch.bfloat16, device=q.device)
            dq, dk, dv = qkv_grad[0, :, :Lq], qkv_grad[1, :, :Lk], qkv_grad[2, :, :Lk]
            delta = torch.empty((B, heads, Lq), dtype=torch.float32, device=q.device)
            bt = bias if has_bias else None
            sb, sh, sq = (0, 0, 0)
            if bt is not None:
                sb, sh, sq = [0 if bt.shape[d] == 1 else bt.stride(d) for d in (0, 1, 2)]
            _lib.check(lib.bdl_attention_bwd(
                q.data_ptr(), q.stride(0), q.stride(1), k.data_ptr(), k.stride(0), k.stride(1), v.data_ptr(), v.stride(0),
                v.stride(1), out.data_ptr(), out.stride(0), out.stride(1), go.data_ptr(), go.stride(0), go.stride(1),
                lse.data_ptr(), delta.data_ptr(), _lib.ptr(bt), sb, sh, sq, dq.data_ptr(), dq.stride(0), dq.stride(1),
                dk.data_ptr(), dk.stride(0), dk.stride(1), dv.data_ptr(), dv.stride(0), dv.stride(1), B, heads, Lq, Lk, D,
                scale, int(causal), _lib.stream_ptr()), "bdl_attention_bwd")
            _lib.count_launch(2)
            return dq, dk, dv, None, None, None, None

        def split(t, L):
            return t.reshape(B, L, heads, D).transpose(1, 2)          # [B, heads, L, D]
        qh, kh, vh = split(q, Lq), split(k, Lk), split(v, Lk)
        do, o = split(d_out.to(torch.bfloat16), Lq), split(out, Lq)
        s = torch.matmul(qh, kh.transpose(-1, -2)).float() * scale
        if has_bias:
            s = s + bias
        if causal:
            keep = torch.ones(Lq, Lk, dtype=torch.bool, device=q.device).tril(Lk - Lq)
            s = s.masked_fill(~keep, float("-inf"))
        p = torch.exp(s - lse.unsqueeze(-1))                           # probabilities, recomputed from the saved LSE
        p16 = p.to(torch.bfloat16)
        dv = torch.matmul(p16.transpose(-1, -2), do)
        dp = torch.matmul(do, vh.transpose(-1, -2)).float()
        delta = (do.float() * o.float()).sum(-1, keepdim=True)
        ds = (p * (dp - delta) * scale).to(torch.bfloat16)
        dq = torch.matmul(ds, kh)
        dk = torch.matmul(ds.transpose(-1, -2), qh)

        def merge(t, L):
            return t.transpose(1, 2).reshape(B, L, W)
        return merge(dq, Lq), merge(dk, Lk), merge(dv, Lk), None, None, None, None


def attention(q, k, v, bias=None, heads: int = 1, scale: float | None = None, causal: bool = False) -> torch.Tensor:
    """softmax(scale * Q K^T + bias) V over ``heads`` heads of ``[B, L, heads * D]`` projections."""
    if scale is None:
        scale = (q.shape[2] // heads) ** -0.5
    return AttentionFn.apply(q, k, v, bias, heads, scale, causal)


def attention_reference(q, k, v, bias=None, heads: int = 1, scale: float | None = None, causal: bool = False):
    """The same computation with torch ops in fp32 (the unfused composition of S/nn/Attention.scala:95-114)."""
    B, Lq, W = q.shape
    Lk, D = k.shape[1], W // heads
    scale = D ** -0.5 if scale is None else scale
    qh = q.float().reshape(B, Lq, heads, D).transpose(1, 2)
    kh = k.float().reshape(B, Lk, heads, D).transpose(1, 2)
    vh = v.float().reshape(B, Lk, heads, D).transpose(1, 2)
    s = torch.matmul(qh, kh.transpose(-1, -2)) * scale
    if bias is not None:
        s = s + bias.float()
    if causal:
        keep = torch.ones(Lq, Lk, dtype=torch.bool, device=q.device).tril(Lk - Lq)
        s = s.masked_fill(~keep, float("-inf"))
    w = torch.softmax(s, dim=-1)
    return torch.matmul(w, vh).transpose(1, 2).reshape(B, Lq, W)
