"""fp8 (e4m3) quantisation ops: abs-max, quantise, fp8 tcgen05 GEMM / implicit-GEMM conv (csrc/quantize.cu,
csrc/gemm_tcgen05.cu kFp8). Reference: core/bigquant/native, S/nn/quantized/Quantization.scala:35-62,
S/nn/MklInt8Convertible.scala:142-175."""
from __future__ import annotations

import ctypes

import torch

from . import _lib

E4M3_MAX = 448.0
BIAS, RELU = 1, 2


def _lib_or_die():
    lib = _lib.cuda_lib()
    if lib is None:
        raise RuntimeError("fp8 ops need the native CUDA library on a GPU")
    return lib


def row_absmax(x2d: torch.Tensor) -> torch.Tensor:
    """max_k |x[r,k]| per row (fp32). GPU: one warp per row; CPU: torch."""
    if not x2d.is_cuda:
        return x2d.detach().float().abs().amax(dim=1)
    lib = _lib_or_die()
    assert x2d.dim() == 2 and x2d.stride(1) == 1 and x2d.dtype in (torch.float32, torch.bfloat16)
    out = torch.empty(x2d.shape[0], dtype=torch.float32, device=x2d.device)
    _lib.check(lib.bdl_row_absmax(x2d.data_ptr(), out.data_ptr(), x2d.shape[0], x2d.shape[1], x2d.stride(0),
                                  int(x2d.dtype == torch.bfloat16), _lib.stream_ptr()), "bdl_row_absmax")
    _lib.count_launch()
    return out


def quantize_e4m3(x2d: torch.Tensor, absmax: torch.Tensor, per_row: bool = False, pad_to: int = 16) -> torch.Tensor:
    """uint8 tensor of e4m3 bit patterns, q = sat(x * 448 / absmax); row pitch padded to ``pad_to`` bytes (zeros)."""
    rows, cols = x2d.shape
    cp = (cols + pad_to - 1) // pad_to * pad_to
    if not x2d.is_cuda:
        am = absmax.float().reshape(-1, 1) if per_row else absmax.float().reshape(1, 1)
        inv = torch.where(am > 0, E4M3_MAX / am, torch.zeros_like(am))
        v = (x2d.detach().float() * inv).clamp(-E4M3_MAX, E4M3_MAX)
        q = torch.zeros(rows, cp, dtype=torch.uint8)
        q[:, :cols] = v.to(torch.float8_e4m3fn).view(torch.uint8)
        return q
    lib = _lib_or_die()
    assert x2d.stride(1) == 1 and x2d.dtype in (torch.float32, torch.bfloat16)
    q = (torch.empty if cp == cols else torch.zeros)((rows, cp), dtype=torch.uint8, device=x2d.device)
    _lib.check(lib.bdl_quantize_e4m3(x2d.data_ptr(), q.data_ptr(), absmax.data_ptr(), int(per_row), rows, cols,
                                     x2d.stride(0), cp, int(x2d.dtype == torch.bfloat16), _lib.stream_ptr()),
               "bdl_quantize_e4m3")
    _lib.count_launch()
    return q


def dequantize_e4m3(q: torch.Tensor, absmax: torch.Tensor, per_row: bool = False) -> torch.Tensor:
    am = absmax.float().reshape(-1, 1) if per_row else absmax.float().reshape(1, 1)
    return q.view(torch.float8_e4m3fn).float() * (am / E4M3_MAX)


def colscale(act_absmax: torch.Tensor, w_absmax: torch.Tensor) -> torch.Tensor:
    """Dequantisation factor per output column = (act_absmax/448) * (w_absmax[n]/448)."""
    if not w_absmax.is_cuda:
        return (act_absmax.float().reshape(1) / E4M3_MAX) * (w_absmax.float() / E4M3_MAX)
    lib = _lib_or_die()
    out = torch.empty_like(w_absmax)
    _lib.check(lib.bdl_fp8_colscale(act_absmax.data_ptr(), w_absmax.data_ptr(), out.data_ptr(), w_absmax.numel(),
                                    _lib.stream_ptr()), "bdl_fp8_colscale")
    _lib.count_launch()
    return out


def gemm_fp8(a_q: torch.Tensor, b_q: torch.Tensor, col_scale: torch.Tensor, *, k: int | None = None, bias=None,
             relu: bool = False, out: torch.Tensor | None = None, out_dtype=torch.bfloat16, out_qscale=None,
             conv_geom=None):
    """D = (a_q @ b_q^T) * col_scale (+bias)(ReLU). ``a_q`` [M,Kp] / ``b_q`` [N,Kp] are uint8 e4m3 with 16-byte pitches.
    With ``conv_geom = (N,H,W,C,Ho,Wo,kh,kw,sh,sw,ph,pw,dh,dw)`` ``a_q`` is the NHWC activation [N,H,W,Cp] and the
    kernel gathers patches with TMA im2col. ``out_dtype=torch.uint8`` writes e4m3 requantised by ``out_qscale``."""
    lib = _lib_or_die()
    assert a_q.dtype == torch.uint8 and b_q.dtype == torch.uint8 and b_q.dim() == 2
    n = b_q.shape[0]
    if conv_geom is not None:
        g = (ctypes.c_int * 14)(*[int(v) for v in conv_geom])
        m = conv_geom[0] * conv_geom[4] * conv_geom[5]
        kk = conv_geom[6] * conv_geom[7] * conv_geom[3]
        lda = a_q.stride(-2)
        gptr = ctypes.cast(g, ctypes.c_void_p)
    else:
        assert a_q.dim() == 2
        m = a_q.shape[0]
        kk = k if k is not None else min(a_q.shape[1], b_q.shape[1])
        lda = a_q.stride(0)
        gptr = None
    if out is None:
        out = torch.empty((m, n), device=a_q.device, dtype=out_dtype)
    kind = {torch.bfloat16: 0, torch.float32: 1, torch.uint8: 2}[out.dtype]
    flags = (BIAS if bias is not None else 0) | (RELU if relu else 0)
    rc = lib.bdl_gemm_fp8(a_q.data_ptr(), b_q.data_ptr(), out.data_ptr(), _lib.ptr(bias), _lib.ptr(col_scale),
                          _lib.ptr(out_qscale), gptr, m, n, kk, lda, b_q.stride(0), out.stride(0), kind, flags,
                          _lib.stream_ptr())
    _lib.check(rc, "bdl_gemm_fp8")
    _lib.count_launch()
    return out


# ---------------------------------------------------------------------------------------------- MX block scaling
def mx_sf_bytes(rows: int, cols: int) -> int:
    return ((rows + 127) // 128) * ((cols + 127) // 128) * 512


def _mx_sf_index(rows: int, cols: int, device="cpu") -> torch.Tensor:
    """[rows, ceil(cols/32)] -> byte offset of each scale inside the atom buffer (csrc/quantize.cu layout)."""
    kb = (cols + 31) // 32
    kg = (cols + 127) // 128
    r = torch.arange(rows, device=device).unsqueeze(1)
    k = torch.arange(kb, device=device).unsqueeze(0)
    return ((r // 128) * kg + k // 4) * 512 + (r % 32) * 16 + ((r % 128) // 32) * 4 + (k % 4)


def quantize_mxfp8(x2d: torch.Tensor):
    """e4m3 bytes [rows, ceil16(cols)] + UE8M0 scale atoms, one scale per (row, 32 consecutive columns):
    scale = smallest power of two with max|block| / scale <= 448. CPU tensors take the same rule in torch."""
    rows, cols = x2d.shape
    cp = (cols + 15) // 16 * 16
    if not x2d.is_cuda:
        kb = (cols + 31) // 32
        xp = torch.zeros(rows, kb * 32)
        xp[:, :cols] = x2d.detach().float()
        am = xp.reshape(rows, kb, 32).abs().amax(-1)
        e = torch.where(am > 0, torch.ceil(torch.log2(am / E4M3_MAX)), torch.zeros_like(am)).clamp(-127, 127) + 127
        inv = torch.exp2(127 - e).unsqueeze(-1)
        v = (xp.reshape(rows, kb, 32) * inv).clamp(-E4M3_MAX, E4M3_MAX).reshape(rows, kb * 32)[:, :cols]
        q = torch.zeros(rows, cp, dtype=torch.uint8)
        q[:, :cols] = v.to(torch.float8_e4m3fn).view(torch.uint8)
        sf = torch.full((mx_sf_bytes(rows, cols),), 127, dtype=torch.uint8)
        sf[_mx_sf_index(rows, cols).reshape(-1)] = e.to(torch.uint8).reshape(-1)
        return q, sf
    lib = _lib_or_die()
    assert x2d.stride(1) == 1 and x2d.dtype in (torch.float32, torch.bfloat16)
    q = (torch.empty if cp == cols else torch.zeros)((rows, cp), dtype=torch.uint8, device=x2d.device)
    sf = torch.empty(mx_sf_bytes(rows, cols), dtype=torch.uint8, device=x2d.device)
    _lib.check(lib.bdl_quantize_mxfp8(x2d.data_ptr(), q.data_ptr(), sf.data_ptr(), rows, cols, x2d.stride(0), cp,
                                      int(x2d.dtype == torch.bfloat16), _lib.stream_ptr()), "bdl_quantize_mxfp8")
    _lib.count_launch()
    return q, sf


def dequantize_mxfp8(q: torch.Tensor, sf: torch.Tensor, cols: int) -> torch.Tensor:
    """fp32 [rows, cols] that the block-scaled operand (q, sf) represents exactly."""
    rows = q.shape[0]
    e = sf[_mx_sf_index(rows, cols, sf.device).reshape(-1)].reshape(rows, -1).float()
    scale = torch.exp2(e - 127).repeat_interleave(32, dim=1)[:, :cols]
    return q[:, :cols].view(torch.float8_e4m3fn).float() * scale


def gemm_mxfp8(a_q: torch.Tensor, sfa: torch.Tensor, b_q: torch.Tensor, sfb: torch.Tensor, k: int, *, bias=None,
               relu: bool = False, out: torch.Tensor | None = None, out_dtype=torch.bfloat16):
    """D = A B^T for block-scaled e4m3 operands (``quantize_mxfp8``); the tensor core applies the per-block scales."""
    lib = _lib_or_die()
    m, n = a_q.shape[0], b_q.shape[0]
    if out is None:
        out = torch.empty((m, n), device=a_q.device, dtype=out_dtype)
    flags = (BIAS if bias is not None else 0) | (RELU if relu else 0)
    rc = lib.bdl_gemm_mxfp8(a_q.data_ptr(), b_q.data_ptr(), sfa.data_ptr(), sfb.data_ptr(), out.data_ptr(), _lib.ptr(bias),
                            m, n, k, a_q.stride(0), b_q.stride(0), out.stride(0), int(out.dtype == torch.float32), flags,
                            _lib.stream_ptr())
    _lib.check(rc, "bdl_gemm_mxfp8")
    _lib.count_launch()
    return out
