"""tcgen05 GEMM binding (csrc/gemm_tcgen05.cu)."""
from __future__ import annotations

import ctypes

import torch

from . import _lib

BIAS, RELU, ATOMIC_ADD = 1, 2, 4
_LOG = [] if __import__("os").environ.get("BDL_LOG_GEMM") else None      # shape log for per-layer roofline analysis


def _log(kind, **kw):
    if _LOG is not None:
        _LOG.append(dict(kind=kind, **kw))
        with open(__import__("os").environ["BDL_LOG_GEMM"], "a") as f:
            f.write(__import__("json").dumps(_LOG[-1]) + "\n")


def _arm_stats(lib, stats, N):
    """Ask the next launch to add the per-column (sum, sum of squares) of its bf16 output into ``stats`` [2N] fp32."""
    if stats is None:
        return
    assert stats.dtype == torch.float32 and stats.numel() == 2 * N and stats.is_contiguous()
    stats._bdl_done = False
    lib.bdl_gemm_set_stats(ctypes.c_void_p(stats.data_ptr()))


def _stats_result(lib, stats):
    if stats is not None:
        stats._bdl_done = bool(lib.bdl_gemm_stats_done())


def _set_residual(lib, residual, M, N, mask: bool = False):
    if mask and residual is not None:
        lib.bdl_gemm_set_residual_mask(1)
    if residual is None:
        return
    assert residual.dtype == torch.bfloat16 and residual.dim() == 2 and tuple(residual.shape) == (M, N) \
        and residual.stride(1) == 1, "residual must be a bf16 [M, N] matrix with unit inner stride"
    _lib.check(lib.bdl_gemm_set_residual(residual.data_ptr(), residual.stride(0)), "bdl_gemm_set_residual")


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False, bias=None, relu=False,
              out: torch.Tensor | None = None, out_dtype=torch.bfloat16, accumulate: bool = False, splits: int = 0,
              residual: torch.Tensor | None = None, stats: torch.Tensor | None = None, res_mask: bool = False):
    """D[M,N] = opA(a) @ opB(b)^T on the tensor cores.

    ``a`` is [M,K] (or [K,M] when ``a_mn``), ``b`` is [N,K] (or [K,N] when ``b_mn``); both bf16, 2-D, unit inner
    stride (row pitch may exceed the row length: channel slices of a wider NHWC tensor are fine).
    ``accumulate`` adds into an fp32 ``out`` with red.global.add (split-K is chosen automatically).
    ``residual`` ([M,N] bf16, unit inner stride) is added before the ReLU in the epilogue (bf16 output only).
    ``stats`` ([2N] fp32, zeroed): the epilogue adds the per-column sum / sum of squares of the output (batch-norm
    statistics for free); ``stats._bdl_done`` says whether this launch could (else run ``bdl_bn_stats``).
    """
    lib = _lib.cuda_lib()
    assert lib is not None, "gemm_bf16 needs a CUDA device"
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.dim() == 2 and b.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    assert K == Kb, f"contraction mismatch {K} vs {Kb}"
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=out_dtype)
    assert out.shape == (M, N) and out.stride(1) == 1
    flags = (BIAS if bias is not None else 0) | (RELU if relu else 0) | (ATOMIC_ADD if accumulate else 0)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.is_contiguous()
    _log("gemm", M=M, N=N, K=K, a_mn=int(a_mn), b_mn=int(b_mn), acc=int(accumulate))
    _set_residual(lib, residual, M, N, res_mask)      # res_mask: out = residual > 0 ? out : 0 (fused ReLU backward)
    _arm_stats(lib, stats, N)
    rc = lib.bdl_gemm_bf16(_lib.ptr(a), _lib.ptr(b), _lib.ptr(out), _lib.ptr(bias), M, N, K,
                           a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn),
                           int(out.dtype == torch.float32), flags, splits, _lib.stream_ptr())
    if stats is not None:
        lib.bdl_gemm_set_stats(None) if rc else _stats_result(lib, stats)
    _lib.check(rc, "bdl_gemm_bf16")
    _lib.count_launch()
    return out


def conv_igemm(mode: int, act: torch.Tensor, ld_act: int, other: torch.Tensor, out: torch.Tensor, O: int, Ho: int,
               Wo: int, kh: int, kw: int, sh: int, sw: int, ph: int, pw: int, dh: int, dw: int, bias=None, relu=False,
               accumulate=False, splits: int = 0, residual=None, stats=None, res_mask: bool = False):
    """Implicit-GEMM convolution (csrc/gemm_tcgen05.cu, TMA im2col). ``act`` is an NHWC view [N,H,W,C] with pixel
    pitch ``ld_act``; ``other`` is the 2-D weight matrix (modes 0/1) or dY [N*Ho*Wo, O] (mode 2); ``out`` is 2-D."""
    lib = _lib.cuda_lib()
    N, H, W, C = act.shape
    flags = (BIAS if bias is not None else 0) | (RELU if relu else 0) | (ATOMIC_ADD if accumulate else 0)
    _log("conv", mode=mode, N=N, H=H, W=W, C=C, O=O, Ho=Ho, Wo=Wo, kh=kh, kw=kw, sh=sh)
    _set_residual(lib, residual, N * Ho * Wo, O, res_mask)
    _arm_stats(lib, stats if mode != 2 else None, O)
    rc = lib.bdl_conv_igemm_bf16(mode, act.data_ptr(), N, H, W, C, ld_act, other.data_ptr(), other.stride(0),
                                 out.data_ptr(), out.stride(0), _lib.ptr(bias), O, Ho, Wo, kh, kw, sh, sw, ph, pw, dh,
                                 dw, int(out.dtype == torch.float32), flags, splits, _lib.stream_ptr())
    if stats is not None and mode != 2:
        lib.bdl_gemm_set_stats(None) if rc else _stats_result(lib, stats)
    _lib.check(rc, "bdl_conv_igemm_bf16")
    _lib.count_launch()
    return out


_ACC_WS: dict = {}


def _acc_workspace(M: int, N: int, device, key) -> torch.Tensor:
    """Persistent zeroed fp32 [M, N] accumulator of one layer (``bdl_bias_act_finish`` re-zeroes it after every use)."""
    k = (key, M, N, device.index)
    t = _ACC_WS.get(k)
    if t is None:
        t = _ACC_WS[k] = torch.zeros(M, N, dtype=torch.float32, device=device)
    return t


def linear_forward(x2: torch.Tensor, w16: torch.Tensor, bias, relu: bool, gather=None, key=None) -> torch.Tensor:
    """y = act(x2 @ w16^T + bias) for a Linear layer (bf16, x2 [M,K], w16 [N,K]).

    Few rows and a long reduction (classifier heads: M = batch, K in the thousands) would leave most SMs idle — one n
    tile each, all of them streaming weights — so those shapes run split-K into a per-layer fp32 accumulator followed
    by one bias/activation pass. ``gather`` (parallel/bucketed.py) names the row blocks of the weight that other ranks
    own: they are read from the owners' memory by the GEMM itself and written through to ``w16``."""
    lib = _lib.cuda_lib()
    M, K = x2.shape
    N = w16.shape[0]
    sms = torch.cuda.get_device_properties(x2.device).multi_processor_count
    bn = 128
    tiles = ((M + 127) // 128) * ((N + bn - 1) // bn)
    split = tiles * 2 <= sms and K >= 2048 and N % 4 == 0
    if gather is None and not split:
        return gemm_bf16(x2, w16, bias=bias, relu=relu)
    y = torch.empty((M, N), device=x2.device, dtype=torch.bfloat16)
    if split:
        acc = _acc_workspace(M, N, x2.device, key if key is not None else w16.data_ptr())
        out, flags, b = acc, ATOMIC_ADD, None
    else:
        out, flags, b = y, (BIAS if bias is not None else 0) | (RELU if relu else 0), bias
    _log("gemm", M=M, N=N, K=K, a_mn=0, b_mn=0, acc=int(split), gathered=int(gather is not None))
    if gather is not None:
        nr = gather["nranks"]
        peers = (ctypes.c_void_p * nr)(*gather["ptrs"])
        t0 = (ctypes.c_int * nr)(*gather["t0"])
        t1 = (ctypes.c_int * nr)(*gather["t1"])
        assert gather["block_n"] == bn
        rc = lib.bdl_gemm_bf16_gathered(_lib.ptr(x2), _lib.ptr(w16), _lib.ptr(out), _lib.ptr(b), M, N, K, x2.stride(0),
                                        w16.stride(0), out.stride(0), int(out.dtype == torch.float32), flags, 0, bn,
                                        peers, t0, t1, nr, gather["rank"], _lib.stream_ptr())
        _lib.check(rc, "bdl_gemm_bf16_gathered")
    else:
        rc = lib.bdl_gemm_bf16(_lib.ptr(x2), _lib.ptr(w16), _lib.ptr(out), _lib.ptr(b), M, N, K, x2.stride(0),
                               w16.stride(0), out.stride(0), 0, 0, int(out.dtype == torch.float32), flags, 0,
                               _lib.stream_ptr())
        _lib.check(rc, "bdl_gemm_bf16")
    _lib.count_launch()
    if split:
        _lib.check(lib.bdl_bias_act_finish(acc.data_ptr(), _lib.ptr(bias), y.data_ptr(), M, N, acc.stride(0), y.stride(0),
                                           int(relu), _lib.stream_ptr()), "bdl_bias_act_finish")
        _lib.count_launch()
    return y
