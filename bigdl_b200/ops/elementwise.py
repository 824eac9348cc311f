"""Bindings for csrc/elementwise.cu, csrc/optim.cu reductions (GPU) with torch fallbacks on CPU tensors."""
from __future__ import annotations

import torch

from . import _lib

RELU, RELU_BWD, ADD, ADD_RELU, MUL, TANH, SIGMOID, TANH_BWD, SIGMOID_BWD, SCALE, AXPY, SUB, MAX, RELU6 = range(14)


def _native(t: torch.Tensor) -> bool:
    return t.is_cuda


def cast_f32_to_bf16(src: torch.Tensor, dst: torch.Tensor | None = None, truncate: bool = False) -> torch.Tensor:
    """fp32 → bf16, round-to-nearest-even or the reference's plain truncation (FP16CompressedTensor.scala:271-278)."""
    assert src.dtype == torch.float32 and src.is_contiguous()
    if dst is None:
        dst = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
    if _native(src):
        lib = _lib.cuda_lib()
        _lib.check(lib.bdl_cast_f32_to_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), int(truncate),
                                            _lib.stream_ptr()), "bdl_cast_f32_to_bf16")
        _lib.count_launch()
    elif truncate:
        dst.view(torch.int16).copy_((src.view(torch.int32) >> 16).to(torch.int16))
    else:
        dst.copy_(src)
    return dst


def cast_bf16_to_f32(src: torch.Tensor, dst: torch.Tensor | None = None) -> torch.Tensor:
    assert src.dtype == torch.bfloat16 and src.is_contiguous()
    if dst is None:
        dst = torch.empty(src.shape, dtype=torch.float32, device=src.device)
    if _native(src):
        lib = _lib.cuda_lib()
        _lib.check(lib.bdl_cast_bf16_to_f32(src.data_ptr(), dst.data_ptr(), src.numel(), _lib.stream_ptr()),
                   "bdl_cast_bf16_to_f32")
        _lib.count_launch()
    else:
        dst.copy_(src)
    return dst


def bf16_trunc_add(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """trunc(expand(a) + expand(b)) on bf16 payloads (FP16CompressedTensor.scala:152-171)."""
    if out is None:
        out = torch.empty_like(a)
    if _native(a):
        lib = _lib.cuda_lib()
        _lib.check(lib.bdl_bf16_trunc_add(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(), _lib.stream_ptr()),
                   "bdl_bf16_trunc_add")
        _lib.count_launch()
    else:
        s = a.float() + b.float()
        out.view(torch.int16).copy_((s.view(torch.int32) >> 16).to(torch.int16))
    return out


def eltwise(op: int, a: torch.Tensor, b: torch.Tensor | None = None, out: torch.Tensor | None = None,
            alpha: float = 1.0) -> torch.Tensor:
    """Dense bf16 elementwise op on tensors that share one memory layout (any dense layout, incl. channels_last)."""
    assert a.dtype == torch.bfloat16 and a.is_cuda
    if out is None:
        out = torch.empty_like(a)
    lib = _lib.cuda_lib()
    _lib.check(lib.bdl_eltwise_bf16(op, a.data_ptr(), _lib.ptr(b), out.data_ptr(), a.numel(), float(alpha),
                                    _lib.stream_ptr()), "bdl_eltwise_bf16")
    _lib.count_launch()
    return out


def addn(tensors, out: torch.Tensor | None = None) -> torch.Tensor:
    """Sum of 2..8 bf16 tensors sharing one dense memory layout, in a single pass (fp32 accumulate)."""
    import ctypes

    a = tensors[0]
    if out is None:
        out = torch.empty_like(a)
    arr = (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])
    lib = _lib.cuda_lib()
    _lib.check(lib.bdl_addn_bf16(arr, len(tensors), out.data_ptr(), a.numel(), _lib.stream_ptr()), "bdl_addn_bf16")
    _lib.count_launch()
    return out


def colsum_sum_relu(tensors, y: torch.Tensor, dz: torch.Tensor, out: torch.Tensor):
    """dz = (y > 0) * sum(tensors); out[c] += column sums of dz. All dense, same layout, [.., C] innermost."""
    import ctypes

    C = out.numel()
    M = y.numel() // C
    arr = (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])
    lib = _lib.cuda_lib()
    _lib.check(lib.bdl_colsum_sumn_relu_bf16(arr, len(tensors), y.data_ptr(), dz.data_ptr(), out.data_ptr(), M, C,
                                             _lib.stream_ptr()), "bdl_colsum_sumn_relu_bf16")
    _lib.count_launch()
    return dz


def colsum(x2d: torch.Tensor, out: torch.Tensor, y2d: torch.Tensor | None = None, dz2d: torch.Tensor | None = None):
    """out[c] += sum_m x[m, c]; with y given, dz = x * (y > 0) is produced and summed (ReLU-bwd + bias-grad fusion)."""
    lib = _lib.cuda_lib()
    M, C = x2d.shape
    _lib.check(lib.bdl_colsum_bf16(x2d.data_ptr(), _lib.ptr(y2d), _lib.ptr(dz2d), out.data_ptr(), M, C, x2d.stride(0),
                                   0 if y2d is None else y2d.stride(0), 0 if dz2d is None else dz2d.stride(0),
                                   _lib.stream_ptr()), "bdl_colsum_bf16")
    _lib.count_launch()
    return out


def copy2d(src2d: torch.Tensor, dst2d: torch.Tensor, accumulate: bool = False):
    lib = _lib.cuda_lib()
    rows, C = src2d.shape
    _lib.check(lib.bdl_copy2d_bf16(src2d.data_ptr(), dst2d.data_ptr(), rows, C, src2d.stride(0), dst2d.stride(0),
                                   int(accumulate), _lib.stream_ptr()), "bdl_copy2d_bf16")
    _lib.count_launch()
    return dst2d


_DROPOUT_CTR: dict = {}


def dropout_counter(device) -> torch.Tensor:
    """Per-device Philox stream position kept in device memory so that CUDA-graph replays draw fresh masks."""
    c = _DROPOUT_CTR.get(device.index)
    if c is None:
        c = torch.zeros(1, dtype=torch.int64, device=device)
        _DROPOUT_CTR[device.index] = c
    return c


def dropout_fwd(x: torch.Tensor, p: float, seed: int, offset: int | None = None):
    lib = _lib.cuda_lib()
    y = torch.empty_like(x)
    mask = torch.empty((x.numel() + 7) // 8, dtype=torch.uint8, device=x.device)
    if offset is None:
        _lib.check(lib.bdl_dropout_fwd_bf16_ctr(x.data_ptr(), y.data_ptr(), mask.data_ptr(), x.numel(), float(p), seed,
                                                dropout_counter(x.device).data_ptr(), _lib.stream_ptr()),
                   "bdl_dropout_fwd_bf16_ctr")
        _lib.count_launch(2)
    else:
        _lib.check(lib.bdl_dropout_fwd_bf16(x.data_ptr(), y.data_ptr(), mask.data_ptr(), x.numel(), float(p), seed,
                                            offset, _lib.stream_ptr()), "bdl_dropout_fwd_bf16")
        _lib.count_launch()
    return y, mask


def dropout_bwd(dy: torch.Tensor, mask: torch.Tensor, p: float):
    lib = _lib.cuda_lib()
    dx = torch.empty_like(dy)
    _lib.check(lib.bdl_dropout_bwd_bf16(dy.data_ptr(), mask.data_ptr(), dx.data_ptr(), dy.numel(), float(p),
                                        _lib.stream_ptr()), "bdl_dropout_bwd_bf16")
    _lib.count_launch()
    return dx


def sumsq(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """out += sum(x^2) (fp32) — building block of L2-norm clipping and LARS (S/parameters/Util.scala:30-50)."""
    if out is None:
        out = torch.zeros(1, dtype=torch.float32, device=x.device)
    if _native(x):
        lib = _lib.cuda_lib()
        _lib.check(lib.bdl_sumsq_f32(x.data_ptr(), x.numel(), out.data_ptr(), _lib.stream_ptr()), "bdl_sumsq_f32")
        _lib.count_launch()
    else:
        out += (x.double() ** 2).sum().float()
    return out


def absmax(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    if out is None:
        out = torch.zeros(1, dtype=torch.float32, device=x.device)
    if _native(x) and x.dtype in (torch.float32, torch.bfloat16) and x.is_contiguous():
        lib = _lib.cuda_lib()
        _lib.check(lib.bdl_absmax(x.data_ptr(), x.numel(), int(x.dtype == torch.bfloat16), out.data_ptr(),
                                  _lib.stream_ptr()), "bdl_absmax")
        _lib.count_launch()
    else:
        torch.maximum(out, x.detach().abs().max().float().reshape(1), out=out)
    return out


def scale_(x: torch.Tensor, alpha: float = 1.0, norm_sq: torch.Tensor | None = None, max_norm: float = 0.0):
    """x *= alpha, or with norm_sq given x *= min(1, max_norm / sqrt(norm_sq)) without a host round trip."""
    if _native(x):
        lib = _lib.cuda_lib()
        _lib.check(lib.bdl_scale_f32(x.data_ptr(), x.numel(), float(alpha), _lib.ptr(norm_sq), float(max_norm),
                                     _lib.stream_ptr()), "bdl_scale_f32")
        _lib.count_launch()
    else:
        if norm_sq is not None:
            nrm = float(norm_sq.sqrt())
            alpha = max_norm / nrm if nrm > max_norm else 1.0
        x.mul_(alpha)
    return x
