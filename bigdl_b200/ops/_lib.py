"""ctypes loader for the native libraries (C ABI declared in include/bigdl_b200.h).

The CUDA library is mandatory on a machine with a GPU: if it is missing there, importing fails loudly instead of
silently falling back to eager PyTorch. On a CPU-only machine ``cuda_lib()`` returns ``None`` and callers use the
CPU oracle path.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

import torch

import os as _os

# BIGDL_B200_LIBDIR: load the native libraries from another directory (A/B runs of two builds on one box)
_LIBDIR = Path(_os.environ.get("BIGDL_B200_LIBDIR") or (Path(__file__).resolve().parent.parent / "lib"))
_cuda = None
_host = None


class NativeError(RuntimeError):
    pass


def _load(name: str):
    path = _LIBDIR / name
    if not path.exists():
        try:
            from .. import build as _build

            if name == "libbigdl_b200.so":
                _build.build_cuda()
            else:
                _build.build_host()
        except Exception as e:  # pragma: no cover - build environment problems
            raise NativeError(f"{name} is missing and could not be built: {e}") from e
    # The CUDA library is loaded globally (its symbols are looked up by name from other extensions); the host runtime
    # stays local so its C++ runtime symbols cannot interpose on those of OpenCV / torch loaded later.
    return ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL if name == "libbigdl_b200.so" else ctypes.RTLD_LOCAL)


def cuda_lib():
    """The sm_100a kernel library, or None when this process has no CUDA device."""
    global _cuda
    if _cuda is None:
        if not torch.cuda.is_available():
            return None
        _cuda = _load("libbigdl_b200.so")
        _cuda.bdl_last_error.restype = ctypes.c_char_p
        _declare(_cuda)
    return _cuda


_P, _I, _L, _F, _Z, _U64 = (ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_size_t,
                            ctypes.c_ulonglong)

# argument types of every exported function (include/bigdl_b200.h); keeps 64-bit values intact through ctypes
_SIGNATURES = {
    "bdl_gemm_bf16": [_P, _P, _P, _P, _I, _I, _I, _L, _L, _L, _I, _I, _I, _I, _I, _P],
    "bdl_conv_igemm_bf16": [_I, _P, _I, _I, _I, _I, _L, _P, _L, _P, _L, _P] + [_I] * 14 + [_P],
    "bdl_gemm_fp8": [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _L, _L, _L, _I, _I, _P],
    "bdl_row_absmax": [_P, _P, _L, _L, _L, _I, _P],
    "bdl_quantize_e4m3": [_P, _P, _P, _I, _L, _L, _L, _L, _I, _P],
    "bdl_fp8_colscale": [_P, _P, _P, _I, _P],
    "bdl_quantize_int8_rows": [_P, _P, _P, _L, _L, _P],
    "bdl_cast_f32_to_bf16": [_P, _P, _Z, _I, _P],
    "bdl_cast_bf16_to_f32": [_P, _P, _Z, _P],
    "bdl_bf16_trunc_add": [_P, _P, _P, _Z, _P],
    "bdl_eltwise_bf16": [_I, _P, _P, _P, _Z, _F, _P],
    "bdl_addn_bf16": [_P, _I, _P, _Z, _P],
    "bdl_colsum_sumn_relu_bf16": [_P, _I, _P, _P, _P, _L, _I, _P],
    "bdl_colsum_bf16": [_P, _P, _P, _P, _L, _I, _L, _L, _L, _P],
    "bdl_dropout_fwd_bf16": [_P, _P, _P, _Z, _F, _U64, _U64, _P],
    "bdl_dropout_bwd_bf16": [_P, _P, _P, _Z, _F, _P],
    "bdl_copy2d_bf16": [_P, _P, _L, _I, _L, _L, _I, _P],
    "bdl_image_batch_u8": [_P, _P, _P, _P] + [_I] * 6 + [_P, _P, _I, _I, _P],
    "bdl_gru_rz_fwd": [_P] * 5 + [_L, _I, _L, _L, _P],
    "bdl_gru_rz_bwd": [_P] * 6 + [_L, _I, _L, _P],
    "bdl_gru_out_fwd": [_P] * 6 + [_L, _I, _L, _L, _L, _P],
    "bdl_gru_out_bwd": [_P] * 7 + [_L, _I, _L, _L, _P],
    "bdl_gemm_set_residual": [_P, _L],
    "bdl_gemm_set_pair_mode": [_I],
    "bdl_lstm_seq_fwd": [_P, _L, _L, _P, _P, _P, _P, _L, _L, _I, _P, _P, _P, _I, _I, _I, _I, _P, _P],
    "bdl_lstm_seq_bwd": [_P, _L, _L, _P, _P, _P, _P, _P, _I, _P, _P, _I, _I, _I, _I, _P, _P],
    "bdl_quantize_mxfp8": [_P, _P, _P, _L, _L, _L, _L, _I, _P],
    "bdl_gemm_mxfp8": [_P, _P, _P, _P, _P, _P, _I, _I, _I, _L, _L, _L, _I, _I, _P],
    "bdl_scatter_acc_clear_f32": [_P, _P, _P, _P, _I, _P],
    "bdl_softmax_fwd": [_P, _P, _L, _I, _L, _L, _I, _I, _I, _P],
    "bdl_softmax_bwd": [_P, _P, _P, _L, _I, _L, _L, _L, _I, _I, _I, _P],
    "bdl_gemm_bf16_gathered": [_P, _P, _P, _P, _I, _I, _I, _L, _L, _L, _I, _I, _I, _I, _P, _P, _P, _I, _I, _P],
    "bdl_bias_act_finish": [_P, _P, _P, _L, _I, _L, _L, _I, _P],
    "bdl_s2d_stem_bf16": [_P, _I, _P] + [_I] * 4 + [_L] * 4 + [_I] * 5 + [_P],
    "bdl_im2col_nhwc_bf16": [_P, _P] + [_I] * 14 + [_L, _I, _P],
    "bdl_col2im_nhwc_bf16": [_P, _P] + [_I] * 14 + [_L, _I, _P],
    "bdl_maxpool_fwd_nhwc_bf16": [_P, _P, _P] + [_I] * 12 + [_L, _L, _P],
    "bdl_maxpool_bwd_nhwc_bf16": [_P, _P, _P] + [_I] * 12 + [_L, _L, _P],
    "bdl_avgpool_fwd_nhwc_bf16": [_P, _P] + [_I] * 12 + [_L, _L, _I, _I, _P],
    "bdl_avgpool_bwd_nhwc_bf16": [_P, _P] + [_I] * 12 + [_L, _L, _I, _I, _P],
    "bdl_lrn_fwd_nhwc_bf16": [_P, _P, _L, _I, _L, _L, _I, _F, _F, _F, _P],
    "bdl_lrn_bwd_nhwc_bf16": [_P, _P, _P, _L, _I, _L, _L, _L, _I, _F, _F, _F, _P],
    "bdl_xent_fwd": [_P, _P, _P, _P, _P, _L, _I, _L, _I, _I, _I, _I, _P],
    "bdl_xent_bwd": [_P, _P, _P, _P, _P, _P, _L, _I, _L, _L, _I, _I, _I, _I, _F, _P],
    "bdl_xent_fwd_noalloc": [_P, _P, _P, _P, _L, _I, _L, _I, _I, _I, _I, _P],
    "bdl_topk_correct": [_P, _P, _P, _L, _I, _L, _I, _I, _P],
    "bdl_optim_update": [_I, _P, _P, _P, _P, _P, _Z, _P, _P, _I, _I, _P],
    "bdl_sumsq_f32": [_P, _Z, _P, _P],
    "bdl_absmax": [_P, _Z, _I, _P, _P],
    "bdl_scale_f32": [_P, _Z, _F, _P, _F, _P],
    "bdl_bn_stats": [_P, _P, _L, _I, _L, _P],
    "bdl_bn_finalize": [_P, _F, _I, _F, _F, _P, _P, _P, _P, _P],
    "bdl_bn_apply": [_P, _P, _P, _P, _P, _P, _L, _I, _L, _L, _F, _I, _I, _P],
    "bdl_bn_bwd_reduce": [_P, _P, _P, _P, _P, _P, _L, _I, _L, _L, _L, _P],
    "bdl_bn_bwd_apply": [_P, _P, _P, _P, _P, _P, _P, _P, _L, _F, _I, _L, _L, _L, _L, _P],
    "bdl_gemm_set_residual_mask": [_I],
    "bdl_gemm_set_stats": [_P],
    "bdl_gemm_stats_done": [],
    "bdl_bn_bwd_reduce_remask": [_P, _P, _P, _P, _P, _P, _P, _L, _I, _L, _L, _P],
    "bdl_bn_bwd_apply_remask": [_P, _P, _P, _P, _P, _P, _P, _P, _L, _F, _I, _L, _L, _L, _P],
    "bdl_bn_apply_res": [_P, _P, _P, _P, _P, _P, _P, _L, _I, _L, _L, _L, _F, _I, _I, _P],
    "bdl_bn_bwd_apply_mask": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _L, _F, _I, _L, _L, _L, _L, _L, _P],
    "bdl_bn_param_grad": [_P, _P, _P, _I, _F, _F, _P],
    "bdl_lstm_gates_fwd": [_P, _P, _P, _P, _P, _L, _I, _L, _P],
    "bdl_lstm_gates_bwd": [_P, _P, _P, _P, _P, _P, _P, _L, _I, _P],
    "bdl_stream_capture_status": [_P],
    "bdl_compress_grads": [_P, _P, _Z, _I, _I, _P],
    "bdl_fused_rs_update": [_P, _P, _P, _I, _I, _L, _L, _P, _P, _P, _P, _I, _P, _P, _I, ctypes.c_uint, _P, _P, _I, _P,
                            _P],
    "bdl_dropout_fwd_bf16_ctr": [_P, _P, _P, _Z, _F, _U64, _P, _P],
    "bdl_counter_add": [_P, _U64, _P],
    "bdl_bucket_rs_update": [_P, _P, _P, _I, _I, _L, _L, _L, _L, _P, _P, _P, _P, _P, _I, _P, _P, _I, ctypes.c_uint,
                             ctypes.c_uint, _P, _L, _P, _P, _P, _L, _L, _I, _P],
    "bdl_pull_gather": [_P, _P, _I, _I, _P, _I, ctypes.c_uint, ctypes.c_uint, _I, _P, _L, _I, _P],
    "bdl_peer_clip_scale": [_P, _P, _I, _I, ctypes.c_uint, ctypes.c_uint, _P, _L, _P, _F, _P, _P, _P],
    "bdl_ipc_alloc": [_Z, _P, _P],
    "bdl_ipc_open": [_P, _P],
    "bdl_ipc_close": [_P],
    "bdl_ipc_free": [_P],
}


def _declare(lib):
    for name, args in _SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.argtypes = args
            fn.restype = ctypes.c_int


def declare(name: str, args):
    """Register the signature of a function added after the table above was written."""
    _SIGNATURES[name] = args
    if _cuda is not None:
        fn = getattr(_cuda, name, None)
        if fn is not None:
            fn.argtypes = args
            fn.restype = ctypes.c_int


def host_lib():
    global _host
    if _host is None:
        _host = _load("libbigdl_b200_host.so")
    return _host


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = _cuda.bdl_last_error().decode() if _cuda is not None else "?"
        raise NativeError(f"{what} failed ({rc}): {msg}")


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    return None if t is None else t.data_ptr()


LAUNCHES = 0  # number of native kernel-launching calls made by this process (bench.py reports it)


def count_launch(n: int = 1):
    global LAUNCHES
    LAUNCHES += n
