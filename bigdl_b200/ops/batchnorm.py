"""BatchNorm autograd Function over csrc/batchnorm.cu (training statistics, fused ReLU, optional cross-rank sync)."""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch.autograd import Function

from . import _lib
from .functions import nhwc_pitched, rows2d, zeros_f32


class BatchNormFn(Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, training, momentum, eps, relu, sync_group,
                residual=None, skip_token=None, pre_stats=None):
        """``residual`` (same shape as x): y = relu?(bn(x) + residual) — the BN closing a residual branch, the shortcut
        add and the ReLU in one pass (and one fused backward that also emits the shortcut's gradient)."""
        lib = _lib.cuda_lib()
        st = _lib.stream_ptr()
        xn, ldx = nhwc_pitched(x)
        N, H, W, C = xn.shape
        M = N * H * W
        rn, ldr = (None, 0) if residual is None else nhwc_pitched(residual)
        ctx.skip_token = skip_token.join() if (residual is not None and skip_token is not None) else None

        def apply(mean_t, scale_t, use_var):
            if rn is None:
                _lib.check(lib.bdl_bn_apply(xn.data_ptr(), y.data_ptr(), mean_t.data_ptr(), scale_t.data_ptr(),
                                            _lib.ptr(g), _lib.ptr(b), M, C, ldx, C, float(eps), use_var, int(relu), st),
                           "bdl_bn_apply")
            else:
                _lib.check(lib.bdl_bn_apply_res(xn.data_ptr(), rn.data_ptr(), y.data_ptr(), mean_t.data_ptr(),
                                                scale_t.data_ptr(), _lib.ptr(g), _lib.ptr(b), M, C, ldx, ldr, C,
                                                float(eps), use_var, int(relu), st), "bdl_bn_apply_res")
        y = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=x.device)
        g = weight.detach() if weight is not None else None
        b = bias.detach() if bias is not None else None
        if training:
            if pre_stats is not None and pre_stats.numel() == 2 * C:
                sums = pre_stats          # summed by the producing convolution's GEMM epilogue: no pass over x
            else:
                sums = zeros_f32(2 * C, x.device)
                _lib.check(lib.bdl_bn_stats(xn.data_ptr(), sums.data_ptr(), M, C, ldx, st), "bdl_bn_stats")
            count = float(M)
            if sync_group is not None and dist.is_initialized() and dist.get_world_size(sync_group) > 1:
                dist.all_reduce(sums, group=sync_group)          # sync-BN: global batch statistics
                count *= dist.get_world_size(sync_group)
            mean = torch.empty(C, dtype=torch.float32, device=x.device)
            invstd = torch.empty(C, dtype=torch.float32, device=x.device)
            _lib.check(lib.bdl_bn_finalize(sums.data_ptr(), count, C, float(eps), float(momentum), mean.data_ptr(),
                                           invstd.data_ptr(), running_mean.data_ptr(), running_var.data_ptr(), st),
                       "bdl_bn_finalize")
            apply(mean, invstd, 0)
            _lib.count_launch(3)
            # BN + ReLU: backward recomputes the mask from x with the forward's own scale / shift, so y is not kept for it
            remask = relu and residual is None
            ctx.save_for_backward(xn, mean, invstd, y if (relu and not remask) else None)
            ctx.cfg = (weight, bias, ldx, count, relu, sync_group, residual is not None)
            ctx.remask = remask
        else:
            apply(running_mean, running_var, 1)
            _lib.count_launch()
            inv = torch.rsqrt(running_var + eps)
            ctx.save_for_backward(xn, running_mean, inv, y if relu else None)
            ctx.cfg = (weight, bias, ldx, None, relu, None, residual is not None)
            ctx.remask = False
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        xn, mean, invstd, y = ctx.saved_tensors
        weight, bias, ldx, count, relu, sync_group, has_res = ctx.cfg
        lib = _lib.cuda_lib()
        st = _lib.stream_ptr()
        N, H, W, C = xn.shape
        M = N * H * W
        dyn, ldd = nhwc_pitched(dy)
        g = weight.detach() if weight is not None else None
        sums = zeros_f32(2 * C, dy.device)
        remask = ctx.remask
        bt = bias.detach() if bias is not None else None
        if remask:
            _lib.check(lib.bdl_bn_bwd_reduce_remask(dyn.data_ptr(), xn.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                                    _lib.ptr(g), _lib.ptr(bt), sums.data_ptr(), M, C, ldd, ldx, st),
                       "bdl_bn_bwd_reduce_remask")
        else:
            _lib.check(lib.bdl_bn_bwd_reduce(dyn.data_ptr(), xn.data_ptr(), _lib.ptr(y), mean.data_ptr(),
                                             invstd.data_ptr(), sums.data_ptr(), M, C, ldd, ldx, C, st),
                       "bdl_bn_bwd_reduce")
        gw = weight.grad if (weight is not None and weight.requires_grad) else None
        gb = bias.grad if (bias is not None and bias.requires_grad) else None
        if gw is not None or gb is not None:
            _lib.check(lib.bdl_bn_param_grad(sums.data_ptr(), _lib.ptr(gw), _lib.ptr(gb), C, 1.0, 1.0, st),
                       "bdl_bn_param_grad")
        dx = d_res = None
        # the shortcut's gradient is dy masked by the fused ReLU; without a ReLU it is dy itself
        want_mask = has_res and relu and ctx.needs_input_grad[10]
        if has_res and not relu and ctx.needs_input_grad[10]:
            d_res = dy
        dmn = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=dy.device) if want_mask else None

        def bwd_apply(sums_t, cnt):
            if remask:
                _lib.check(lib.bdl_bn_bwd_apply_remask(dyn.data_ptr(), xn.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                                       _lib.ptr(g), _lib.ptr(bt), sums_t.data_ptr(), dxn.data_ptr(), M,
                                                       cnt, C, ldd, ldx, C, st), "bdl_bn_bwd_apply_remask")
            elif dmn is None:
                _lib.check(lib.bdl_bn_bwd_apply(dyn.data_ptr(), xn.data_ptr(), _lib.ptr(y), mean.data_ptr(),
                                                invstd.data_ptr(), _lib.ptr(g), sums_t.data_ptr(), dxn.data_ptr(), M, cnt,
                                                C, ldd, ldx, C, C, st), "bdl_bn_bwd_apply")
            else:
                _lib.check(lib.bdl_bn_bwd_apply_mask(dyn.data_ptr(), xn.data_ptr(), y.data_ptr(), mean.data_ptr(),
                                                     invstd.data_ptr(), _lib.ptr(g), sums_t.data_ptr(), dxn.data_ptr(),
                                                     dmn.data_ptr(), M, cnt, C, ldd, ldx, C, C, C, st),
                           "bdl_bn_bwd_apply_mask")

        if ctx.needs_input_grad[0] or want_mask:
            dxn = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=dy.device)
            if count is None:
                # inference-mode statistics are constants: dx = gamma * invstd * dy'
                zero = torch.zeros(2 * C, dtype=torch.float32, device=dy.device)
                bwd_apply(zero, 1.0)
            else:
                if sync_group is not None and dist.is_initialized() and dist.get_world_size(sync_group) > 1:
                    dist.all_reduce(sums, group=sync_group)
                bwd_apply(sums, float(count))
            dx = dxn.permute(0, 3, 1, 2)
            if dmn is not None:
                d_res = dmn.permute(0, 3, 1, 2)
        _lib.count_launch(3)
        tok = ctx.skip_token
        if tok is not None and tok.paired and d_res is not None:
            other = tok.exchange(lambda: d_res)
            d_res = None if other is None else d_res + other      # normally first: the paired convolution adds it
        return dx, None, None, None, None, None, None, None, None, None, d_res, None, None
