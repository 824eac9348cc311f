"""BatchNorm autograd Function over csrc/batchnorm.cu (training statistics, fused ReLU, optional cross-rank sync)."""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch.autograd import Function

from . import _lib
from .functions import nhwc_pitched, rows2d


class BatchNormFn(Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, training, momentum, eps, relu, sync_group):
        lib = _lib.cuda_lib()
        st = _lib.stream_ptr()
        xn, ldx = nhwc_pitched(x)
        N, H, W, C = xn.shape
        M = N * H * W
        y = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=x.device)
        g = weight.detach() if weight is not None else None
        b = bias.detach() if bias is not None else None
        if training:
            sums = torch.zeros(2 * C, dtype=torch.float32, device=x.device)
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
            _lib.check(lib.bdl_bn_apply(xn.data_ptr(), y.data_ptr(), mean.data_ptr(), invstd.data_ptr(), _lib.ptr(g),
                                        _lib.ptr(b), M, C, ldx, C, float(eps), 0, int(relu), st), "bdl_bn_apply")
            _lib.count_launch(3)
            ctx.save_for_backward(xn, mean, invstd, y if relu else None)
            ctx.cfg = (weight, bias, ldx, count, relu, sync_group)
        else:
            _lib.check(lib.bdl_bn_apply(xn.data_ptr(), y.data_ptr(), running_mean.data_ptr(), running_var.data_ptr(),
                                        _lib.ptr(g), _lib.ptr(b), M, C, ldx, C, float(eps), 1, int(relu), st),
                       "bdl_bn_apply")
            _lib.count_launch()
            inv = torch.rsqrt(running_var + eps)
            ctx.save_for_backward(xn, running_mean, inv, y if relu else None)
            ctx.cfg = (weight, bias, ldx, None, relu, None)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        xn, mean, invstd, y = ctx.saved_tensors
        weight, bias, ldx, count, relu, sync_group = ctx.cfg
        lib = _lib.cuda_lib()
        st = _lib.stream_ptr()
        N, H, W, C = xn.shape
        M = N * H * W
        dyn, ldd = nhwc_pitched(dy)
        g = weight.detach() if weight is not None else None
        sums = torch.zeros(2 * C, dtype=torch.float32, device=dy.device)
        _lib.check(lib.bdl_bn_bwd_reduce(dyn.data_ptr(), xn.data_ptr(), _lib.ptr(y), mean.data_ptr(), invstd.data_ptr(),
                                         sums.data_ptr(), M, C, ldd, ldx, C, st), "bdl_bn_bwd_reduce")
        gw = weight.grad if (weight is not None and weight.requires_grad) else None
        gb = bias.grad if (bias is not None and bias.requires_grad) else None
        if gw is not None or gb is not None:
            _lib.check(lib.bdl_bn_param_grad(sums.data_ptr(), _lib.ptr(gw), _lib.ptr(gb), C, 1.0, 1.0, st),
                       "bdl_bn_param_grad")
        dx = None
        if ctx.needs_input_grad[0]:
            dxn = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=dy.device)
            if count is None:
                # inference-mode statistics are constants: dx = gamma * invstd * dy'
                zero = torch.zeros(2 * C, dtype=torch.float32, device=dy.device)
                _lib.check(lib.bdl_bn_bwd_apply(dyn.data_ptr(), xn.data_ptr(), _lib.ptr(y), mean.data_ptr(),
                                                invstd.data_ptr(), _lib.ptr(g), zero.data_ptr(), dxn.data_ptr(), M, 1.0,
                                                C, ldd, ldx, C, C, st), "bdl_bn_bwd_apply")
            else:
                if sync_group is not None and dist.is_initialized() and dist.get_world_size(sync_group) > 1:
                    dist.all_reduce(sums, group=sync_group)
                _lib.check(lib.bdl_bn_bwd_apply(dyn.data_ptr(), xn.data_ptr(), _lib.ptr(y), mean.data_ptr(),
                                                invstd.data_ptr(), _lib.ptr(g), sums.data_ptr(), dxn.data_ptr(), M,
                                                float(count), C, ldd, ldx, C, C, st), "bdl_bn_bwd_apply")
            dx = dxn.permute(0, 3, 1, 2)
        _lib.count_launch(3)
        return dx, None, None, None, None, None, None, None, None, None
