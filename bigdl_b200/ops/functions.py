"""GPU autograd Functions built on the sm_100a kernels (the "mkldnn engine" of this framework).

Activations are bf16, logically NCHW but stored NHWC (torch ``channels_last``), so every convolution is a GEMM over
``[N*Ho*Wo, kh*kw*C]`` and every 1x1 convolution is a GEMM on the activation tensor itself. Parameter gradients are
accumulated *directly into the flat fp32 gradient buffer* (``param.grad``) by the split-K wgrad GEMM — the
``accGradParameters`` contract of the reference — so backward returns ``None`` for parameter inputs.

Reference call stack being replaced: S/nn/SpatialConvolution.scala:253-361 (updateOutput), :364-427 (updateGradInput),
:435-521 (accGradParameters); S/nn/Linear.scala:78-167; S/nn/mkldnn/SpatialConvolution.scala:435-466.
"""
from __future__ import annotations

import torch
from torch.autograd import Function

from . import _lib
from . import elementwise as ew
from .gemm import conv_igemm, gemm_bf16

_WS: dict = {}

# ---- weight-gradient side stream ----------------------------------------------------------------------------------
# dgrad is on the critical path of backward (the next layer needs it), wgrad is not: it only has to finish before the
# optimizer step. wgrad GEMMs are therefore issued on a per-device side stream, where they fill the SMs that the small
# dgrad kernels of the following layers leave idle. ``join_side_streams`` must run before anything reads the gradient
# buffer (training loops call it right after ``autograd.backward``; it is part of the captured CUDA graph).
_WGRAD_STREAM: dict = {}
_WGRAD_KEEP: list = []          # operands of in-flight wgrad GEMMs, kept alive until the join


def _wgrad_side(device):
    from ..utils import config

    if not config.get("bigdl.wgrad.stream"):
        return None
    st = _WGRAD_STREAM.get(device.index)
    if st is None:
        st = _WGRAD_STREAM[device.index] = torch.cuda.Stream(device=device)
    return st


# ---- pre-zeroed fp32 scratch ------------------------------------------------------------------------------------
# Accumulator vectors that kernels add into with red.global (batch-norm sums, ...) used to be a torch.zeros each: one
# fill launch per use, 113 per ResNet-50 step. They are now slices of one zeroed buffer per device, handed out in order
# and re-zeroed with ONE fill when the step's backward has finished (join_side_streams, part of the captured graph).
_ZERO_POOL: dict = {}
_ZERO_POOL_LIMIT = 16 << 20          # floats; beyond this (nobody resets the pool) callers get a plain torch.zeros


def zeros_f32(n: int, device) -> torch.Tensor:
    st = _ZERO_POOL.get(device.index)
    if st is None:
        st = _ZERO_POOL[device.index] = {"chunks": [], "cur": -1, "off": 0, "total": 0}
    n4 = (n + 3) // 4 * 4
    while True:
        if st["cur"] >= 0:
            buf = st["chunks"][st["cur"]]
            if st["off"] + n4 <= buf.numel():
                out = buf[st["off"]:st["off"] + n]
                st["off"] += n4
                st["total"] += n4
                return out
        if st["total"] + n4 > _ZERO_POOL_LIMIT:
            return torch.zeros(n, dtype=torch.float32, device=device)
        if st["cur"] + 1 < len(st["chunks"]) and st["chunks"][st["cur"] + 1].numel() >= n4:
            st["cur"] += 1
        else:
            st["chunks"].insert(st["cur"] + 1, torch.zeros(max(n4, 1 << 18), dtype=torch.float32, device=device))
            st["cur"] += 1
        st["off"] = 0


def reset_zero_pool(device=None):
    """Re-zero what was handed out since the last reset; every slice must be dead by now (end of backward)."""
    for idx, st in _ZERO_POOL.items():
        if device is not None and device.index not in (None, idx):
            continue
        for k in range(st["cur"] + 1):
            buf = st["chunks"][k]
            used = st["off"] if k == st["cur"] else buf.numel()
            if used:
                buf[:used].zero_()
        st["cur"], st["off"], st["total"] = (0 if st["chunks"] else -1), 0, 0


def join_side_streams(device=None):
    """Make the current stream wait for every outstanding side-stream weight-gradient GEMM (and drop per-backward
    side-channel state)."""
    _PREMASKED.clear()
    reset_zero_pool(device)
    if not _WGRAD_KEEP:
        return
    cur = torch.cuda.current_stream(device)
    for idx, st in _WGRAD_STREAM.items():
        if device is None or device.index in (None, idx):
            cur.wait_stream(st)
    _WGRAD_KEEP.clear()


def workspace(numel: int, device, slot: int = 0) -> torch.Tensor:
    """Grow-only bf16 scratch buffer shared by the conv layers of a device that run on the SAME stream (im2col / dgrad
    columns). Keyed by stream: Concat branches on side streams (``bigdl.concat.streams``) get their own buffer, so two
    branches that both take the explicit im2col path cannot overwrite each other's columns."""
    key = (device.index, slot, torch.cuda.current_stream(device).cuda_stream)
    buf = _WS.get(key)
    if buf is None or buf.numel() < numel:
        buf = torch.empty(max(numel, 1 << 20), dtype=torch.bfloat16, device=device)
        _WS[key] = buf
    return buf[:numel]


def _nhwc(x: torch.Tensor) -> torch.Tensor:
    """[N,C,H,W] logical → contiguous [N,H,W,C] view (copying only if the tensor is not channels_last yet)."""
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    x = x.contiguous(memory_format=torch.channels_last)
    return x.permute(0, 2, 3, 1)


def nhwc_pitched(x: torch.Tensor):
    """[N,C,H,W] logical → ([N,H,W,C] view, pixel pitch). A channel slice of a dense NHWC tensor (what autograd hands
    back for one branch of a channel concat) is used in place — the kernels take the pitch — instead of being copied."""
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    xp = x.permute(0, 2, 3, 1)
    N, H, W, C = xp.shape
    st = xp.stride()
    ld = st[2] if W > 1 else (st[1] if H > 1 else (st[0] if N > 1 else C))
    ok = (st[3] == 1 or C == 1) and ld >= C and ld % 8 == 0 and (W == 1 or st[2] == ld) and \
        (H == 1 or st[1] == W * ld) and (N == 1 or st[0] == H * W * ld) and xp.data_ptr() % 16 == 0
    if ok:
        return xp, ld
    return x.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1), C


def rows2d(xp: torch.Tensor, ld: int) -> torch.Tensor:
    """[N,H,W,C] (pixel pitch ld) → [N*H*W, C] view with row pitch ld."""
    N, H, W, C = xp.shape
    return torch.as_strided(xp, (N * H * W, C), (ld, 1), xp.storage_offset())


_OUT_HINT: list = [None]
_IGEMM: list = [True]      # implicit-GEMM (TMA im2col) convolution path; False forces the explicit im2col path


def set_out_hint(buf: torch.Tensor | None):
    """The next ConvFn.forward writes its output into ``buf`` ([N,Ho,Wo,O] view, e.g. a channel slice of a concat
    buffer) instead of allocating — how Concat/JoinTable become zero-copy (cf. S/nn/mkldnn/Fusion.scala JoinTable)."""
    _OUT_HINT[0] = buf


def _take_out_hint():
    h, _OUT_HINT[0] = _OUT_HINT[0], None
    return h


_RES_HINT: list = [None]


def set_residual_hint(res):
    """``(shortcut [N,O,Ho,Wo] bf16, relu)``: the next ConvFn.forward that can (1x1 direct or implicit-GEMM path, no
    autograd) adds the shortcut and applies the ReLU in its GEMM epilogue — the conv + sum fusion of a residual block
    (S/nn/mkldnn/Fusion.scala fusionCAddTable). ``take_residual_hint`` returns it if nobody consumed it."""
    _RES_HINT[0] = res


def take_residual_hint():
    h, _RES_HINT[0] = _RES_HINT[0], None
    return h


# Gradients that already had a consumer's ReLU mask applied and its bias gradient summed, keyed by (data_ptr, shape):
# ConcatViewFn.backward does that once for all branch tails of a Concat; the tails' ConvFn.backward picks its slice up
# here instead of running its own column-sum pass. Cleared after every backward (join_side_streams).
_PREMASKED: dict = {}
_BIAS_DONE = object()            # _PREMASKED value: masked AND the bias gradient is already in the flat buffer
_NEED_COLSUM = object()          # _PREMASKED value: masked, but the bias-gradient column sums are still to be taken
_RELU_DGRAD: list = [True]       # fold the producer's ReLU backward (+ bias-gradient sums) into the consumer's dgrad epilogue
_SUMS: dict = {}                 # persistent, self-clearing fp32 column-sum vectors, one per Concat (see _concat_sums)


def _concat_sums(key, Ct: int, device) -> torch.Tensor:
    """Zeroed fp32 [Ct] scratch owned by one Concat: ``_scatter_bias_grads`` clears it again after use, so no fill kernel
    runs per step (it is allocated during the eager warm-up iterations that precede a graph capture)."""
    k = (key, Ct, device.index)
    t = _SUMS.get(k)
    if t is None:
        t = _SUMS[k] = torch.zeros(Ct, dtype=torch.float32, device=device)
    return t


def _tail_bias_grads(tails):
    """Bias-gradient views of a Concat's branch tails when ONE scatter kernel may feed them all, else None."""
    if not tails or len(tails) > 8:
        return None
    out = []
    for t in tails:
        b = t._parameters.get("bias")
        if b is None or not b.requires_grad or b.grad is None or t.scaleB != 1.0 or not getattr(t, "_fuse_relu", False):
            return None
        out.append(b.grad)
    return out


def _scatter_bias_grads(sums: torch.Tensor, grads, sizes):
    import ctypes

    lib = _lib.cuda_lib()
    n = len(grads)
    dst = (ctypes.c_void_p * n)(*[g.data_ptr() for g in grads])
    offs, c = [], 0
    for sz in sizes:
        offs.append(c)
        c += sz
    _lib.check(lib.bdl_scatter_acc_clear_f32(sums.data_ptr(), dst, (ctypes.c_int * n)(*offs), (ctypes.c_int * n)(*sizes),
                                             n, _lib.stream_ptr()), "bdl_scatter_acc_clear_f32")
    _lib.count_launch()


class ConcatViewFn(Function):
    """The branches already wrote into channel slices of ``big`` ([N,H,W,Ctot]); expose it as one NCHW tensor whose
    gradient is handed back to the branches as (pitched) channel slices — no copies in either direction.

    When every branch ends in a conv with fused bias + ReLU (``relu_all``: ``big`` then IS the ReLU output of all of
    them), the backward runs ONE fused ReLU-mask + bias-gradient column sum over the full width instead of one narrow,
    pitched pass per branch (four per Inception module), and hands out slices of the masked gradient."""

    @staticmethod
    def forward(ctx, big, relu_all, tails, *outs):
        ctx.sizes = [o.shape[1] for o in outs]
        ctx.relu_all = bool(relu_all)
        ctx.tails = list(tails) if tails else None        # the branches' last convolutions (bias-gradient owners)
        if ctx.relu_all:
            ctx.save_for_backward(big)
        return big.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, g):
        ready = ctx.__dict__.pop("_ready", None)
        tail_grads = _tail_bias_grads(ctx.tails) if ctx.relu_all else None
        if ready is not None and ready[0].data_ptr() == g.data_ptr() and tuple(g.shape[2:]) == tuple(ready[0].shape[1:3]):
            sums = ready[1]              # the consumer's fan-out node already summed, masked and column-summed
        elif ctx.relu_all and g.is_cuda and g.dtype == torch.bfloat16:
            (big,) = ctx.saved_tensors
            gn, ld = nhwc_pitched(g)
            N, H, W, Ct = gn.shape
            dz = torch.empty((N, H, W, Ct), dtype=torch.bfloat16, device=g.device)
            sums = _concat_sums(id(ctx.tails[0]), Ct, g.device) if tail_grads is not None else \
                torch.zeros(Ct, dtype=torch.float32, device=g.device)
            if ready is not None and tail_grads is not None:
                sums.zero_()             # a parked result that was not ours to use left its sums in the shared vector
            ew.colsum(rows2d(gn, ld), sums, rows2d(big, Ct), rows2d(dz, Ct))
            g = dz.permute(0, 3, 1, 2)
        else:
            sums = None
        done = False
        if sums is not None and tail_grads is not None:
            # one kernel feeds every tail's bias gradient and clears the vector (the tails then add nothing themselves)
            _scatter_bias_grads(sums, tail_grads, ctx.sizes)
            done = True
        grads, c = [], 0
        for n in ctx.sizes:
            sl = g[:, c:c + n]
            if sums is not None:
                _PREMASKED[(sl.data_ptr(), tuple(sl.shape))] = _BIAS_DONE if done else sums[c:c + n]
            grads.append(sl)
            c += n
        return (None, None, None, *grads)


def _acc_grad(param: torch.Tensor, scale: float):
    return param.grad if (param is not None and param.requires_grad and scale != 0.0) else None


class LinearFn(Function):
    @staticmethod
    def forward(ctx, x, weight, bias, w16, relu, scale_w, scale_b, gather=None):
        K = w16.shape[1]
        x2 = x.reshape(-1, K)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        x2 = x2.contiguous()
        if K % 8:
            raise _lib.NativeError("Linear on the tensor-core path needs inputSize % 8 == 0")
        b32 = bias.detach() if bias is not None else None
        from .gemm import linear_forward

        y = linear_forward(x2, w16, b32, relu, gather=gather, key=id(weight))
        ctx.save_for_backward(x2, w16, y if relu else None)
        ctx.params = (weight, bias, relu, scale_w, scale_b, x.shape)
        return y.reshape(*x.shape[:-1], w16.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w16, y = ctx.saved_tensors
        weight, bias, relu, scale_w, scale_b, xshape = ctx.params
        N = w16.shape[0]
        dy2 = dy.reshape(-1, N)
        if dy2.dtype != torch.bfloat16:
            dy2 = dy2.to(torch.bfloat16)
        dy2 = dy2.contiguous()
        gb = _acc_grad(bias, scale_b)
        if relu:
            dz = torch.empty_like(dy2)
            gtmp = gb if (gb is not None and scale_b == 1.0) else torch.zeros(N, device=dy.device)
            ew.colsum(dy2, gtmp, y, dz)
            if gb is not None and scale_b != 1.0:
                gb.add_(gtmp, alpha=scale_b)
        else:
            dz = dy2
            if gb is not None:
                if scale_b == 1.0:
                    ew.colsum(dy2, gb)
                else:
                    gb.add_(ew.colsum(dy2, torch.zeros(N, device=dy.device)), alpha=scale_b)
        gw = _acc_grad(weight, scale_w)
        wk = w16
        if N % 8:
            # dY is an MN-major operand of both backward GEMMs, so its row pitch must be a multiple of 8 elements:
            # zero-pad the output dimension (small classifiers such as 10-way heads)
            n8 = (N + 7) // 8 * 8
            dz = torch.nn.functional.pad(dz, (0, n8 - N))
            wk = torch.nn.functional.pad(w16, (0, 0, 0, n8 - N))
        if gw is not None:
            if scale_w == 1.0 and gw.is_contiguous() and N % 8 == 0:
                side = _wgrad_side(dz.device)
                if side is not None:
                    side.wait_stream(torch.cuda.current_stream(dz.device))
                    with torch.cuda.stream(side):
                        gemm_bf16(dz, x2, a_mn=True, b_mn=True, out=gw, accumulate=True)
                    _WGRAD_KEEP.append((dz, x2))
                else:
                    gemm_bf16(dz, x2, a_mn=True, b_mn=True, out=gw, accumulate=True)
            else:
                tmp = torch.zeros((dz.shape[1], x2.shape[1]), device=gw.device)
                gemm_bf16(dz, x2, a_mn=True, b_mn=True, out=tmp, accumulate=True)
                gw.add_(tmp[:N], alpha=scale_w)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = gemm_bf16(dz, wk, b_mn=True).reshape(xshape)
        return dx, None, None, None, None, None, None, None


_STEM_S2D: list = [True]


def _stem_s2d_plan(x, w16, geom, needs_dx):
    """Decide whether a stride-2 few-channel convolution (the 7x7/2 RGB stems of Inception and ResNet) runs as a
    space-to-depth + stride-1 implicit GEMM. The explicit path writes and re-reads a [N*Ho*Wo, 7*7*4] patch matrix
    (0.4 GB at batch 128); the folded input is 16 channels wide, which the TMA-im2col kernel accepts directly."""
    kh, kw, sh, sw, pt, pl, pb, pr, dh, dw, groups = geom
    N, C, H, W = x.shape
    O = w16.shape[0]
    if not (_STEM_S2D[0] and _IGEMM[0] and groups == 1 and C <= 4 and sh == 2 and sw == 2 and dh == 1 and dw == 1
            and pt == pb and pl == pr and O % 8 == 0 and not needs_dx and kh >= 3 and kw >= 3
            and x.dtype in (torch.float32, torch.bfloat16)):
        return None
    th, tw = (kh + 1) // 2, (kw + 1) // 2
    R, S = (H + 2 * pt + 1) // 2, (W + 2 * pl + 1) // 2
    Ho, Wo = (H + 2 * pt - kh) // 2 + 1, (W + 2 * pl - kw) // 2 + 1
    if R - th + 1 != Ho or S - tw + 1 != Wo:
        return None
    return th, tw, R, S, Ho, Wo


def _stem_s2d_weight(w4, th, tw):
    """[O,kh,kw,C] → [O, th*tw*16] with tap (i,j), channel (a*2+b)*C+c holding w[o, 2i+a, 2j+b, c]."""
    O, kh, kw, C = w4.shape
    wp = torch.nn.functional.pad(w4, (0, 0, 0, 2 * tw - kw, 0, 2 * th - kh))
    wp = wp.reshape(O, th, 2, tw, 2, C).permute(0, 1, 3, 2, 4, 5).reshape(O, th, tw, 4 * C)
    return torch.nn.functional.pad(wp, (0, 16 - 4 * C)).reshape(O, th * tw * 16)


def _stem_s2d_forward(ctx, x, weight, bias, w16, geom, relu, scale_w, scale_b, plan):
    kh, kw = geom[0], geom[1]
    pt, pl = geom[4], geom[5]
    th, tw, R, S, Ho, Wo = plan
    N, C, H, W = x.shape
    O = w16.shape[0]
    lib = _lib.cuda_lib()
    # horizontal taps folded into channels as well: [N, R, Wo, tw*16] and a th x 1 convolution (full 128-byte gathers)
    x2 = torch.empty((N, R, Wo, tw * 16), dtype=torch.bfloat16, device=x.device)
    st = x.stride()
    _lib.check(lib.bdl_s2d_stem_bf16(x.data_ptr(), int(x.dtype == torch.float32), x2.data_ptr(), N, C, H, W, st[0],
                                     st[1], st[2], st[3], pt, pl, R, Wo, tw, _lib.stream_ptr()), "bdl_s2d_stem_bf16")
    _lib.count_launch()
    wk = _stem_s2d_weight(w16.permute(0, 2, 3, 1), th, tw)
    hint = _take_out_hint()
    if hint is not None and tuple(hint.shape) == (N, Ho, Wo, O):
        y, ldy = hint, hint.stride(2)
    else:
        y, ldy = torch.empty((N, Ho, Wo, O), dtype=torch.bfloat16, device=x.device), O
    b32 = bias.detach() if bias is not None else None
    conv_igemm(0, x2, tw * 16, wk, rows2d(y, ldy), O, Ho, Wo, th, 1, 1, 1, 0, 0, 1, 1, bias=b32, relu=relu)
    ctx.s2d = (plan, (N, C, H, W, O, kh, kw), ldy)
    ctx.save_for_backward(x2, y if relu else None)
    ctx.params = (weight, bias, geom, relu, scale_w, scale_b, None, x.shape)
    return y.permute(0, 3, 1, 2)


def _stem_s2d_backward(ctx, dy):
    x2, y = ctx.saved_tensors
    weight, bias, geom, relu, scale_w, scale_b, _, _ = ctx.params
    (th, tw, R, S, Ho, Wo), (N, C, H, W, O, kh, kw), ldy = ctx.s2d
    M = N * Ho * Wo
    dy4, ldd = nhwc_pitched(dy)
    dyn = rows2d(dy4, ldd)
    gb = _acc_grad(bias, scale_b)
    if relu:
        dz = torch.empty((M, O), dtype=torch.bfloat16, device=dy.device)
        gtmp = gb if (gb is not None and scale_b == 1.0) else torch.zeros(O, device=dy.device)
        ew.colsum(dyn, gtmp, rows2d(y, ldy), dz)
        if gb is not None and scale_b != 1.0:
            gb.add_(gtmp, alpha=scale_b)
    else:
        dz = dyn
        if gb is not None:
            if scale_b == 1.0:
                ew.colsum(dyn, gb)
            else:
                gb.add_(ew.colsum(dyn, torch.zeros(O, device=dy.device)), alpha=scale_b)
    gw = _acc_grad(weight, scale_w)
    if gw is not None:
        tmp = torch.zeros((O, th * tw * 16), device=dy.device)
        conv_igemm(2, x2, tw * 16, dz, tmp, O, Ho, Wo, th, 1, 1, 1, 0, 0, 1, 1, accumulate=True)
        g = tmp.view(O, th, tw, 16)[..., :4 * C].reshape(O, th, tw, 2, 2, C).permute(0, 1, 3, 2, 4, 5)
        g = g.reshape(O, 2 * th, 2 * tw, C)[:, :kh, :kw]
        gw.permute(0, 2, 3, 1).add_(g, alpha=scale_w)
    return None, None, None, None, None, None, None, None


class SkipToken:
    """Pairs the two consumers of a residual block's input so their input gradients are summed without a pass of
    their own. Whichever backward runs first deposits its gradient and returns None to autograd; the second adds the
    deposit — a convolution does it in its dgrad GEMM epilogue — and returns the sum. (Autograd alone would run an
    element-wise add over the block input: 16 per ResNet-50 step.)"""

    __slots__ = ("grad", "holders")

    def __init__(self):
        self.grad = None
        self.holders = 0          # consumers that took the token in forward; the protocol needs exactly two

    def join(self):
        self.holders += 1
        return self

    @property
    def paired(self) -> bool:
        return self.holders == 2

    def exchange(self, mine_ready):
        """Returns the other side's deposit, or None after storing ``mine_ready()`` as this side's."""
        if self.grad is None:
            self.grad = mine_ready()
            return None
        other, self.grad = self.grad, None
        return other


class ConvFn(Function):
    """2-D convolution, NHWC bf16, explicit-GEMM on the tcgen05 kernel (1x1/s1/p0 convs skip im2col entirely)."""

    @staticmethod
    def forward(ctx, x, weight, bias, w16, geom, relu, scale_w, scale_b, skip_token=None, stats=None,
                mask_input=False):
        kh, kw, sh, sw, pt, pl, pb, pr, dh, dw, groups = geom
        ctx.s2d = None
        ctx.skip_token = skip_token.join() if skip_token is not None else None
        # the input is the (fused) ReLU output of another convolution: our dgrad can apply that layer's ReLU backward
        # and sum its bias gradient in the epilogue (see backward)
        ctx.mask_input = bool(mask_input)
        plan = _stem_s2d_plan(x, w16, geom, ctx.needs_input_grad[0])
        if plan is not None:
            return _stem_s2d_forward(ctx, x, weight, bias, w16, geom, relu, scale_w, scale_b, plan)
        xn, ldx = nhwc_pitched(x)
        N, H, W, C = xn.shape
        O = w16.shape[0]
        Ho = (H + pt + pb - dh * (kh - 1) - 1) // sh + 1
        Wo = (W + pl + pr - dw * (kw - 1) - 1) // sw + 1
        M = N * Ho * Wo
        c_real = C
        w4 = w16.permute(0, 2, 3, 1)                        # OHWI physical order
        if groups == 1 and C % 8 != 0 and C < 8:
            # tiny channel counts (the RGB stem): pad to 4 (or 8) channels so a vector im2col path applies; 4 keeps the
            # patch matrix half the size of the 8-channel padding
            cp = 4 if C <= 4 else 8
            xn = torch.nn.functional.pad(xn, (0, cp - C))
            w4 = torch.nn.functional.pad(w4, (0, cp - C))
            C, ldx = cp, cp
        asym = (pt != pb) or (pl != pr)
        if asym:
            # TF-style SAME padding can be asymmetric: materialise the extra bottom/right rows once
            xn = torch.nn.functional.pad(xn, (0, 0, 0, pr - pl if pr > pl else 0, 0, pb - pt if pb > pt else 0))
            N, H, W, C = xn.shape
            ldx = C
        Cg, Og = C // groups, O // groups
        hint = _take_out_hint()
        if hint is not None and tuple(hint.shape) == (N, Ho, Wo, O):
            y, ldy = hint, hint.stride(2)
        else:
            y, ldy = torch.empty((N, Ho, Wo, O), dtype=torch.bfloat16, device=x.device), O
        y2 = rows2d(y, ldy)
        x2 = rows2d(xn, ldx)
        w2 = w4.reshape(O, kh * kw * Cg)                     # [O, K]
        direct = (kh == 1 and kw == 1 and sh == 1 and sw == 1 and pt == pl == pb == pr == 0 and Cg % 8 == 0)
        K = kh * kw * Cg
        Kp = (K + 7) // 8 * 8
        lib = _lib.cuda_lib()
        b32 = bias.detach() if bias is not None else None
        implicit = (not direct) and groups == 1 and C % 8 == 0 and C >= 16 and O % 8 == 0 and _IGEMM[0]
        kept_cols = None
        res2d = None
        if _RES_HINT[0] is not None and groups == 1 and (implicit or direct):     # only set under no_grad (ConcatTable)
            rs, rrelu = take_residual_hint()
            if tuple(rs.shape) == (N, O, Ho, Wo) and rs.dtype == torch.bfloat16:
                rn, ldr = nhwc_pitched(rs)
                res2d, relu = rows2d(rn, ldr), bool(relu or rrelu)
                ctx.fused_residual = True
            else:
                set_residual_hint((rs, rrelu))          # shapes disagree: leave it to the caller's add
        if stats is not None and (relu or res2d is not None or groups != 1 or not (implicit or direct)):
            stats = None                  # column statistics are only taken of a plain conv output (the BN's input)
        if implicit:
            conv_igemm(0, xn, ldx, w2, y2, O, Ho, Wo, kh, kw, sh, sw, pt, pl, dh, dw, bias=b32, relu=relu,
                       residual=res2d, stats=stats)
        for g in range(groups if not implicit else 0):
            wg = w2[g * Og:(g + 1) * Og]
            if Kp != K:
                wpad = torch.zeros((Og, Kp), dtype=torch.bfloat16, device=x.device)
                wpad[:, :K] = wg
                wg = wpad
            bg = b32[g * Og:(g + 1) * Og] if b32 is not None else None
            if direct:
                a = x2[:, g * Cg:(g + 1) * Cg]
            else:
                # The only convolutions left on this path are the RGB stems (C < 16) and grouped ones. The stem's patch
                # matrix is identical in fprop and wgrad, so when a backward will follow it is kept (a dedicated
                # buffer instead of the shared workspace) and the second 0.6 ms im2col pass is skipped.
                keep = groups == 1 and ctx.needs_input_grad[1] and M * Kp * 2 <= (4 << 30)
                a = (torch.empty(M * Kp, dtype=torch.bfloat16, device=x.device) if keep
                     else workspace(M * Kp, x.device)).view(M, Kp)
                xg = xn[..., g * Cg:(g + 1) * Cg]
                _lib.check(lib.bdl_im2col_nhwc_bf16(xg.data_ptr(), a.data_ptr(), N, H, W, Cg, Ho, Wo, kh, kw, sh, sw,
                                                    pt, pl, dh, dw, ldx, Kp, _lib.stream_ptr()), "bdl_im2col_nhwc_bf16")
                _lib.count_launch()
                if keep:
                    kept_cols = a
            gemm_bf16(a, wg, bias=bg, relu=relu, out=y2[:, g * Og:(g + 1) * Og], residual=res2d if direct else None,
                      stats=stats if direct else None)
        ctx.kept_cols = kept_cols
        ctx.save_for_backward(xn, w4, y if relu else None)
        ctx.params = (weight, bias, geom, relu, scale_w, scale_b,
                      (N, H, W, C, Ho, Wo, O, K, Kp, direct, ldx, ldy, c_real, implicit), x.shape)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        if ctx.s2d is not None:
            return tuple(_stem_s2d_backward(ctx, dy)) + (None, None, None)
        xn, w4, y = ctx.saved_tensors
        weight, bias, geom, relu, scale_w, scale_b, dims, xshape = ctx.params
        kh, kw, sh, sw, pt, pl, pb, pr, dh, dw, groups = geom
        N, H, W, C, Ho, Wo, O, K, Kp, direct, ldx, ldy, c_real, implicit = dims
        Cg, Og = C // groups, O // groups
        M = N * Ho * Wo
        lib = _lib.cuda_lib()
        dy4, ldd = nhwc_pitched(dy)
        dyn = rows2d(dy4, ldd)
        gb = _acc_grad(bias, scale_b)
        pre = _PREMASKED.pop((dy.data_ptr(), tuple(dy.shape)), None) if (relu and _PREMASKED) else None
        if pre is not None:
            # the Concat above (or the consuming convolution's dgrad epilogue) already masked this gradient with our
            # ReLU output and, unless it says otherwise, summed its columns
            dz = dyn
            if gb is not None and pre is _NEED_COLSUM:
                if scale_b == 1.0:
                    ew.colsum(dyn, gb)
                else:
                    gb.add_(ew.colsum(dyn, torch.zeros(O, device=dy.device)), alpha=scale_b)
            elif gb is not None and pre is not _BIAS_DONE:
                pre.record_stream(torch.cuda.current_stream(dy.device))
                gb.add_(pre, alpha=scale_b)
        elif relu:
            dz = torch.empty((M, O), dtype=torch.bfloat16, device=dy.device)
            gtmp = gb if (gb is not None and scale_b == 1.0) else torch.zeros(O, device=dy.device)
            ew.colsum(dyn, gtmp, rows2d(y, ldy), dz)
            if gb is not None and scale_b != 1.0:
                gb.add_(gtmp, alpha=scale_b)
        else:
            dz = dyn
            if gb is not None:
                if scale_b == 1.0:
                    ew.colsum(dyn, gb)
                else:
                    gb.add_(ew.colsum(dyn, torch.zeros(O, device=dy.device)), alpha=scale_b)
        x2 = rows2d(xn, ldx)
        w2 = w4.reshape(O, kh * kw * Cg)
        gw = _acc_grad(weight, scale_w)
        padded_c = c_real != C and groups == 1 and C in (4, 8) and c_real < C
        gw2 = None
        if gw is not None and not padded_c:
            gw2 = gw.permute(0, 2, 3, 1).reshape(O, kh * kw * Cg)
        need_dx = ctx.needs_input_grad[0]
        dxn = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=dy.device) if need_dx else None
        done_w = done_x = False
        # residual block input: the other consumer's gradient (if it ran first) is added in the dgrad epilogue
        token = ctx.skip_token if (need_dx and ctx.skip_token is not None and ctx.skip_token.paired) else None
        skip = token.grad if token is not None else None
        res2d = None
        if skip is not None:
            token.grad = None
            if c_real == C and tuple(xshape) == (N, C, H, W) and tuple(skip.shape) == tuple(xshape):
                skip_n, lds = nhwc_pitched(skip)
                res2d = rows2d(skip_n, lds)
        fused_skip = False
        masked_dx = False
        # ReLU backward of the producing conv + its bias-gradient column sums, folded into this dgrad's epilogue: the
        # mask operand is our own saved input (x > 0 <=> the producer's ReLU was active)
        mask2d = colsum_stats = None
        if (need_dx and ctx.mask_input and res2d is None and skip is None and groups == 1 and c_real == C
                and tuple(xshape) == (N, C, H, W) and _RELU_DGRAD[0]):
            mask2d = rows2d(xn, ldx)
            colsum_stats = zeros_f32(2 * C, dy.device)
        if implicit:
            ldz = dz.stride(0)
            if gw is not None and not padded_c and C % 8 == 0 and C >= 16 and scale_w == 1.0 and gw2.stride(1) == 1:
                side = _wgrad_side(dz.device)
                if side is not None:
                    side.wait_stream(torch.cuda.current_stream(dz.device))
                    with torch.cuda.stream(side):
                        conv_igemm(2, xn, ldx, dz, gw2, O, Ho, Wo, kh, kw, sh, sw, pt, pl, dh, dw, accumulate=True)
                    _WGRAD_KEEP.append((dz, xn))
                else:
                    conv_igemm(2, xn, ldx, dz, gw2, O, Ho, Wo, kh, kw, sh, sw, pt, pl, dh, dw, accumulate=True)
                done_w = True
            if need_dx and sh == 1 and sw == 1 and dh * (kh - 1) - pt >= 0 and dw * (kw - 1) - pl >= 0:
                dz_nhwc = torch.as_strided(dz, (N, Ho, Wo, O), (Ho * Wo * ldz, Wo * ldz, ldz, 1), dz.storage_offset())
                conv_igemm(1, dz_nhwc, ldz, w2, dxn.view(N * H * W, C), C, H, W, kh, kw, 1, 1, dh * (kh - 1) - pt,
                           dw * (kw - 1) - pl, dh, dw, residual=res2d if mask2d is None else mask2d,
                           res_mask=mask2d is not None, stats=colsum_stats)
                done_x = True
                fused_skip = res2d is not None
                masked_dx = mask2d is not None
        for g in range(groups):
            if (gw is None or done_w) and (not need_dx or done_x):
                break
            dzg = dz[:, g * Og:(g + 1) * Og]
            if direct:
                a = x2[:, g * Cg:(g + 1) * Cg]
            elif gw is not None and not done_w and ctx.kept_cols is not None:
                a = ctx.kept_cols                      # patch matrix kept by forward
            elif gw is not None and not done_w:
                a = workspace(M * Kp, dy.device).view(M, Kp)
                xg = xn[..., g * Cg:(g + 1) * Cg]
                _lib.check(lib.bdl_im2col_nhwc_bf16(xg.data_ptr(), a.data_ptr(), N, H, W, Cg, Ho, Wo, kh, kw, sh, sw,
                                                    pt, pl, dh, dw, ldx, Kp, _lib.stream_ptr()), "bdl_im2col_nhwc_bf16")
                _lib.count_launch()
            if gw is not None and not done_w:
                if padded_c:
                    tmp = torch.zeros((O, Kp), device=dy.device)
                    gemm_bf16(dzg, a, a_mn=True, b_mn=True, out=tmp, accumulate=True)
                    gw.permute(0, 2, 3, 1).add_(tmp[:, :K].reshape(O, kh, kw, C)[..., :c_real], alpha=scale_w)
                else:
                    tgt = gw2[g * Og:(g + 1) * Og]
                    if scale_w == 1.0 and Kp == K and tgt.stride(1) == 1:
                        side = _wgrad_side(dz.device) if (direct or ctx.kept_cols is not None) else None
                        if side is not None:          # operands are not in the shared workspace: safe to defer
                            side.wait_stream(torch.cuda.current_stream(dz.device))
                            with torch.cuda.stream(side):
                                gemm_bf16(dzg, a, a_mn=True, b_mn=True, out=tgt, accumulate=True)
                            _WGRAD_KEEP.append((dz, a, xn))
                        else:
                            gemm_bf16(dzg, a, a_mn=True, b_mn=True, out=tgt, accumulate=True)
                    else:
                        tmp = torch.zeros((Og, Kp), device=dy.device)
                        gemm_bf16(dzg, a, a_mn=True, b_mn=True, out=tmp, accumulate=True)
                        tgt.add_(tmp[:, :K], alpha=scale_w)
            if need_dx and not done_x:
                wg = w2[g * Og:(g + 1) * Og]
                if direct:
                    gemm_bf16(dzg, wg, b_mn=True, out=dxn.view(N * H * W, C)[:, g * Cg:(g + 1) * Cg],
                              residual=(res2d if mask2d is None else mask2d) if groups == 1 else None,
                              res_mask=mask2d is not None, stats=colsum_stats)
                    fused_skip = res2d is not None and groups == 1
                    masked_dx = mask2d is not None
                else:
                    if Kp != K:
                        wpad = torch.zeros((Og, Kp), dtype=torch.bfloat16, device=dy.device)
                        wpad[:, :K] = wg
                        wg = wpad
                    dcol = workspace(M * Kp, dy.device).view(M, Kp)
                    gemm_bf16(dzg, wg, b_mn=True, out=dcol)
                    dxg = dxn[..., g * Cg:(g + 1) * Cg]
                    _lib.check(lib.bdl_col2im_nhwc_bf16(dcol.data_ptr(), dxg.data_ptr(), N, H, W, Cg, Ho, Wo, kh, kw,
                                                        sh, sw, pt, pl, dh, dw, C, Kp, _lib.stream_ptr()),
                               "bdl_col2im_nhwc_bf16")
                    _lib.count_launch()
        dx = None
        if need_dx:
            dx = dxn.permute(0, 3, 1, 2)
            if dx.shape[1] != xshape[1]:
                dx = dx[:, :xshape[1]]
            if tuple(dx.shape) != tuple(xshape):     # asymmetric SAME padding: drop the materialised border
                dx = dx[:, :, :xshape[2], :xshape[3]]
            if masked_dx:
                # tell the producing conv that this gradient is already masked by its ReLU; hand it the column sums when
                # the epilogue produced them (else it runs its own column-sum pass over the masked gradient)
                done = colsum_stats is not None and getattr(colsum_stats, "_bdl_done", False)
                _PREMASKED[(dx.data_ptr(), tuple(dx.shape))] = colsum_stats[:C] if done else _NEED_COLSUM
            if skip is not None and not fused_skip:
                dx = dx + skip                       # a dgrad path without a residual epilogue (strided / grouped)
            elif token is not None and skip is None:
                token.grad = dx                      # first of the pair: the other consumer adds this
                dx = None
        return dx, None, None, None, None, None, None, None, None, None, None


class FanOutFn(Function):
    """Identity that hands the same activation to ``n`` consumers; its backward sums their gradients in ONE kernel
    instead of autograd's chain of n-1 binary adds (an Inception module input feeds four branches: 5 tensor passes
    instead of 9). Gradients that are not dense bf16 in a common layout fall back to torch adds.

    ``producer``: the ConcatViewFn node that produced ``x`` when ``x`` is the fused-ReLU output of a zero-copy Concat.
    Then the sum, that Concat's ReLU mask and its bias-gradient column sums are one kernel; the masked gradient is
    parked on the producer node, which recognises it (by address) and skips its own pass."""

    @staticmethod
    def forward(ctx, x, n, producer=None):
        ctx.n = n
        ctx.producer = producer
        if producer is not None:
            ctx.save_for_backward(x)
        return tuple(x.view_as(x) for _ in range(n))

    @staticmethod
    def backward(ctx, *grads):
        gs = [g for g in grads if g is not None]
        if not gs:
            return None, None, None
        g0 = gs[0]
        same = all(g.dtype == torch.bfloat16 and g.is_cuda and g.shape == g0.shape and g.stride() == g0.stride()
                   and g.data_ptr() % 16 == 0 for g in gs)
        dense = g0.is_contiguous() or (g0.dim() == 4 and g0.is_contiguous(memory_format=torch.channels_last))
        prod = ctx.producer
        if (prod is not None and same and 2 <= len(gs) <= 4 and g0.dim() == 4
                and g0.is_contiguous(memory_format=torch.channels_last) and g0.shape[1] % 8 == 0):
            (x,) = ctx.saved_tensors
            if x.is_contiguous(memory_format=torch.channels_last) and x.dtype == torch.bfloat16:
                N, Ct, H, W = g0.shape
                dz = torch.empty((N, H, W, Ct), dtype=torch.bfloat16, device=g0.device)
                # the producer Concat scatters (and clears) this vector when all its tails qualify — same test as there
                tails = getattr(prod, "tails", None)
                sums = _concat_sums(id(tails[0]), Ct, g0.device) if _tail_bias_grads(tails) is not None else \
                    torch.zeros(Ct, dtype=torch.float32, device=g0.device)
                ew.colsum_sum_relu(gs, x, dz, sums)
                prod._ready = (dz, sums)
                return dz.permute(0, 3, 1, 2), None, None
        if len(gs) == 1:
            return gs[0], None, None
        if same and dense and g0.numel() % 8 == 0 and len(gs) <= 8:
            return ew.addn(gs), None, None
        total = gs[0]
        for g in gs[1:]:
            total = total + g
        return total, None, None


class MaxPoolFn(Function):
    @staticmethod
    def forward(ctx, x, kh, kw, sh, sw, ph, pw, Ho, Wo):
        xn, ldx = nhwc_pitched(x)
        N, H, W, C = xn.shape
        lib = _lib.cuda_lib()
        y = torch.empty((N, Ho, Wo, C), dtype=torch.bfloat16, device=x.device)
        idx = torch.empty((N, Ho, Wo, C), dtype=torch.uint8, device=x.device)
        _lib.check(lib.bdl_maxpool_fwd_nhwc_bf16(xn.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W, C, Ho, Wo, kh,
                                                 kw, sh, sw, ph, pw, ldx, C, _lib.stream_ptr()), "maxpool_fwd")
        _lib.count_launch()
        ctx.save_for_backward(idx)
        ctx.geom = (N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw = ctx.geom
        lib = _lib.cuda_lib()
        dyn, ldy = nhwc_pitched(dy)
        dx = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=dy.device)
        _lib.check(lib.bdl_maxpool_bwd_nhwc_bf16(dyn.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, C, Ho, Wo, kh,
                                                 kw, sh, sw, ph, pw, C, ldy, _lib.stream_ptr()), "maxpool_bwd")
        _lib.count_launch()
        return (dx.permute(0, 3, 1, 2),) + (None,) * 8


class AvgPoolFn(Function):
    @staticmethod
    def forward(ctx, x, kh, kw, sh, sw, ph, pw, Ho, Wo, count_include_pad, divide):
        xn = _nhwc(x)
        N, H, W, C = xn.shape
        lib = _lib.cuda_lib()
        y = torch.empty((N, Ho, Wo, C), dtype=torch.bfloat16, device=x.device)
        _lib.check(lib.bdl_avgpool_fwd_nhwc_bf16(xn.data_ptr(), y.data_ptr(), N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph,
                                                 pw, C, C, int(count_include_pad), int(divide), _lib.stream_ptr()),
                   "avgpool_fwd")
        _lib.count_launch()
        ctx.geom = (N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw, count_include_pad, divide)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph, pw, cip, divide = ctx.geom
        lib = _lib.cuda_lib()
        dyn = _nhwc(dy)
        dx = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=dy.device)
        _lib.check(lib.bdl_avgpool_bwd_nhwc_bf16(dyn.data_ptr(), dx.data_ptr(), N, H, W, C, Ho, Wo, kh, kw, sh, sw, ph,
                                                 pw, C, C, int(cip), int(divide), _lib.stream_ptr()), "avgpool_bwd")
        _lib.count_launch()
        return (dx.permute(0, 3, 1, 2),) + (None,) * 10


class LRNFn(Function):
    @staticmethod
    def forward(ctx, x, size, alpha, beta, k):
        xn = _nhwc(x)
        N, H, W, C = xn.shape
        lib = _lib.cuda_lib()
        y = torch.empty_like(xn)
        _lib.check(lib.bdl_lrn_fwd_nhwc_bf16(xn.data_ptr(), y.data_ptr(), N * H * W, C, C, C, size, alpha, beta, k,
                                             _lib.stream_ptr()), "lrn_fwd")
        _lib.count_launch()
        ctx.save_for_backward(xn)
        ctx.hp = (size, alpha, beta, k)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        (xn,) = ctx.saved_tensors
        size, alpha, beta, k = ctx.hp
        N, H, W, C = xn.shape
        lib = _lib.cuda_lib()
        dyn = _nhwc(dy)
        dx = torch.empty_like(xn)
        _lib.check(lib.bdl_lrn_bwd_nhwc_bf16(xn.data_ptr(), dyn.data_ptr(), dx.data_ptr(), N * H * W, C, C, C, C, size,
                                             alpha, beta, k, _lib.stream_ptr()), "lrn_bwd")
        _lib.count_launch()
        return dx.permute(0, 3, 1, 2), None, None, None, None


def _dense_like(a: torch.Tensor) -> torch.Tensor:
    """Return ``a`` if its memory is dense (contiguous or channels_last), else a contiguous copy."""
    if a.is_contiguous() or (a.dim() == 4 and a.is_contiguous(memory_format=torch.channels_last)):
        return a
    return a.contiguous()


class ReLUFn(Function):
    @staticmethod
    def forward(ctx, x):
        x = _dense_like(x)
        y = ew.eltwise(ew.RELU, x)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        if dy.stride() != y.stride():
            dy = dy.contiguous(memory_format=torch.channels_last) if (
                y.dim() == 4 and not y.is_contiguous()) else dy.contiguous()
        return ew.eltwise(ew.RELU_BWD, dy, y)


class AddFn(Function):
    """out = a + b (optionally ReLU'd) for two tensors with the same dense layout (residual connections)."""

    @staticmethod
    def forward(ctx, a, b, relu):
        a = _dense_like(a)
        if b.stride() != a.stride():
            b = b.contiguous(memory_format=torch.channels_last) if (a.dim() == 4 and not a.is_contiguous()) \
                else b.contiguous()
        y = ew.eltwise(ew.ADD_RELU if relu else ew.ADD, a, b)
        ctx.relu = relu
        if relu:
            ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        if ctx.relu:
            (y,) = ctx.saved_tensors
            if dy.stride() != y.stride():
                dy = dy.contiguous(memory_format=torch.channels_last) if (
                    y.dim() == 4 and not y.is_contiguous()) else dy.contiguous()
            dz = ew.eltwise(ew.RELU_BWD, dy, y)
            return dz, dz, None
        return dy, dy, None


_dropout_offset = [0]


class DropoutFn(Function):
    @staticmethod
    def forward(ctx, x, p, seed):
        x = _dense_like(x)
        y, mask = ew.dropout_fwd(x, p, seed)          # stream position lives on the device (graph-replay safe)
        ctx.save_for_backward(mask)
        ctx.p = p
        ctx.xstride = x.stride()
        return y

    @staticmethod
    def backward(ctx, dy):
        (mask,) = ctx.saved_tensors
        if dy.stride() != ctx.xstride:
            dy = dy.contiguous(memory_format=torch.channels_last) if len(ctx.xstride) == 4 and ctx.xstride[1] == 1 \
                else dy.contiguous()
        return ew.dropout_bwd(dy, mask, ctx.p), None, None


class SoftmaxFn(Function):
    """SoftMax / LogSoftMax over the last dimension on the native row kernel (csrc/loss_softmax.cu); any leading shape.
    ``out_dtype``: dtype of the result (LogSoftMax returns fp32 like the reference's float tensors)."""

    @staticmethod
    def forward(ctx, x, log_form, out_dtype):
        lib = _lib.cuda_lib()
        C = x.shape[-1]
        x2 = x.reshape(-1, C)
        if x2.stride(-1) != 1:
            x2 = x2.contiguous()
        if x2.dtype not in (torch.bfloat16, torch.float32):
            x2 = x2.float()
        y = torch.empty(x2.shape, dtype=out_dtype, device=x.device)
        _lib.check(lib.bdl_softmax_fwd(x2.data_ptr(), y.data_ptr(), x2.shape[0], C, x2.stride(0), y.stride(0),
                                       int(x2.dtype == torch.bfloat16), int(out_dtype == torch.bfloat16), int(log_form),
                                       _lib.stream_ptr()), "bdl_softmax_fwd")
        _lib.count_launch()
        ctx.save_for_backward(y)
        ctx.cfg = (log_form, x.shape, x.dtype)
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        log_form, shape, in_dtype = ctx.cfg
        lib = _lib.cuda_lib()
        C = y.shape[-1]
        d2 = dy.reshape(-1, C)
        if d2.dtype not in (torch.bfloat16, torch.float32):
            d2 = d2.float()
        d2 = d2.contiguous()
        dx = torch.empty_like(d2)
        _lib.check(lib.bdl_softmax_bwd(y.data_ptr(), d2.data_ptr(), dx.data_ptr(), y.shape[0], C, y.stride(0),
                                       d2.stride(0), dx.stride(0), int(y.dtype == torch.bfloat16),
                                       int(d2.dtype == torch.bfloat16), int(log_form), _lib.stream_ptr()), "bdl_softmax_bwd")
        _lib.count_launch()
        return dx.reshape(shape).to(in_dtype), None, None


class XentFn(Function):
    """CrossEntropyCriterion = LogSoftMax + ClassNLL, fused; returns the scalar loss (fp32)."""

    @staticmethod
    def forward(ctx, logits, target, class_w, size_average, padding_value):
        lib = _lib.cuda_lib()
        lg = logits if logits.stride(-1) == 1 else logits.contiguous()
        B, C = lg.shape
        tgt = target.reshape(-1)
        if tgt.dtype not in (torch.float32, torch.int64):
            tgt = tgt.float()
        tgt = tgt.contiguous()
        scratch = torch.empty(3 * B, dtype=torch.float32, device=lg.device)
        out = torch.empty(2, dtype=torch.float32, device=lg.device)
        _lib.check(lib.bdl_xent_fwd(lg.data_ptr(), tgt.data_ptr(), _lib.ptr(class_w), scratch.data_ptr(),
                                    out.data_ptr(), B, C, lg.stride(0), int(lg.dtype == torch.bfloat16),
                                    int(tgt.dtype == torch.int64), padding_value, int(size_average),
                                    _lib.stream_ptr()), "bdl_xent_fwd")
        _lib.count_launch(2)
        ctx.save_for_backward(lg, tgt, scratch, out)
        ctx.cfg = (class_w, size_average, padding_value)
        return out[0]

    @staticmethod
    def backward(ctx, gloss):
        lg, tgt, scratch, out = ctx.saved_tensors
        class_w, size_average, padding_value = ctx.cfg
        lib = _lib.cuda_lib()
        B, C = lg.shape
        d = torch.empty((B, C), dtype=lg.dtype, device=lg.device)
        # gloss is almost always the constant 1; keep it on the device-free fast path
        g = 1.0 if not isinstance(gloss, torch.Tensor) else None
        if g is None:
            g = 1.0
            scale_t = gloss
        else:
            scale_t = None
        _lib.check(lib.bdl_xent_bwd(lg.data_ptr(), tgt.data_ptr(), _lib.ptr(class_w), scratch.data_ptr(),
                                    out.data_ptr(), d.data_ptr(), B, C, lg.stride(0), d.stride(0),
                                    int(lg.dtype == torch.bfloat16), int(tgt.dtype == torch.int64), padding_value,
                                    int(size_average), g, _lib.stream_ptr()), "bdl_xent_bwd")
        _lib.count_launch()
        if scale_t is not None and not _is_one(scale_t):
            d = d * scale_t.to(d.dtype)
        return d, None, None, None, None


def _is_one(t: torch.Tensor) -> bool:
    # backward seeds created by the training loops are python-constructed ones; avoid a device sync for them
    return getattr(t, "_bdl_is_one", False)


def topk_correct(logits: torch.Tensor, target: torch.Tensor, counts: torch.Tensor | None = None) -> torch.Tensor:
    """counts[0] += top-1 hits, counts[1] += top-5 hits (uint32, device resident)."""
    lib = _lib.cuda_lib()
    lg = logits if logits.stride(-1) == 1 else logits.contiguous()
    tgt = target.reshape(-1)
    if tgt.dtype not in (torch.float32, torch.int64):
        tgt = tgt.float()
    if counts is None:
        counts = torch.zeros(2, dtype=torch.int32, device=lg.device)
    B, C = lg.shape
    _lib.check(lib.bdl_topk_correct(lg.data_ptr(), tgt.contiguous().data_ptr(), counts.data_ptr(), B, C, lg.stride(0),
                                    int(lg.dtype == torch.bfloat16), int(tgt.dtype == torch.int64), _lib.stream_ptr()),
               "bdl_topk_correct")
    _lib.count_launch()
    return counts
