"""Implicit-GEMM convolution (TMA im2col feeding tcgen05) vs torch conv2d in fp32: fprop, dgrad, wgrad."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _close(a, b, rtol, atol):
    err = (a.float() - b.float()).abs()
    assert bool((err <= atol + rtol * b.float().abs()).all()), f"max err {err.max().item():.4g}, ref max {b.abs().max().item():.4g}"


CASES = [
    # N, C, H, W, O, kh, kw, sh, sw, ph, pw, dh, dw
    (4, 64, 16, 16, 128, 3, 3, 1, 1, 1, 1, 1, 1),
    (2, 128, 14, 14, 96, 1, 1, 1, 1, 0, 0, 1, 1),
    (2, 64, 13, 17, 64, 3, 5, 1, 1, 1, 2, 1, 1),      # non-square kernel/pad: settles the W/H order of the box corners
    (3, 96, 12, 12, 80, 3, 3, 1, 1, 1, 1, 1, 1),      # C not a multiple of 64 (zero-filled channel tail)
    (2, 16, 10, 10, 32, 5, 5, 1, 1, 2, 2, 1, 1),
    (2, 64, 15, 15, 64, 3, 3, 2, 2, 1, 1, 1, 1),      # strided traversal
    (2, 64, 12, 12, 64, 3, 3, 1, 1, 2, 2, 2, 2),      # dilation
    (2, 32, 14, 14, 64, 5, 5, 1, 1, 2, 2, 1, 1),      # C <= 32: 32-channel atoms, 64B swizzle, 2 taps per k block
    (2, 24, 14, 14, 64, 5, 5, 1, 1, 2, 2, 1, 1),      # C = 24 inside a 32-channel atom (zero-filled tail)
    (3, 16, 19, 19, 48, 4, 4, 1, 1, 0, 0, 1, 1),      # 16-channel atoms, 32B swizzle, 4 taps per k block (s2d stem)
    (2, 8, 12, 12, 32, 3, 3, 1, 1, 1, 1, 1, 1),       # C = 8 inside a 16-channel atom; 9 taps = 2 k blocks + 1 tap
    (2, 32, 15, 15, 40, 3, 3, 2, 2, 1, 1, 1, 1),      # strided, 64B swizzle
]


@pytest.mark.parametrize("case", CASES)
def test_igemm_fprop(case):
    from bigdl_b200.ops.gemm import conv_igemm
    N, C, H, W, O, kh, kw, sh, sw, ph, pw, dh, dw = case
    torch.manual_seed(0)
    x = torch.randn(N, H, W, C, device="cuda").bfloat16()
    w = (torch.randn(O, kh, kw, C, device="cuda") * 0.1).bfloat16()
    b = torch.randn(O, device="cuda")
    Ho = (H + 2 * ph - dh * (kh - 1) - 1) // sh + 1
    Wo = (W + 2 * pw - dw * (kw - 1) - 1) // sw + 1
    y = torch.empty(N * Ho * Wo, O, device="cuda", dtype=torch.bfloat16)
    conv_igemm(0, x, C, w.reshape(O, -1), y, O, Ho, Wo, kh, kw, sh, sw, ph, pw, dh, dw, bias=b, relu=True)
    torch.cuda.synchronize()
    ref = F.relu(F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), b, (sh, sw), (ph, pw), (dh, dw)))
    _close(y.reshape(N, Ho, Wo, O).permute(0, 3, 1, 2), ref, 2e-2, 3e-2)


@pytest.mark.parametrize("case", [c for c in CASES if c[7] == 1 and c[8] == 1])
def test_igemm_dgrad(case):
    from bigdl_b200.ops.gemm import conv_igemm
    N, C, H, W, O, kh, kw, sh, sw, ph, pw, dh, dw = case
    torch.manual_seed(1)
    Ho = H + 2 * ph - dh * (kh - 1)
    Wo = W + 2 * pw - dw * (kw - 1)
    dy = torch.randn(N, Ho, Wo, O, device="cuda").bfloat16()
    w = (torch.randn(O, kh, kw, C, device="cuda") * 0.1).bfloat16()
    dx = torch.empty(N * H * W, C, device="cuda", dtype=torch.bfloat16)
    # transposed-conv geometry: gather from dY with pad' = dil*(k-1) - pad, flipped taps, output size = input size
    conv_igemm(1, dy, O, w.reshape(O, -1), dx, C, H, W, kh, kw, 1, 1, dh * (kh - 1) - ph, dw * (kw - 1) - pw, dh, dw)
    torch.cuda.synchronize()
    xr = torch.zeros(N, C, H, W, device="cuda", requires_grad=True)
    F.conv2d(xr, w.float().permute(0, 3, 1, 2), None, 1, (ph, pw), (dh, dw)).backward(dy.float().permute(0, 3, 1, 2))
    _close(dx.reshape(N, H, W, C).permute(0, 3, 1, 2), xr.grad, 3e-2, 5e-2)


WGRAD_EXTRA = [
    (2, 24, 9, 9, 48, 5, 5, 1, 1, 2, 2, 1, 1),        # C < 64: one partly-filled channel atom per tap
    (2, 160, 7, 7, 224, 3, 3, 1, 1, 1, 1, 1, 1),      # C = 2.5 atoms, O not a multiple of 128
    (2, 72, 8, 8, 40, 3, 3, 2, 2, 1, 1, 1, 1),        # C % 16 == 8: a 16-column chunk is only half valid
    (8, 64, 56, 56, 192, 3, 3, 1, 1, 1, 1, 1, 1),     # O > 128 with a long pixel axis: two stacked m tiles per CTA
    (8, 32, 40, 40, 320, 3, 3, 1, 1, 1, 1, 1, 1),     # same with 32-channel atoms and a half-empty last tile pair
]


@pytest.mark.parametrize("case", CASES + WGRAD_EXTRA)
def test_igemm_wgrad(case):
    from bigdl_b200.ops.gemm import conv_igemm
    N, C, H, W, O, kh, kw, sh, sw, ph, pw, dh, dw = case
    torch.manual_seed(2)
    Ho = (H + 2 * ph - dh * (kh - 1) - 1) // sh + 1
    Wo = (W + 2 * pw - dw * (kw - 1) - 1) // sw + 1
    x = torch.randn(N, H, W, C, device="cuda").bfloat16()
    dy = torch.randn(N * Ho * Wo, O, device="cuda").bfloat16()
    gw = torch.ones(O, kh * kw * C, device="cuda")
    conv_igemm(2, x, C, dy, gw, O, Ho, Wo, kh, kw, sh, sw, ph, pw, dh, dw, accumulate=True)
    torch.cuda.synchronize()
    wr = torch.zeros(O, C, kh, kw, device="cuda", requires_grad=True)
    F.conv2d(x.float().permute(0, 3, 1, 2), wr, None, (sh, sw), (ph, pw), (dh, dw)).backward(
        dy.float().reshape(N, Ho, Wo, O).permute(0, 3, 1, 2))
    ref = 1.0 + wr.grad.permute(0, 2, 3, 1).reshape(O, -1)
    _close(gw, ref, 2e-2, 0.3)


@pytest.mark.parametrize("shape", [(4, 3, 64, 64, 64, 7, 3), (2, 3, 56, 40, 32, 5, 2), (2, 1, 28, 28, 16, 3, 1),
                                   (3, 3, 33, 35, 24, 7, 3)])
@pytest.mark.parametrize("fp32_in", [True, False])
def test_stem_space_to_depth_conv(shape, fp32_in):
    """Stride-2 RGB stem through the space-to-depth + stride-1 implicit GEMM: output, weight and bias gradients vs
    fp32 conv2d, and vs the explicit im2col path it replaces."""
    from bigdl_b200.ops import functions as fn
    N, C, H, W, O, k, p = shape
    torch.manual_seed(1)
    x = torch.randn(N, C, H, W, device="cuda")
    if not fp32_in:
        x = x.bfloat16().contiguous(memory_format=torch.channels_last)
    geom = (k, k, 2, 2, p, p, p, p, 1, 1, 1)
    plan = fn._stem_s2d_plan(x, torch.empty(O, C, k, k), geom, False)
    outs = []
    for use in (True, False):
        fn._STEM_S2D[0] = use
        try:
            w = (torch.randn(O, C, k, k, device="cuda", generator=torch.Generator("cuda").manual_seed(2)) * 0.1
                 ).requires_grad_()
            b = torch.zeros(O, device="cuda").requires_grad_()
            w.grad, b.grad = torch.zeros_like(w), torch.zeros_like(b)
            w16 = w.detach().bfloat16().contiguous(memory_format=torch.channels_last)
            y = fn.ConvFn.apply(x, w, b, w16, geom, True, 1.0, 1.0)
            gy = torch.randn(y.shape, device="cuda", generator=torch.Generator("cuda").manual_seed(3)).bfloat16()
            y.backward(gy)
            torch.cuda.synchronize()
            outs.append((y.detach().float(), w.grad.clone(), b.grad.clone()))
        finally:
            fn._STEM_S2D[0] = True
    wr = w16.float().detach().requires_grad_()
    br = torch.zeros(O, device="cuda", requires_grad=True)
    # the kernel rounds x to bf16: give the oracle the same values so the ReLU masks agree
    yr = F.relu(F.conv2d(x.bfloat16().float(), wr, br, 2, p))
    yr.backward(gy.float())
    for name, (y, gw, gb) in zip(("s2d", "im2col"), outs):
        if name == "s2d" and plan is None:
            continue
        for what, a, b, atol in (("y", y, yr, 3e-2), ("gw", gw, wr.grad, 0.02 * wr.grad.abs().max().item()),
                                 ("gb", gb, br.grad, 0.02 * br.grad.abs().max().item())):
            err = (a.float() - b.float()).abs()
            assert bool((err <= atol + 3e-2 * b.abs()).all()), \
                f"{name}/{what}: max err {err.max().item():.4g}, ref max {b.abs().max().item():.4g}"
    if shape[0] == 4:
        assert plan is not None


PAIR_CASES = [
    (8, 64, 16, 16, 128, 3, 3, 1, 1, 1, 1, 1, 1),
    (4, 128, 14, 14, 256, 1, 1, 1, 1, 0, 0, 1, 1),
    (5, 96, 12, 12, 64, 3, 3, 1, 1, 1, 1, 1, 1),      # C tail inside a 64-channel block, ragged last 256-pixel tile
    (4, 64, 15, 15, 128, 3, 3, 2, 2, 1, 1, 1, 1),     # strided traversal
    (3, 192, 17, 17, 320, 3, 3, 1, 1, 1, 1, 1, 1),    # two n tiles (block_n 192 + 128 or 256 + 64 depending on waves)
    (64, 128, 28, 28, 256, 3, 3, 1, 1, 1, 1, 1, 1),   # long reduction, many tiles: 512-pixel pair tiles (two stacked MMAs)
]


@pytest.mark.parametrize("case", PAIR_CASES)
def test_igemm_cta_pair_fprop_dgrad(case):
    """Implicit-GEMM fprop and (stride-1) dgrad on the cta_group::2 kernels (forced) vs fp32 conv2d and vs the 1-CTA
    kernels."""
    from bigdl_b200.ops import _lib
    from bigdl_b200.ops.gemm import conv_igemm
    N, C, H, W, O, kh, kw, sh, sw, ph, pw, dh, dw = case
    lib = _lib.cuda_lib()
    torch.manual_seed(5)
    x = torch.randn(N, H, W, C, device="cuda").bfloat16()
    w = (torch.randn(O, kh, kw, C, device="cuda") * 0.1).bfloat16()
    b = torch.randn(O, device="cuda")
    Ho = (H + 2 * ph - dh * (kh - 1) - 1) // sh + 1
    Wo = (W + 2 * pw - dw * (kw - 1) - 1) // sw + 1
    ref = F.relu(F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), b, (sh, sw), (ph, pw), (dh, dw)))
    old = lib.bdl_gemm_set_pair_mode(1)
    try:
        got = {}
        for mode in (1, 0):
            lib.bdl_gemm_set_pair_mode(mode)
            y = torch.empty(N * Ho * Wo, O, device="cuda", dtype=torch.bfloat16)
            conv_igemm(0, x, C, w.reshape(O, -1), y, O, Ho, Wo, kh, kw, sh, sw, ph, pw, dh, dw, bias=b, relu=True)
            torch.cuda.synchronize()
            _close(y.reshape(N, Ho, Wo, O).permute(0, 3, 1, 2), ref, 2e-2, 3e-2)
            got[mode] = y
        assert torch.equal(got[0], got[1])
        if sh == 1 and sw == 1:
            dy = torch.randn(N, Ho, Wo, O, device="cuda").bfloat16()
            xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
            F.conv2d(xr, w.float().permute(0, 3, 1, 2), None, 1, (ph, pw), (dh, dw)).backward(dy.float().permute(0, 3, 1, 2))
            for mode in (1, 0):
                lib.bdl_gemm_set_pair_mode(mode)
                dx = torch.empty(N * H * W, C, device="cuda", dtype=torch.bfloat16)
                conv_igemm(1, dy, O, w.reshape(O, -1), dx, C, H, W, kh, kw, 1, 1, dh * (kh - 1) - ph, dw * (kw - 1) - pw,
                           dh, dw)
                torch.cuda.synchronize()
                _close(dx.reshape(N, H, W, C).permute(0, 3, 1, 2), xr.grad, 3e-2, 6e-2)
    finally:
        lib.bdl_gemm_set_pair_mode(old)
