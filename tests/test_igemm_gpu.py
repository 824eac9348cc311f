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
