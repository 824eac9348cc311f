"""fp8 (e4m3) tcgen05 GEMM / implicit-GEMM conv and the quantised layers, against fp32 references built from the SAME
quantised operands (so the only difference is accumulation order) and against the float model (quantisation error)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _q(x2d, per_row=False):
    from bigdl_b200.ops import quant as Q
    from bigdl_b200.ops.elementwise import absmax

    am = Q.row_absmax(x2d) if per_row else absmax(x2d.contiguous())
    return Q.quantize_e4m3(x2d, am, per_row=per_row), am


@pytest.mark.parametrize("M,N,K", [(128, 128, 128), (256, 512, 1024), (200, 72, 336), (1000, 1000, 2048), (33, 16, 48)])
def test_gemm_fp8_matches_dequantised_reference(M, N, K):
    from bigdl_b200.ops import quant as Q

    torch.manual_seed(0)
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K, device="cuda") * 0.1
    bias = torch.randn(N, device="cuda")
    aq, a_am = _q(a)
    bq, b_am = _q(b, per_row=True)
    assert torch.equal(aq.cpu(), Q.quantize_e4m3(a.cpu(), a_am.cpu()))          # GPU and CPU quantisers agree bit-exactly
    cs = Q.colscale(a_am, b_am)
    out = Q.gemm_fp8(aq, bq, cs, k=K, bias=bias, relu=True, out_dtype=torch.float32)
    ad = Q.dequantize_e4m3(aq, a_am)[:, :K]
    bd = Q.dequantize_e4m3(bq, b_am, per_row=True)[:, :K]
    ref = F.relu(ad.double() @ bd.double().t() + bias.double()).float()
    assert torch.allclose(out, ref, rtol=2e-4, atol=2e-4 * float(ref.abs().max()))
    # and close to the unquantised product
    full = F.relu(a.float() @ b.t() + bias)
    assert float((out - full).norm() / full.norm()) < 0.06
    out16 = Q.gemm_fp8(aq, bq, cs, k=K, bias=bias, relu=True)
    assert torch.allclose(out16.float(), ref, rtol=1e-2, atol=1e-2 * float(ref.abs().max()))


def test_gemm_fp8_requantised_output():
    from bigdl_b200.ops import quant as Q

    torch.manual_seed(1)
    M, N, K = 256, 128, 256
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K, device="cuda") * 0.05
    aq, a_am = _q(a)
    bq, b_am = _q(b, per_row=True)
    cs = Q.colscale(a_am, b_am)
    y = Q.gemm_fp8(aq, bq, cs, k=K, out_dtype=torch.float32)
    y_am = y.abs().max().reshape(1)
    qs = (Q.E4M3_MAX / y_am).float()
    yq = Q.gemm_fp8(aq, bq, cs, k=K, out_dtype=torch.uint8, out_qscale=qs)
    want = Q.quantize_e4m3(y, y_am)
    # identical except where fp32 rounding of v*qs lands on a rounding boundary
    diff = (yq.view(torch.float8_e4m3fn).float() - want.view(torch.float8_e4m3fn).float()).abs()
    assert float((diff > 0).float().mean()) < 0.01
    assert float(diff.max()) <= 32.0


@pytest.mark.parametrize("C,O,k,s,p,d,HW", [(64, 64, 3, 1, 1, 1, 14), (16, 32, 3, 2, 1, 1, 17), (128, 256, 1, 1, 0, 1, 7),
                                            (48, 80, 5, 1, 2, 1, 12), (256, 64, 3, 1, 2, 2, 9), (64, 128, 1, 2, 0, 1, 8),
                                            (3, 64, 7, 2, 3, 1, 32)])
def test_quantized_conv_matches_emulation(C, O, k, s, p, d, HW):
    from bigdl_b200.nn import layer as L
    from bigdl_b200 import quantized as QZ

    torch.manual_seed(2)
    conv = L.SpatialConvolution(C, O, k, k, s, s, p, p, dilation_w=d, dilation_h=d)
    conv._fuse_relu = True
    q = QZ.SpatialConvolution.from_float(conv)
    x = torch.randn(4, C, HW, HW)
    ref = q.forward(x).clone()                           # CPU emulation of the same arithmetic
    qg = q.cloneModule().to("cuda")
    out = qg.forward(x.cuda().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)).float().cpu()
    assert out.shape == ref.shape
    if C >= 16:
        # activations reach the GPU layer as bf16, the CPU path quantises fp32: allow the bf16 input rounding
        assert float((out - ref).norm() / ref.norm()) < 0.02
    else:
        assert "weight_hp" in qg._buffers          # RGB stem stays in bf16 on the GPU
    full = F.relu(F.conv2d(x, conv.weight.detach(), conv.bias.detach(), s, p, d))
    assert float((out - full).norm() / full.norm()) < 0.08


def test_quantized_model_static_chain_gpu():
    from bigdl_b200.nn import layer as L
    from bigdl_b200.quantized import Quantizer
    from bigdl_b200.ops import _lib

    torch.manual_seed(3)
    m = (L.Sequential()
         .add(L.SpatialConvolution(16, 32, 3, 3, 1, 1, 1, 1)).add(L.SpatialBatchNormalization(32)).add(L.ReLU())
         .add(L.SpatialConvolution(32, 64, 1, 1)).add(L.ReLU())
         .add(L.SpatialConvolution(64, 32, 3, 3, 2, 2, 1, 1))
         .add(L.Reshape([32 * 4 * 4])).add(L.Linear(512, 10)))
    m.evaluate()
    x = torch.randn(8, 16, 8, 8)
    ref = m.forward(x).clone()
    q = m.quantize().to("cuda")
    xg = x.cuda().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    dyn = q.forward(xg).float().cpu()
    assert float((dyn - ref).norm() / ref.norm()) < 0.08
    Quantizer.calibrate(q, [xg])
    qs = Quantizer.quantized_layers(q)
    assert qs[0]._emit_q is qs[1] and qs[1]._emit_q is qs[2]
    q.forward(xg)                          # first static call caches the per-layer dequantisation vectors
    n0 = _lib.LAUNCHES
    st = q.forward(xg).float().cpu()
    launches = _lib.LAUNCHES - n0
    assert float((st - ref).norm() / ref.norm()) < 0.1
    # 1 quantise + 3 conv GEMMs (e4m3 handed from epilogue to consumer) + quantise + linear GEMM
    assert launches == 6, launches
    assert q.modules[0].output.dtype == torch.uint8


def test_resnet_block_shared_quantisation():
    from bigdl_b200.models.resnet import ResNet
    from bigdl_b200.quantized import Quantizer

    torch.manual_seed(4)
    m = ResNet(10, depth=20, shortcut_type="B", dataset="cifar10").evaluate()
    x = torch.randn(4, 3, 32, 32)
    ref = m.forward(x).clone()
    q = m.quantize().to("cuda")
    xg = x.cuda().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    Quantizer.calibrate(q, [xg])
    out = q.forward(xg).float().cpu()
    assert torch.isfinite(out).all()
    assert float((out - ref).norm() / ref.norm()) < 0.25


@pytest.mark.parametrize("M,N,K", [(128, 128, 128), (256, 128, 512), (300, 200, 416), (1000, 64, 96), (512, 1000, 2048)])
def test_mxfp8_block_scaled_gemm(M, N, K):
    """tcgen05.mma.kind::mxf8f6f4.block_scale vs an fp32 GEMM of the SAME block-quantised operands (what the tensor core
    must compute exactly, up to fp32 accumulation order), and the quantiser vs its torch restatement."""
    from bigdl_b200.ops import quant as Q
    torch.manual_seed(4)
    # per-row / per-block magnitudes spread over several octaves: one scale per tensor would lose the small blocks
    a = torch.randn(M, K, device="cuda") * torch.exp2(torch.randint(-6, 6, (M, 1), device="cuda").float())
    b = torch.randn(N, K, device="cuda") * torch.exp2(torch.randint(-4, 4, (N, (K + 31) // 32), device="cuda").float()
                                                      ).repeat_interleave(32, dim=1)[:, :K]
    bias = torch.randn(N, device="cuda")
    aq, sfa = Q.quantize_mxfp8(a.bfloat16())
    bq, sfb = Q.quantize_mxfp8(b.bfloat16())
    aq_c, sfa_c = Q.quantize_mxfp8(a.bfloat16().cpu())
    assert torch.equal(aq.cpu(), aq_c) and torch.equal(sfa.cpu(), sfa_c)
    ad, bd = Q.dequantize_mxfp8(aq, sfa, K), Q.dequantize_mxfp8(bq, sfb, K)
    # block scaling keeps ~2 significant digits in EVERY block
    assert float(((ad - a.bfloat16().float()).abs() / (a.abs() + 1e-3 * a.abs().amax(1, keepdim=True))).max()) < 0.13
    ref = ad @ bd.t()
    out = Q.gemm_mxfp8(aq, sfa, bq, sfb, K, out_dtype=torch.float32)
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    assert err <= 1e-4 * ref.abs().max().item() + 1e-5, (err, ref.abs().max().item())
    out16 = Q.gemm_mxfp8(aq, sfa, bq, sfb, K, bias=bias, relu=True)
    torch.cuda.synchronize()
    assert torch.allclose(out16.float(), torch.relu(ref + bias), rtol=1e-2, atol=1e-2 * ref.abs().max().item())
    # against the unquantised product: block scales keep the relative error at the e4m3 level for every row
    true = a.bfloat16().float() @ b.bfloat16().float().t()
    rel = ((out - true).norm(dim=1) / (true.norm(dim=1) + 1e-6)).max().item()
    assert rel < 0.09, rel


def test_top1_agreement_on_gpu_kernels():
    """The trained classifier of the CPU test, quantised and run on the tcgen05 fp8 kernels: per-tensor scales
    (kind::f8f6f4) and MX block scales (kind::mxf8f6f4.block_scale) both keep >= 99.5 % top-1 agreement with the float
    model (reference: T/integration/Quantization.scala:88-148, 'same top-1 accuracy')."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_quantized_cpu import _trained_mlp
    from bigdl_b200.quantized.quantizer import Quantizer

    m, batch = _trained_mlp()
    x, y = batch(4096)
    ref = m.forward(x).float()
    m.cuda()
    xg = x.cuda()
    ref_gpu = m.forward(xg).float().cpu()
    assert float((ref_gpu.argmax(1) == ref.argmax(1)).float().mean()) >= 0.995
    errs = {}
    for mode in (False, True):
        q = Quantizer.quantize(m, block_scale=mode)
        out = q.forward(xg).float().cpu()
        agree = float((out.argmax(1) == ref.argmax(1)).float().mean())
        assert agree >= 0.995, (mode, agree)
        errs[mode] = float((out - ref).norm() / ref.norm())
    assert errs[True] <= errs[False] * 1.05, errs
