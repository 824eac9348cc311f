"""BatchNorm + shortcut add + ReLU fused into one forward kernel and one backward kernel pair (csrc/batchnorm.cu
bn_apply_kernel<true>, bn_bwd_apply_kernel<true>) against the three separate layers and against fp32 torch."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-9))


@pytest.mark.parametrize("relu", [True, False])
@pytest.mark.parametrize("shape", [(4, 64, 14, 14), (2, 256, 7, 9)])
def test_function_matches_fp32(relu, shape):
    from bigdl_b200.ops.batchnorm import BatchNormFn
    torch.manual_seed(0)
    N, C, H, W = shape
    x = torch.randn(N, H, W, C, device=DEV).bfloat16().permute(0, 3, 1, 2).requires_grad_(True)
    r = torch.randn(N, H, W, C, device=DEV).bfloat16().permute(0, 3, 1, 2).requires_grad_(True)
    w = (torch.rand(C, device=DEV) + 0.5).requires_grad_(True)
    b = torch.randn(C, device=DEV).requires_grad_(True)
    w.grad, b.grad = torch.zeros_like(w), torch.zeros_like(b)
    rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    y = BatchNormFn.apply(x, w, b, rm, rv, True, 0.1, 1e-5, relu, None, r)
    go = torch.randn_like(y)
    y.backward(go)
    xf, rf = x.detach().float().requires_grad_(True), r.detach().float().requires_grad_(True)
    wf, bf = w.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    ref = F.batch_norm(xf, None, None, wf, bf, True, 0.1, 1e-5) + rf
    ref = F.relu(ref) if relu else ref
    ref.backward(go.float())
    assert _rel(y, ref) < 1e-2
    assert _rel(x.grad, xf.grad) < 2e-2 and _rel(r.grad, rf.grad) < 1e-2
    assert _rel(w.grad, wf.grad) < 2e-2 and _rel(b.grad, bf.grad) < 2e-2


def _block(cin, mid):
    from bigdl_b200.nn import layer as L
    main = (L.Sequential().add(L.SpatialConvolution(cin, mid, 1, 1)).add(L.SpatialBatchNormalization(mid)).add(L.ReLU())
            .add(L.SpatialConvolution(mid, mid, 3, 3, 1, 1, 1, 1)).add(L.SpatialBatchNormalization(mid)).add(L.ReLU())
            .add(L.SpatialConvolution(mid, cin, 1, 1)).add(L.SpatialBatchNormalization(cin)))
    return L.Sequential().add(L.ConcatTable().add(main).add(L.Identity())).add(L.CAddTable()).add(L.ReLU())


def test_residual_block_fused_equals_unfused():
    from bigdl_b200.nn import fusion
    from bigdl_b200.nn import layer as L
    from bigdl_b200.ops import _lib
    from bigdl_b200.utils import config
    torch.manual_seed(1)
    model = L.Sequential().add(_block(64, 16)).add(_block(64, 16)).cuda().training()
    x = torch.randn(8, 64, 12, 12, device=DEV).bfloat16()
    go = torch.randn(8, 64, 12, 12, device=DEV).bfloat16()
    results = {}
    for mode in (False, True):
        config.set("bigdl.fusion.bnadd", mode)
        fusion.unfuse(model)
        fusion.fuse(model)
        w, g = model.getParameters()
        g.zero_()
        before = _lib.LAUNCHES
        xin = x.clone().requires_grad_(True)
        with torch.enable_grad():
            y = model._call(xin)
        torch.autograd.backward(y, go)
        results[mode] = (y.detach().float(), xin.grad.float(), g.clone().float(), _lib.LAUNCHES - before)
    config.set("bigdl.fusion.bnadd", True)
    (y0, dx0, g0, n0), (y1, dx1, g1, n1) = results[False], results[True]
    assert _rel(y1, y0) < 1e-2 and _rel(dx1, dx0) < 2e-2 and _rel(g1, g0) < 2e-2
    assert n1 < n0                                       # two add+ReLU forward and two backward launches are gone


def _proj_block(cin, mid, cout):
    from bigdl_b200.nn import layer as L
    main = (L.Sequential().add(L.SpatialConvolution(cin, mid, 1, 1)).add(L.SpatialBatchNormalization(mid)).add(L.ReLU())
            .add(L.SpatialConvolution(mid, mid, 3, 3, 2, 2, 1, 1)).add(L.SpatialBatchNormalization(mid)).add(L.ReLU())
            .add(L.SpatialConvolution(mid, cout, 1, 1)).add(L.SpatialBatchNormalization(cout)))
    short = L.Sequential().add(L.SpatialConvolution(cin, cout, 1, 1, 2, 2)).add(L.SpatialBatchNormalization(cout))
    return L.Sequential().add(L.ConcatTable().add(main).add(short)).add(L.CAddTable()).add(L.ReLU())


def test_skip_gradient_pairing_matches_autograd_sum():
    """Identity and projection shortcuts: the block input's two gradients are summed in a dgrad epilogue (SkipToken)
    instead of by autograd; results must equal the unpaired run, with fewer at:: add kernels on the tape."""
    from bigdl_b200.nn import fusion
    from bigdl_b200.nn import layer as L
    from bigdl_b200.utils import config
    torch.manual_seed(2)
    model = (L.Sequential().add(L.SpatialConvolution(16, 64, 3, 3, 1, 1, 1, 1)).add(_block(64, 16))
             .add(_proj_block(64, 32, 128)).add(_block(128, 32))).cuda().training()
    x = torch.randn(4, 16, 12, 12, device=DEV).bfloat16()
    go = torch.randn(4, 128, 6, 6, device=DEV).bfloat16()
    res = {}
    for mode in (False, True):
        config.set("bigdl.fusion.skipgrad", mode)
        fusion.unfuse(model)
        fusion.fuse(model)
        pairs = [m.__dict__.get("_skip_pair") for m in model.all_modules() if "_skip_pair" in m.__dict__]
        assert (len(pairs) == 3) == mode
        w, g = model.getParameters()
        g.zero_()
        xin = x.clone().requires_grad_(True)
        with torch.enable_grad():
            y = model._call(xin)
        torch.autograd.backward(y, go)
        res[mode] = (y.detach().float(), xin.grad.float(), g.clone().float())
    config.set("bigdl.fusion.skipgrad", True)
    for a, b in zip(res[True], res[False]):
        assert _rel(a, b) < 2e-2


@pytest.mark.parametrize("affine", [True, False])
def test_bn_relu_backward_recomputes_the_mask(affine):
    """BN + fused ReLU: backward derives the ReLU mask from x (no read of the output); must equal fp32 autograd."""
    from bigdl_b200.ops.batchnorm import BatchNormFn
    torch.manual_seed(4)
    N, C, H, W = 6, 96, 10, 11
    x = torch.randn(N, H, W, C, device=DEV).bfloat16().permute(0, 3, 1, 2).requires_grad_(True)
    w = (torch.rand(C, device=DEV) + 0.5).requires_grad_(True) if affine else None
    b = torch.randn(C, device=DEV).requires_grad_(True) if affine else None
    if affine:
        w.grad, b.grad = torch.zeros_like(w), torch.zeros_like(b)
    rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    y = BatchNormFn.apply(x, w, b, rm, rv, True, 0.1, 1e-5, True, None)
    go = torch.randn_like(y)
    y.backward(go)
    xf = x.detach().float().requires_grad_(True)
    wf = w.detach().clone().requires_grad_(True) if affine else None
    bf = b.detach().clone().requires_grad_(True) if affine else None
    ref = F.relu(F.batch_norm(xf, None, None, wf, bf, True, 0.1, 1e-5))
    # the kernel masks with ITS y (bf16 of the fp32 result): use the same mask in the reference backward
    mask = (y.detach().float() > 0).float()
    pre = F.batch_norm(xf, None, None, wf, bf, True, 0.1, 1e-5)
    (pre * mask).backward(go.float())
    assert _rel(y, ref) < 1e-2 and _rel(x.grad, xf.grad) < 2e-2
    if affine:
        assert _rel(w.grad, wf.grad) < 2e-2 and _rel(b.grad, bf.grad) < 2e-2


@pytest.mark.parametrize("M,N,K", [(4096, 64, 192), (5000, 256, 64), (3000, 512, 128), (130, 96, 64)])
def test_gemm_epilogue_column_statistics(M, N, K):
    """bdl_gemm_set_stats: the TMA-store epilogue adds per-column sum and sum of squares of the bf16 output."""
    from bigdl_b200.ops.gemm import gemm_bf16
    torch.manual_seed(5)
    a = torch.randn(M, K, device=DEV).bfloat16()
    b = (torch.randn(N, K, device=DEV) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device=DEV)
    stats = torch.zeros(2 * N, device=DEV)
    out = gemm_bf16(a, b, bias=bias, stats=stats)
    assert stats._bdl_done
    o = out.float()
    assert torch.allclose(stats[:N], o.sum(0), rtol=2e-3, atol=2e-2)
    assert torch.allclose(stats[N:], (o * o).sum(0), rtol=2e-3, atol=2e-2)
    # a launch that cannot stage its output (fp32 result) reports that it did not produce them
    st2 = torch.zeros(2 * N, device=DEV)
    gemm_bf16(a, b, out_dtype=torch.float32, stats=st2)
    assert not st2._bdl_done and float(st2.abs().sum()) == 0.0


def test_conv_bn_statistics_from_the_epilogue_match_the_separate_pass():
    from bigdl_b200.nn import fusion
    from bigdl_b200.nn import layer as L
    from bigdl_b200.ops import _lib
    from bigdl_b200.utils import config
    torch.manual_seed(6)
    model = (L.Sequential().add(L.SpatialConvolution(16, 64, 3, 3, 1, 1, 1, 1)).add(L.SpatialBatchNormalization(64))
             .add(L.ReLU()).add(L.SpatialConvolution(64, 128, 1, 1)).add(L.SpatialBatchNormalization(128)).add(L.ReLU())
             .add(_block(128, 32))).cuda().training()
    x = torch.randn(8, 16, 20, 20, device=DEV).bfloat16()
    go = torch.randn(8, 128, 20, 20, device=DEV).bfloat16()
    res = {}
    for mode in (False, True):
        config.set("bigdl.fusion.bnstats", mode)
        fusion.unfuse(model)
        fusion.fuse(model)
        for m in model.all_modules():                       # same running statistics at the start of both runs
            if "runningMean" in getattr(m, "_buffers", {}):
                m._buffers["runningMean"].zero_()
                m._buffers["runningVar"].fill_(1.0)
        w, g = model.getParameters()
        g.zero_()
        before = _lib.LAUNCHES
        xin = x.clone().requires_grad_(True)
        with torch.enable_grad():
            y = model._call(xin)
        torch.autograd.backward(y, go)
        rms = [m._buffers["runningMean"].clone() for m in model.all_modules() if "runningMean" in getattr(m, "_buffers", {})]
        res[mode] = (y.detach().float(), xin.grad.float(), g.clone().float(), rms, _lib.LAUNCHES - before)
    config.set("bigdl.fusion.bnstats", True)
    a, b = res[True], res[False]
    assert _rel(a[0], b[0]) < 5e-3 and _rel(a[1], b[1]) < 2e-2 and _rel(a[2], b[2]) < 2e-2
    for r1, r0 in zip(a[3], b[3]):
        assert torch.allclose(r1, r0, rtol=1e-3, atol=1e-4)
