"""conv+ReLU -> conv: the first convolution's ReLU backward and bias-gradient column sums folded into the second
convolution's dgrad GEMM epilogue (mask operand + column statistics), against the separate mask / column-sum pass; and
the skip-gradient pairing of a basic (3x3-first) residual block, whose dgrad is the implicit-GEMM path."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-9))


def _run(model, x, go, key, values):
    from bigdl_b200.nn import fusion
    from bigdl_b200.ops import _lib
    from bigdl_b200.utils import config
    out = {}
    for mode in values:
        config.set(key, mode)
        fusion.unfuse(model)
        fusion.fuse(model)
        w, g = model.getParameters()
        g.zero_()
        before = _lib.LAUNCHES
        xin = x.clone().requires_grad_(True)
        with torch.enable_grad():
            y = model._call(xin)
        torch.autograd.backward(y, go)
        torch.cuda.synchronize()
        out[mode] = (y.detach().float(), xin.grad.float(), g.clone().float(), _lib.LAUNCHES - before)
    config.set(key, key != "bigdl.fusion.reludgrad")      # back to the defaults
    fusion.unfuse(model)
    return out


def test_relu_backward_in_the_consumer_dgrad():
    from bigdl_b200.nn import layer as L
    torch.manual_seed(0)
    model = (L.Sequential()
             .add(L.SpatialConvolution(16, 64, 3, 3, 1, 1, 1, 1)).add(L.ReLU())
             .add(L.SpatialConvolution(64, 64, 3, 3, 1, 1, 1, 1)).add(L.ReLU())          # implicit dgrad consumer
             .add(L.SpatialConvolution(64, 128, 1, 1)).add(L.ReLU())                      # direct (1x1) dgrad consumer
             .add(L.SpatialMaxPooling(2, 2, 2, 2))
             .add(L.SpatialConvolution(128, 32, 3, 3, 1, 1, 1, 1))).cuda().training()
    n_marked = 0
    x = torch.randn(8, 16, 24, 24, device=DEV).bfloat16()
    go = torch.randn(8, 32, 12, 12, device=DEV).bfloat16()
    from bigdl_b200.nn import fusion
    from bigdl_b200.utils import config
    config.set("bigdl.fusion.reludgrad", True)
    fusion.fuse(model)
    n_marked = sum(1 for m in model.all_modules() if m.__dict__.get("_mask_input"))
    res = _run(model, x, go, "bigdl.fusion.reludgrad", (False, True))
    assert n_marked == 2                                   # conv2 and conv3; conv4 follows a pooling layer
    (y0, dx0, g0, n0), (y1, dx1, g1, n1) = res[False], res[True]
    assert _rel(y1, y0) < 1e-3 and _rel(dx1, dx0) < 1e-2 and _rel(g1, g0) < 1e-2
    assert n1 <= n0 - 2                                    # two mask + column-sum passes gone


def test_basic_block_skip_gradient_uses_implicit_dgrad_residual():
    from bigdl_b200.nn import layer as L
    torch.manual_seed(1)

    def basic(c):
        main = (L.Sequential().add(L.SpatialConvolution(c, c, 3, 3, 1, 1, 1, 1)).add(L.SpatialBatchNormalization(c))
                .add(L.ReLU()).add(L.SpatialConvolution(c, c, 3, 3, 1, 1, 1, 1)).add(L.SpatialBatchNormalization(c)))
        return L.Sequential().add(L.ConcatTable().add(main).add(L.Identity())).add(L.CAddTable()).add(L.ReLU())
    model = (L.Sequential().add(L.SpatialConvolution(16, 32, 3, 3, 1, 1, 1, 1)).add(basic(32)).add(basic(32))).cuda().training()
    x = torch.randn(4, 16, 16, 16, device=DEV).bfloat16()
    go = torch.randn(4, 32, 16, 16, device=DEV).bfloat16()
    res = _run(model, x, go, "bigdl.fusion.skipgrad", (False, True))
    for a, b in zip(res[True][:3], res[False][:3]):
        assert _rel(a, b) < 2e-2
