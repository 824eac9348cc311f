"""sm_100a kernels vs plain fp32 PyTorch references of the same op (conv fprop/dgrad/wgrad, pooling, LRN, fused
cross-entropy, ReLU/add/dropout, casts, optimizer updates)."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _close(a, b, rtol, atol):
    a, b = a.float(), b.float()
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    assert bool((err <= tol).all()), f"max err {err.max().item():.4g} (ref max {b.abs().max().item():.4g})"


@pytest.mark.parametrize("cfg", [
    # (N, Cin, H, W, Cout, k, stride, pad, groups, dil)
    (4, 64, 28, 28, 96, 1, 1, 0, 1, 1),
    (4, 96, 28, 28, 128, 3, 1, 1, 1, 1),
    (2, 16, 14, 14, 32, 5, 1, 2, 1, 1),
    (2, 3, 64, 64, 64, 7, 2, 3, 1, 1),
    (2, 64, 17, 17, 64, 3, 2, 1, 2, 1),
    (2, 32, 20, 20, 48, 3, 1, 2, 1, 2),
])
def test_conv_fwd_bwd(cfg):
    from bigdl_b200.nn.layers_core import SpatialConvolution
    N, Cin, H, W, Cout, k, s, p, g, d = cfg
    torch.manual_seed(0)
    conv = SpatialConvolution(Cin, Cout, k, k, s, s, p, p, g, dilation_w=d, dilation_h=d)
    ref_w = conv.weight.detach().clone().contiguous()
    ref_b = conv.bias.detach().clone()
    conv.cuda()
    x = torch.randn(N, Cin, H, W, device=DEV)
    xb = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = conv.forward(xb)
    # reference on bf16-rounded inputs/weights in fp32
    xr = xb.float().requires_grad_(True)
    wr = ref_w.to(DEV).to(torch.bfloat16).float().requires_grad_(True)
    br = ref_b.to(DEV).requires_grad_(True)
    yr = F.conv2d(xr, wr, br, s, p, d, g)
    _close(y, yr, 2e-2, 2e-2)
    go = torch.randn_like(yr)
    gx = conv.backward(xb, go.to(torch.bfloat16))
    yr.backward(go.to(torch.bfloat16).float())
    scale = math.sqrt(Cout * k * k / g)
    _close(gx, xr.grad, 3e-2, 3e-2 * scale)
    gw = conv.parameters()[1][0]
    _close(gw, wr.grad, 3e-2, 3e-2 * math.sqrt(N * H * W / (s * s)))
    _close(conv.parameters()[1][1], br.grad, 3e-2, 3e-2 * math.sqrt(N * H * W / (s * s)))


def test_linear_fwd_bwd():
    from bigdl_b200.nn.layers_core import Linear
    torch.manual_seed(1)
    lin = Linear(1024, 1000)
    w, b = lin.weight.detach().clone(), lin.bias.detach().clone()
    lin.cuda()
    x = torch.randn(128, 1024, device=DEV).to(torch.bfloat16)
    y = lin.forward(x)
    xr = x.float().requires_grad_(True)
    wr = w.to(DEV).to(torch.bfloat16).float().requires_grad_(True)
    br = b.to(DEV).requires_grad_(True)
    yr = F.linear(xr, wr, br)
    _close(y, yr, 2e-2, 2e-2)
    go = torch.randn_like(yr).to(torch.bfloat16)
    gx = lin.backward(x, go)
    yr.backward(go.float())
    _close(gx, xr.grad, 3e-2, 0.2)
    _close(lin.parameters()[1][0], wr.grad, 3e-2, 0.3)
    _close(lin.parameters()[1][1], br.grad, 3e-2, 0.3)


@pytest.mark.parametrize("k,s,p,ceil", [(3, 2, 0, True), (3, 1, 1, True), (2, 2, 0, False), (3, 2, 1, False)])
def test_maxpool(k, s, p, ceil):
    from bigdl_b200.nn.layers_core import SpatialMaxPooling
    torch.manual_seed(2)
    pool = SpatialMaxPooling(k, k, s, s, p, p, to_ceil=ceil)
    x = torch.randn(3, 64, 29, 29, device=DEV).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = pool.forward(x)
    xr = x.float().cpu().requires_grad_(True)
    yr = pool._call(xr)            # CPU path of the same module = torch reference
    assert y.shape == yr.shape
    _close(y.cpu(), yr, 0, 0)
    go = torch.randn_like(yr).to(torch.bfloat16)
    gx = pool.backward(x, go.to(DEV))
    yr.backward(go.float())
    # the 3x3 backward accumulates overlapping windows in packed bf16: a few roundings instead of one
    _close(gx.cpu(), xr.grad, 2e-2, 1e-2 * float(xr.grad.abs().max()))


@pytest.mark.parametrize("H", [7, 14, 28, 8])
@pytest.mark.parametrize("s,p,ceil", [(1, 1, False), (2, 0, True), (2, 1, False), (1, 0, False)])
def test_maxpool3_sliding_window(H, s, p, ceil):
    """The strip-walking 3x3 kernels at map sizes that are / are not multiples of the 7-row strip."""
    from bigdl_b200.nn.layers_core import SpatialMaxPooling
    torch.manual_seed(20)
    pool = SpatialMaxPooling(3, 3, s, s, p, p, to_ceil=ceil)
    x = torch.randn(2, 24, H, H + 3, device=DEV).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = pool.forward(x)
    xr = x.float().cpu().requires_grad_(True)
    yr = pool._call(xr)
    assert y.shape == yr.shape
    _close(y.cpu(), yr, 0, 0)
    go = torch.randn_like(yr).to(torch.bfloat16)
    gx = pool.backward(x, go.to(DEV))
    yr.backward(go.float())
    # the 3x3 backward accumulates overlapping windows in packed bf16: a few roundings instead of one
    _close(gx.cpu(), xr.grad, 2e-2, 1e-2 * float(xr.grad.abs().max()))


@pytest.mark.parametrize("k,s,p,ceil,cip", [(7, 1, 0, False, True), (5, 3, 0, True, True), (3, 2, 1, False, False)])
def test_avgpool(k, s, p, ceil, cip):
    from bigdl_b200.nn.layers_core import SpatialAveragePooling
    torch.manual_seed(3)
    pool = SpatialAveragePooling(k, k, s, s, p, p, ceil_mode=ceil, count_include_pad=cip)
    x = torch.randn(2, 32, 14, 14, device=DEV).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = pool.forward(x)
    xr = x.float().cpu().requires_grad_(True)
    yr = pool._call(xr)
    assert y.shape == yr.shape
    _close(y.cpu(), yr, 1e-2, 1e-2)
    go = torch.randn_like(yr).to(torch.bfloat16)
    gx = pool.backward(x, go.to(DEV))
    yr.backward(go.float())
    _close(gx.cpu(), xr.grad, 2e-2, 2e-2)


def test_lrn():
    from bigdl_b200.nn.layers_core import SpatialCrossMapLRN
    torch.manual_seed(4)
    lrn = SpatialCrossMapLRN(5, 1e-4, 0.75, 1.0)
    x = (torch.randn(2, 64, 9, 9, device=DEV) * 30).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = lrn.forward(x)
    xr = x.float().cpu().requires_grad_(True)
    yr = lrn._call(xr)
    _close(y.cpu(), yr, 1e-2, 1e-2)
    yt = F.local_response_norm(x.float().cpu(), 5, 1e-4, 0.75, 1.0)
    _close(yr.detach(), yt, 1e-5, 1e-5)
    go = torch.randn_like(yr).to(torch.bfloat16)
    gx = lrn.backward(x, go.to(DEV))
    yr.backward(go.float())
    _close(gx.cpu(), xr.grad, 2e-2, 2e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fused_cross_entropy(dtype):
    from bigdl_b200.nn.criterion import CrossEntropyCriterion
    torch.manual_seed(5)
    B, C = 128, 1000
    logits = (torch.randn(B, C, device=DEV) * 3).to(dtype)
    target = torch.randint(1, C + 1, (B,), device=DEV).float()
    w = torch.rand(C) + 0.5
    for weights in (None, w):
        crit = CrossEntropyCriterion(weights)
        loss = crit.forward(logits, target)
        ref = F.cross_entropy(logits.float(), target.long() - 1, weight=None if weights is None else w.to(DEV))
        assert abs(loss - float(ref.detach())) < 2e-3 * max(1.0, abs(float(ref.detach())))
        g = crit.backward(logits, target)
        lr = logits.float().clone().requires_grad_(True)
        F.cross_entropy(lr, target.long() - 1, weight=None if weights is None else w.to(DEV)).backward()
        _close(g, lr.grad, 2e-2, 1e-5)


def test_topk_counters():
    from bigdl_b200.optim.validation import Top1Accuracy, Top5Accuracy
    torch.manual_seed(6)
    out = torch.randn(256, 1000, device=DEV)
    tgt = torch.randint(1, 1001, (256,), device=DEV).float()
    tgt[:40] = (out[:40].argmax(1) + 1).float()
    r1, r5 = Top1Accuracy()(out, tgt), Top5Accuracy()(out, tgt)
    c1 = int((out.argmax(1) + 1 == tgt.long()).sum())
    c5 = int((out.topk(5, 1).indices + 1 == tgt.long()[:, None]).any(1).sum())
    assert (r1.correct, r5.correct) == (c1, c5)


def test_relu_add_dropout():
    from bigdl_b200.nn.layers_core import Dropout, ReLU
    from bigdl_b200.ops import elementwise as ew
    torch.manual_seed(7)
    x = torch.randn(4, 32, 7, 7, device=DEV).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    r = ReLU()
    y = r.forward(x)
    assert torch.equal(y, torch.relu(x))
    g = r.backward(x, torch.ones_like(x))
    assert torch.equal(g, (x > 0).to(torch.bfloat16))
    a = torch.randn(1000, device=DEV).to(torch.bfloat16)
    b = torch.randn(1000, device=DEV).to(torch.bfloat16)
    _close(ew.eltwise(ew.ADD_RELU, a, b), torch.relu(a.float() + b.float()), 1e-2, 1e-2)
    d = Dropout(0.4)
    big = torch.ones(1 << 20, device=DEV, dtype=torch.bfloat16)
    yd = d.forward(big)
    keep = float((yd != 0).float().mean())
    assert abs(keep - 0.6) < 0.01
    assert abs(float(yd.float().mean()) - 1.0) < 0.02
    gd = d.backward(big, torch.ones_like(big))
    assert torch.equal(gd != 0, yd != 0)


def test_casts_and_truncation():
    from bigdl_b200.ops import elementwise as ew
    x = torch.tensor([1.111111, 4.111111, -3.3333, 1e-3] * 5, device=DEV)
    t = ew.cast_f32_to_bf16(x, truncate=True).float()
    # vectors from T/parameters/FP16ParameterSpec.scala:47-66
    assert abs(float(t[0]) - 1.109375) < 1e-7 and abs(float(t[1]) - 4.09375) < 1e-7
    r = ew.cast_f32_to_bf16(x, truncate=False)
    assert torch.equal(r, x.to(torch.bfloat16))
    back = ew.cast_bf16_to_f32(r)
    assert torch.equal(back, r.float())
    s = ew.bf16_trunc_add(t.to(torch.bfloat16), t.to(torch.bfloat16)).float()
    assert torch.equal(s, (t + t))


@pytest.mark.parametrize("name", ["sgd", "sgd_nesterov", "adam", "adagrad", "rmsprop", "adadelta", "adamax", "ftrl"])
def test_optim_kernels_match_cpu(name):
    from bigdl_b200.optim import optim_method as om
    mk = {
        "sgd": lambda: om.SGD(0.1, 0.01, 1e-4, 0.9),
        "sgd_nesterov": lambda: om.SGD(0.1, 0.0, 0.0, 0.9, 0.0, True),
        "adam": lambda: om.Adam(1e-2),
        "adagrad": lambda: om.Adagrad(1e-1, 0.01, 1e-3),
        "rmsprop": lambda: om.RMSprop(1e-2),
        "adadelta": lambda: om.Adadelta(),
        "adamax": lambda: om.Adamax(2e-3),
        "ftrl": lambda: om.Ftrl(0.1, -0.5, 0.1, 0.01, 0.01, 0.001),
    }[name]
    torch.manual_seed(8)
    n = 4099
    x0 = torch.randn(n)
    a, b = mk(), mk()
    xc, xg = x0.clone(), x0.clone().to(DEV)
    w16 = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    for step in range(4):
        g = torch.randn(n)
        a.update(xc, g.clone())
        b.update(xg, g.to(DEV), w16=w16)
    _close(xg.cpu(), xc, 1e-5, 1e-5)   # host libm vs device rsqrt differ in the last ulps
    assert torch.equal(w16, xg.to(torch.bfloat16))


@pytest.mark.parametrize("relu", [False, True])
@pytest.mark.parametrize("train", [True, False])
def test_batchnorm_fwd_bwd(relu, train):
    from bigdl_b200.nn.layers_core import SpatialBatchNormalization
    from bigdl_b200.utils.random_generator import RNG
    RNG.setSeed(11)
    torch.manual_seed(11)
    bn = SpatialBatchNormalization(32, 1e-3)
    bn._fuse_relu = relu
    ref = bn.cloneModule()
    ref._fuse_relu = relu
    if not train:
        bn.evaluate()
        ref.evaluate()
        with torch.no_grad():
            for m in (bn, ref):
                m._buffers["runningMean"].copy_(torch.linspace(-1, 1, 32))
                m._buffers["runningVar"].copy_(torch.linspace(0.5, 2, 32))
    bn.cuda()
    x = (torch.randn(8, 32, 14, 14) * 2 + 0.5)
    xb = x.cuda().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = bn.forward(xb)
    yr = ref.forward(xb.float().cpu())
    _close(y.cpu(), yr, 2e-2, 2e-2)
    if train:
        _close(bn.runningMean.cpu(), ref.runningMean, 1e-3, 1e-3)
        _close(bn.runningVar.cpu(), ref.runningVar, 1e-3, 1e-3)
    go = torch.randn_like(yr).to(torch.bfloat16)
    gx = bn.backward(xb, go.cuda())
    gr = ref.backward(xb.float().cpu(), go.float())
    _close(gx.cpu(), gr, 5e-2, 5e-2)
    _close(bn.gradWeight.cpu(), ref.gradWeight, 3e-2, 0.5)
    _close(bn.gradBias.cpu(), ref.gradBias, 3e-2, 0.5)


def test_lstm_gates_kernel():
    from bigdl_b200.ops.recurrent import lstm_gates
    torch.manual_seed(12)
    B, H = 64, 96
    g = torch.randn(B, 4 * H, device=DEV).to(torch.bfloat16).requires_grad_(True)
    c0 = torch.randn(B, H, device=DEV).requires_grad_(True)
    h, c = lstm_gates(g, c0)
    gf = g.detach().float().requires_grad_(True)
    cf = c0.detach().clone().requires_grad_(True)
    i, z, f, o = torch.sigmoid(gf[:, :H]), torch.tanh(gf[:, H:2 * H]), torch.sigmoid(gf[:, 2 * H:3 * H]), torch.sigmoid(gf[:, 3 * H:])
    cr = i * z + f * cf
    hr = torch.tanh(cr) * o
    _close(h, hr, 2e-2, 2e-2)
    _close(c, cr, 2e-2, 2e-2)
    dh, dc = torch.randn_like(hr), torch.randn_like(cr)
    torch.autograd.backward([h, c], [dh.to(torch.bfloat16), dc])
    torch.autograd.backward([hr, cr], [dh.to(torch.bfloat16).float(), dc])
    _close(g.grad, gf.grad, 4e-2, 4e-2)
    _close(c0.grad, cf.grad, 4e-2, 4e-2)


def test_recurrent_lstm_on_gpu_matches_cpu():
    from bigdl_b200.nn.layer import LSTM, Recurrent
    from bigdl_b200.utils.random_generator import RNG
    RNG.setSeed(13)
    r = Recurrent().add(LSTM(64, 128))
    ref = r.cloneModule()
    r.cuda()
    x = torch.randn(4, 6, 64)
    y = r.forward(x.cuda().to(torch.bfloat16))
    yr = ref.forward(x)
    _close(y.cpu(), yr, 5e-2, 5e-2)
    go = torch.randn_like(yr)
    gx = r.backward(x.cuda().to(torch.bfloat16), go.cuda().to(torch.bfloat16))
    gr = ref.backward(x, go)
    _close(gx.cpu(), gr, 8e-2, 8e-2)


@pytest.mark.parametrize("layout", ["nhwc", "nchw"])
def test_gpu_image_batch_matches_torch(layout):
    """uint8 frames -> crop + flip + normalise on the device vs the torch restatement of the same pipeline."""
    from bigdl_b200.ops.data import _image_batch_torch, image_batch
    torch.manual_seed(30)
    frames = torch.randint(0, 256, (5, 40, 48, 3), dtype=torch.uint8)
    crop = torch.tensor([[0, 0], [16, 8], [3, 5], [16, 0], [7, 8]], dtype=torch.int32)
    flip = torch.tensor([0, 1, 1, 0, 1], dtype=torch.uint8)
    mean, std = (104.0, 117.0, 123.0), (58.0, 57.0, 57.5)
    cl = layout == "nhwc"
    for dt in (torch.float32, torch.bfloat16):
        got = image_batch(frames.to(DEV), 32, 32, mean, std, crop, flip, dt, cl)
        ref = _image_batch_torch(frames, 32, 32, mean, std, crop, flip, torch.float32, cl)
        assert got.shape == ref.shape == (5, 3, 32, 32)
        _close(got.float().cpu(), ref, 1e-2 if dt == torch.bfloat16 else 1e-6, 1e-5)
    centre = image_batch(frames.to(DEV), 32, 32, mean, std, None, None, torch.float32, cl)
    _close(centre.cpu(), _image_batch_torch(frames, 32, 32, mean, std, None, None, torch.float32, cl), 1e-6, 1e-5)


def test_gru_fused_step_matches_reference():
    from bigdl_b200.ops.recurrent import gru_out, gru_rz
    torch.manual_seed(13)
    B, H = 48, 64
    xt = torch.randn(B, 3 * H, device=DEV).bfloat16().requires_grad_()
    hg = torch.randn(B, 2 * H, device=DEV).bfloat16().requires_grad_()
    hc = torch.randn(B, H, device=DEV).bfloat16().requires_grad_()
    h = torch.randn(B, H, device=DEV).bfloat16().requires_grad_()
    z, hr = gru_rz(xt[:, :2 * H], hg, h)
    out = gru_out(xt[:, 2 * H:], hc, z, h)
    f = [t.detach().float().requires_grad_() for t in (xt, hg, hc, h)]
    rz = torch.sigmoid(f[0][:, :2 * H] + f[1])
    r_ref, z_ref = rz[:, :H], rz[:, H:]
    hr_ref = f[3] * r_ref
    hh = torch.tanh(f[0][:, 2 * H:] + f[2])
    out_ref = (1 - z_ref) * hh + z_ref * f[3]
    _close(hr, hr_ref, 2e-2, 2e-2)
    _close(out, out_ref, 2e-2, 2e-2)
    g1, g2 = torch.randn_like(out_ref), torch.randn_like(hr_ref)
    torch.autograd.backward([out, hr], [g1.bfloat16(), g2.bfloat16()])
    torch.autograd.backward([out_ref, hr_ref], [g1.bfloat16().float(), g2.bfloat16().float()])
    for a, b in zip((xt, hg, hc, h), f):
        _close(a.grad, b.grad, 4e-2, 4e-2)


@pytest.mark.parametrize("log_form", [False, True])
@pytest.mark.parametrize("shape,dtype", [((128, 1000), torch.bfloat16), ((7, 33), torch.float32), ((4, 12, 35, 35), torch.bfloat16)])
def test_softmax_rows_native(shape, dtype, log_form):
    """SoftMax / LogSoftMax row kernel (fwd + bwd from the output) vs fp32 torch."""
    from bigdl_b200.ops.functions import SoftmaxFn
    torch.manual_seed(3)
    x = (torch.randn(*shape, device="cuda") * 3).to(dtype).requires_grad_(True)
    out_dtype = torch.float32 if log_form else dtype
    y = SoftmaxFn.apply(x, log_form, out_dtype)
    dy = torch.randn(*shape, device="cuda").to(out_dtype)
    y.backward(dy)
    xr = x.detach().float().requires_grad_(True)
    yr = torch.log_softmax(xr, -1) if log_form else torch.softmax(xr, -1)
    yr.backward(dy.float())
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
    assert torch.allclose(y.float(), yr, rtol=tol, atol=tol)
    assert torch.allclose(x.grad.float(), xr.grad, rtol=5 * tol, atol=5 * tol)


def test_scatter_acc_clear():
    from bigdl_b200.ops.functions import _scatter_bias_grads
    src = torch.arange(40, device="cuda", dtype=torch.float32)
    a, b, c = torch.ones(8, device="cuda"), torch.ones(24, device="cuda"), torch.ones(8, device="cuda")
    ref = src.clone()
    _scatter_bias_grads(src, [a, b, c], [8, 24, 8])
    torch.cuda.synchronize()
    assert torch.equal(a, 1 + ref[:8]) and torch.equal(b, 1 + ref[8:32]) and torch.equal(c, 1 + ref[32:])
    assert float(src.abs().max()) == 0.0
