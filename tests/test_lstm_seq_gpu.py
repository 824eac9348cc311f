"""Persistent whole-sequence LSTM kernel (csrc/lstm_seq.cu) vs a plain fp32 PyTorch LSTM recurrence: outputs, final
state and every gradient (input projection, recurrent weights, initial state), PTB-like and ragged shapes, both time
directions; and the Recurrent layer taking the fused path end to end."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref_lstm(xw, whh, h0, c0, reverse):
    B, T, H4 = xw.shape
    H = H4 // 4
    h, c, outs = h0, c0, [None] * T
    order = range(T - 1, -1, -1) if reverse else range(T)
    for t in order:
        g = xw[:, t] + h @ whh.t()
        i, z, f, o = torch.sigmoid(g[:, :H]), torch.tanh(g[:, H:2 * H]), torch.sigmoid(g[:, 2 * H:3 * H]), torch.sigmoid(g[:, 3 * H:])
        c = i * z + f * c
        h = o * torch.tanh(c)
        outs[t] = h
    return torch.stack(outs, 1), h, c


@pytest.mark.parametrize("B,T,H,reverse,with_state", [(20, 35, 200, False, False), (20, 35, 650, False, True),
                                                      (7, 5, 72, True, True), (40, 6, 96, False, False),
                                                      (20, 8, 1500, False, True)])
def test_lstm_sequence_kernel_matches_fp32_recurrence(B, T, H, reverse, with_state):
    from bigdl_b200.ops.recurrent import LstmSeqFn, lstm_seq_supported
    assert lstm_seq_supported(H, torch.device("cuda"))
    torch.manual_seed(0)
    dev = "cuda"
    xw = (torch.randn(B, T, 4 * H, device=dev) * 0.5).bfloat16().requires_grad_(True)
    whh = (torch.randn(4 * H, H, device=dev) / H ** 0.5).requires_grad_(True)
    whh.grad = torch.zeros_like(whh)
    h0 = (torch.randn(B, H, device=dev) * 0.3).bfloat16().requires_grad_(True) if with_state else None
    c0 = (torch.randn(B, H, device=dev) * 0.3).requires_grad_(True) if with_state else None
    hs, h_last, c_last = LstmSeqFn.apply(xw, whh, whh.detach().bfloat16(), h0, c0, reverse)
    gout = (torch.randn(B, T, H, device=dev) * 0.1).bfloat16()
    glast = (torch.randn(B, H, device=dev) * 0.1).bfloat16()
    torch.autograd.backward([hs, h_last], [gout, glast])
    torch.cuda.synchronize()

    xr = xw.detach().float().requires_grad_(True)
    wr = whh.detach().bfloat16().float().requires_grad_(True)
    h0r = h0.detach().float().requires_grad_(True) if with_state else torch.zeros(B, H, device=dev)
    c0r = c0.detach().float().requires_grad_(True) if with_state else torch.zeros(B, H, device=dev)
    rs, rh, rc = _ref_lstm(xr, wr, h0r, c0r, reverse)
    torch.autograd.backward([rs, rh], [gout.float(), glast.float()])

    def rel(a, b):
        return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-9))

    assert rel(hs, rs) < 2e-2 and rel(h_last, rh) < 2e-2 and rel(c_last, rc) < 2e-2
    assert rel(xw.grad, xr.grad) < 4e-2, rel(xw.grad, xr.grad)
    assert rel(whh.grad, wr.grad) < 4e-2, rel(whh.grad, wr.grad)
    if with_state:
        assert rel(h0.grad, h0r.grad) < 4e-2 and rel(c0.grad, c0r.grad) < 4e-2


def test_recurrent_layer_takes_the_persistent_path_and_matches_the_step_loop():
    from bigdl_b200.nn.layer import LSTM, Recurrent
    from bigdl_b200.ops import _lib
    from bigdl_b200.utils import config
    from bigdl_b200.utils.random_generator import RNG

    RNG.setSeed(4)
    m = Recurrent().add(LSTM(64, 128)).cuda()
    m.training()
    x = torch.randn(6, 9, 64, device="cuda").bfloat16()
    res = {}
    for fused in (True, False):
        config.set("bigdl.rnn.persistent", fused)
        try:
            m.zeroGradParameters()
            xin = x.clone().requires_grad_(True)
            n0 = _lib.LAUNCHES
            with torch.enable_grad():
                y = m._call(xin)
            torch.autograd.backward(y, torch.ones_like(y) * 0.1)
            torch.cuda.synchronize()
            res[fused] = (y.detach().float(), xin.grad.float(), torch.cat([p.grad.flatten().clone() for p in m.torch_parameters()]),
                          _lib.LAUNCHES - n0)
        finally:
            config.unset("bigdl.rnn.persistent")

    def rel(a, b):
        return float((a - b).norm() / (b.norm() + 1e-9))

    assert rel(res[True][0], res[False][0]) < 2e-2 and rel(res[True][1], res[False][1]) < 4e-2
    assert rel(res[True][2], res[False][2]) < 4e-2
    assert res[True][3] < res[False][3] / 2, (res[True][3], res[False][3])       # far fewer launches than T x (GEMM + gates)
