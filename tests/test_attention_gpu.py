"""Fused attention kernel (csrc/attention.cu) against the fp32 composition of S/nn/Attention.scala:95-114."""
import pytest
import torch

from bigdl_b200.ops import attention as A

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-9))


def _qkv(B, Lq, Lk, heads, D, seed=0):
    g = torch.Generator(device=DEV).manual_seed(seed)
    mk = lambda L: (torch.randn(B, L, heads * D, device=DEV, generator=g)).bfloat16()
    return mk(Lq), mk(Lk), mk(Lk)


@pytest.mark.parametrize("B,Lq,Lk,heads,D", [(2, 128, 128, 4, 64), (3, 77, 200, 2, 64), (1, 384, 384, 8, 64),
                                             (2, 128, 256, 2, 128), (2, 50, 50, 3, 128), (1, 1, 300, 4, 64)])
def test_forward_matches_fp32(B, Lq, Lk, heads, D):
    q, k, v = _qkv(B, Lq, Lk, heads, D)
    out = A.attention(q, k, v, None, heads)
    ref = A.attention_reference(q, k, v, None, heads)
    assert out.shape == ref.shape
    assert _rel(out, ref) < 1.5e-2


@pytest.mark.parametrize("B,Lq,Lk,heads,D,causal", [(2, 256, 256, 2, 64, False), (1, 300, 300, 3, 64, True),
                                                    (2, 130, 400, 2, 64, False), (1, 512, 512, 2, 128, True),
                                                    (1, 257, 140, 2, 128, False)])
def test_two_query_tiles_per_cta(B, Lq, Lk, heads, D, causal):
    """The 256-query CTA (two softmax groups ping-ponging on one tensor core) against fp32 and against the 1-tile kernel."""
    q, k, v = _qkv(B, Lq, Lk, heads, D, seed=9)
    old = A.set_two_tile_mode(1)
    try:
        out2 = A.attention(q, k, v, None, heads, causal=causal)
        A.set_two_tile_mode(0)
        out1 = A.attention(q, k, v, None, heads, causal=causal)
    finally:
        A.set_two_tile_mode(old)
    ref = A.attention_reference(q, k, v, None, heads, causal=causal)
    assert _rel(out2, ref) < 1.5e-2 and _rel(out1, ref) < 1.5e-2
    assert _rel(out2, out1) < 2e-3


@pytest.mark.parametrize("kind", ["padding", "full", "causal_bias", "causal_flag"])
def test_bias_and_masks(kind):
    B, L, heads, D = 2, 200, 4, 64
    q, k, v = _qkv(B, L, L, heads, D, seed=1)
    bias, causal = None, False
    if kind == "padding":
        bias = torch.zeros(B, 1, 1, L, device=DEV)
        bias[0, ..., 150:] = -1e9
        bias[1, ..., 10:40] = -1e9
    elif kind == "full":
        bias = torch.randn(B, heads, L, L, device=DEV)
    elif kind == "causal_bias":
        bias = torch.full((1, 1, L, L), -1e9, device=DEV).triu(1)
    else:
        causal = True
    out = A.attention(q, k, v, bias, heads, causal=causal)
    ref = A.attention_reference(q, k, v, bias, heads, causal=causal)
    assert _rel(out, ref) < 1.5e-2
    if kind == "causal_flag":
        ref2 = A.attention_reference(q, k, v, torch.full((1, 1, L, L), -1e9, device=DEV).triu(1), heads)
        assert _rel(out, ref2) < 1.5e-2


def test_strided_inputs_and_backward():
    B, L, heads, D = 2, 130, 2, 64
    W = heads * D
    g = torch.Generator(device=DEV).manual_seed(3)
    qkv = torch.randn(B, L, 3 * W, device=DEV, generator=g).bfloat16().requires_grad_(True)   # one fused projection
    q, k, v = qkv[..., :W], qkv[..., W:2 * W], qkv[..., 2 * W:]
    bias = torch.full((1, 1, L, L), -1e9, device=DEV).triu(1)
    out = A.attention(q, k, v, bias, heads)
    go = torch.randn_like(out)
    out.backward(go)
    got = qkv.grad.clone()
    x = qkv.detach().float().requires_grad_(True)
    ref = A.attention_reference(x[..., :W], x[..., W:2 * W], x[..., 2 * W:], bias, heads)
    ref.backward(go.float())
    assert _rel(out, ref) < 1.5e-2
    assert _rel(got, x.grad) < 3e-2


@pytest.mark.parametrize("B,Lq,Lk,heads,D,kind", [(2, 128, 128, 2, 64, "none"), (2, 200, 200, 3, 64, "causal_flag"),
                                                  (1, 77, 300, 2, 64, "padding"), (2, 256, 130, 2, 128, "full"),
                                                  (1, 300, 300, 2, 128, "causal_bias"), (2, 40, 40, 1, 64, "none")])
def test_native_backward_matches_fp32(B, Lq, Lk, heads, D, kind):
    q, k, v = _qkv(B, Lq, Lk, heads, D, seed=5)
    bias, causal = None, False
    if kind == "padding":
        bias = torch.zeros(B, 1, 1, Lk, device=DEV)
        bias[0, ..., Lk - 50:] = -1e9
    elif kind == "full":
        bias = torch.randn(B, heads, Lq, Lk, device=DEV)
    elif kind == "causal_bias":
        bias = torch.full((1, 1, Lq, Lk), -1e9, device=DEV).triu(1)
    elif kind == "causal_flag":
        causal = True
    qs = [t.clone().requires_grad_(True) for t in (q, k, v)]
    out = A.attention(*qs, bias, heads, causal=causal)
    go = torch.randn_like(out)
    out.backward(go)
    fs = [t.detach().float().requires_grad_(True) for t in (q, k, v)]
    ref = A.attention_reference(*fs, bias, heads, causal=causal)
    ref.backward(go.float())
    for name, got, want in zip("qkv", qs, fs):
        assert _rel(got.grad, want.grad) < 3e-2, name


def test_attention_layer_uses_the_kernel():
    from bigdl_b200.nn.layers_attention import Attention
    from bigdl_b200.ops import _lib
    from bigdl_b200.utils.table import Table
    torch.manual_seed(0)
    m = Attention(256, 4, 0.0).cuda().evaluate()
    x = torch.randn(2, 96, 256, device=DEV).bfloat16()
    bias = torch.full((1, 1, 96, 96), -1e9, device=DEV).triu(1)
    before = _lib.LAUNCHES
    y = m.forward(Table(x, x, bias))
    assert _lib.LAUNCHES - before >= 5          # 4 projections + the attention kernel
    m._cache = {}                               # the cached (incremental decoding) path is the unfused composition
    y2 = m.forward(Table(x, x, bias))
    m._cache = None
    assert _rel(y, y2) < 2e-2
