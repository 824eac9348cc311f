"""tcgen05 GEMM vs a plain fp32 PyTorch reference (all four operand-major combinations, ragged shapes, split-K)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn):
    A = a.float().t() if a_mn else a.float()
    B = b.float() if b_mn else b.float().t()
    return A @ B


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 512, 256), (384, 200, 136), (1000, 96, 72), (130, 1000, 1024),
                                   (4096, 2048, 512)])
def test_gemm_majors(a_mn, b_mn, M, N, K):
    from bigdl_b200.ops.gemm import gemm_bf16
    torch.manual_seed(0)
    dev = "cuda"
    # leading dims must be multiples of 8 elements: pad the storage, slice the view
    def mk(rows, cols):
        ld = (cols + 7) // 8 * 8
        return torch.randn(rows, ld, device=dev).to(torch.bfloat16)[:, :cols]
    a = mk(K, M) if a_mn else mk(M, K)
    b = mk(K, N) if b_mn else mk(N, K)
    out = gemm_bf16(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32)
    torch.cuda.synchronize()
    ref = _ref(a, b, a_mn, b_mn)
    err = (out - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err <= 2e-3 * scale + 1e-3, (err, scale)


def test_gemm_bias_relu_bf16_out():
    from bigdl_b200.ops.gemm import gemm_bf16
    torch.manual_seed(1)
    a = torch.randn(512, 320, device="cuda").bfloat16()
    b = torch.randn(192, 320, device="cuda").bfloat16()
    bias = torch.randn(192, device="cuda")
    wide = torch.zeros(512, 256, device="cuda", dtype=torch.bfloat16)
    out = gemm_bf16(a, b, bias=bias, relu=True, out=wide[:, 32:224])
    torch.cuda.synchronize()
    ref = torch.relu(a.float() @ b.float().t() + bias)
    assert torch.allclose(out.float(), ref, rtol=2e-2, atol=2e-1)
    assert wide[:, :32].abs().max().item() == 0 and wide[:, 224:].abs().max().item() == 0


def test_gemm_splitk_accumulate():
    from bigdl_b200.ops.gemm import gemm_bf16
    torch.manual_seed(2)
    # wgrad shape: dW[Cout, K] = dY[M, Cout]^T X[M, K]; tiny output, huge contraction -> split-K with red.add
    M, Cout, Kd = 128 * 196, 64, 576
    dy = torch.randn(M, Cout, device="cuda").bfloat16()
    x = torch.randn(M, Kd, device="cuda").bfloat16()
    acc = torch.ones(Cout, Kd, device="cuda")
    gemm_bf16(dy, x, a_mn=True, b_mn=True, out=acc, accumulate=True)
    torch.cuda.synchronize()
    ref = 1.0 + dy.float().t() @ x.float()
    err = (acc - ref).abs().max().item()
    assert err <= 2e-3 * ref.abs().max().item() + 1e-2, err


@pytest.mark.parametrize("Cout,Kd", [(192, 576), (320, 264), (256, 1152), (136, 96)])
def test_gemm_splitk_two_stacked_m_tiles(Cout, Kd):
    """Weight-gradient shapes with more than 128 output rows: one CTA owns two stacked 128-row accumulators (dual_m),
    including a last pair whose second half is partly or wholly out of range, and wide n tiles."""
    from bigdl_b200.ops.gemm import gemm_bf16
    torch.manual_seed(4)
    M = 128 * 160
    dy = torch.randn(M, Cout, device="cuda").bfloat16()
    x = torch.randn(M, Kd, device="cuda").bfloat16()
    acc = torch.ones(Cout, Kd, device="cuda")
    gemm_bf16(dy, x, a_mn=True, b_mn=True, out=acc, accumulate=True)
    torch.cuda.synchronize()
    ref = 1.0 + dy.float().t() @ x.float()
    err = (acc - ref).abs().max().item()
    assert err <= 2e-3 * ref.abs().max().item() + 1e-2, err


def test_linear_fn_output_size_not_multiple_of_8():
    """10-way classifier head: backward pads dY for the MN-major GEMM operands."""
    import torch
    from bigdl_b200.nn.layer import Linear

    torch.manual_seed(0)
    lin = Linear(64, 10)
    ref = lin.cloneModule()
    lin.cuda()
    x = torch.randn(32, 64)
    go = torch.randn(32, 10)
    y = lin.forward(x.cuda().to(torch.bfloat16))
    gi = lin.backward(x.cuda().to(torch.bfloat16), go.cuda().to(torch.bfloat16))
    yr = ref.forward(x)
    gir = ref.backward(x, go)
    assert torch.allclose(y.float().cpu(), yr, atol=5e-2, rtol=5e-2)
    assert torch.allclose(gi.float().cpu(), gir, atol=5e-2, rtol=5e-2)
    gw = [p.grad for p in lin.torch_parameters()]
    gwr = [p.grad for p in ref.torch_parameters()]
    for a, b in zip(gw, gwr):
        assert torch.allclose(a.float().cpu(), b, atol=0.15, rtol=5e-2)


@pytest.fixture
def pair_mode():
    """Force (1) or forbid (0) the CTA-pair kernels for one test, restoring the heuristic afterwards."""
    from bigdl_b200.ops import _lib
    lib = _lib.cuda_lib()
    old = []

    def set_mode(m):
        old.append(lib.bdl_gemm_set_pair_mode(m))

    yield set_mode
    if old:
        lib.bdl_gemm_set_pair_mode(old[0])


@pytest.mark.parametrize("b_mn", [False, True])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (300, 128, 200), (1024, 512, 1024), (2048, 96, 320), (8192, 384, 512),
                                   (4096, 1000, 256),
                                   (16384, 768, 1024), (19000, 520, 1032)])   # 512-row pair tiles + banded tile order
def test_gemm_cta_pair_matches_single_cta_and_fp32(pair_mode, b_mn, M, N, K):
    """cta_group::2 kernels (forced) vs the fp32 reference and vs the 1-CTA kernel: bf16 and fp32 outputs, bias + ReLU,
    ragged M / N / K, both B majors."""
    from bigdl_b200.ops.gemm import gemm_bf16
    torch.manual_seed(3)
    ldn = (N + 7) // 8 * 8
    a = torch.randn(M, (K + 7) // 8 * 8, device="cuda").bfloat16()[:, :K]
    b = torch.randn(K, ldn, device="cuda").bfloat16()[:, :N] if b_mn else torch.randn(N, (K + 7) // 8 * 8, device="cuda").bfloat16()[:, :K]
    bias = torch.randn(N, device="cuda")
    ref = a.float() @ (b.float() if b_mn else b.float().t())
    outs = {}
    for mode in (1, 0):
        pair_mode(mode)
        o32 = gemm_bf16(a, b, b_mn=b_mn, out_dtype=torch.float32)
        o16 = torch.zeros(M, ldn, device="cuda", dtype=torch.bfloat16)
        gemm_bf16(a, b, b_mn=b_mn, bias=bias, relu=True, out=o16[:, :N])
        torch.cuda.synchronize()
        outs[mode] = (o32, o16)
        err = (o32 - ref).abs().max().item()
        assert err <= 2e-3 * ref.abs().max().item() + 1e-3, (mode, err)
        assert torch.allclose(o16[:, :N].float(), torch.relu(ref + bias), rtol=2e-2, atol=2e-1), mode
        assert o16[:, N:].abs().max().item() == 0 if ldn > N else True
    # same accumulation order per output element in both kernels -> identical results
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("M,N,K,ranks", [(128, 1024, 4096, 4), (96, 1000, 2048, 3), (384, 512, 512, 2), (128, 4096, 1024, 8)])
def test_linear_forward_with_fused_allgather(M, N, K, ranks):
    """The all-gather-fused GEMM on ONE GPU with N virtual owners (separate buffers stand in for the peers' memory —
    the kernel only ever issues plain TMA loads through them): tiles read from the owner's buffer, written through to
    the local copy, split-K + bias/ReLU finish and the single-pass form, vs the fp32 reference."""
    from bigdl_b200.ops.gemm import linear_forward
    torch.manual_seed(11)
    dev = "cuda"
    bn = 128
    tiles = (N + bn - 1) // bn
    x = torch.randn(M, K, device=dev).bfloat16()
    w_true = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
    bias = torch.randn(N, device=dev)
    # owner r holds tiles [t0[r], t1[r]); a tile in between two owners is "boundary" (already in the local copy)
    cuts = [round(i * tiles / ranks) for i in range(ranks + 1)]
    t0, t1 = [cuts[r] for r in range(ranks)], [cuts[r + 1] for r in range(ranks)]
    if tiles >= 2 * ranks:
        t1[0] -= 1                                   # leave one boundary tile that nobody serves remotely
    me = 1 if ranks > 1 else 0
    bufs = []
    for r in range(ranks):
        b = torch.full((N, K), float("nan"), device=dev, dtype=torch.bfloat16)
        b[t0[r] * bn:t1[r] * bn] = w_true[t0[r] * bn:t1[r] * bn]
        bufs.append(b)
    local = bufs[me]
    served = torch.zeros(N, dtype=torch.bool, device=dev)
    for r in range(ranks):
        served[t0[r] * bn:t1[r] * bn] = True
    local[~served] = w_true[~served]                 # boundary rows were pushed by their owners
    plan = {"nranks": ranks, "rank": me, "block_n": bn, "t0": t0, "t1": t1, "ptrs": [b.data_ptr() for b in bufs]}
    ref = torch.relu(x.float() @ w_true.float().t() + bias)
    y = linear_forward(x, local, bias, True, gather=plan, key=("t", M, N, K))
    torch.cuda.synchronize()
    assert torch.allclose(y.float(), ref, rtol=2e-2, atol=2e-1)
    assert torch.equal(local, w_true)                # every remote tile was written through to the local copy
    y2 = linear_forward(x, local, bias, True, gather=plan, key=("t", M, N, K))      # accumulator was re-zeroed
    torch.cuda.synchronize()
    assert torch.allclose(y2.float(), ref, rtol=2e-2, atol=2e-1)      # (split-K adds arrive in any order: not bit-equal)
    y3 = linear_forward(x, w_true, bias, True, gather=None, key=("u", M, N, K))     # same shapes without the gather
    torch.cuda.synchronize()
    assert torch.allclose(y3.float(), ref, rtol=2e-2, atol=2e-1)
