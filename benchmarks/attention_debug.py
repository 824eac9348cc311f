import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    import torch
    from bigdl_b200.ops import attention as A
    B, L, heads, D, bias_kind, strided = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5], int(sys.argv[6])
    W = heads * D
    if strided:
        qkv = torch.randn(B, L, 3 * W, device="cuda").bfloat16().requires_grad_(True)
        q, k, v = qkv[..., :W], qkv[..., W:2 * W], qkv[..., 2 * W:]
    else:
        q, k, v = (torch.randn(B, L, W, device="cuda").bfloat16().requires_grad_(True) for _ in range(3))
    bias = None
    if bias_kind == "causal":
        bias = torch.full((1, 1, L, L), -1e9, device="cuda").triu(1)
    elif bias_kind == "pad":
        bias = torch.zeros(B, 1, 1, L, device="cuda")
    out = A.attention(q, k, v, bias, heads)
    torch.cuda.synchronize()
    print("fwd ok", flush=True)
    out.backward(torch.ones_like(out))
    torch.cuda.synchronize()
    print("bwd ok", flush=True)
    sys.exit(0)
for cfg in [(2, 128, 2, 64, "none", 0), (2, 130, 2, 64, "none", 0), (2, 128, 2, 64, "causal", 0), (2, 128, 2, 64, "none", 1),
            (2, 130, 2, 64, "causal", 1), (1, 300, 2, 128, "pad", 0), (2, 40, 1, 64, "none", 0)]:
    try:
        r = subprocess.run([sys.executable, __file__] + [str(c) for c in cfg], capture_output=True, text=True, timeout=60)
        print(cfg, "rc", r.returncode, r.stdout.strip().replace("\n", " | "), r.stderr.strip()[-300:], flush=True)
    except subprocess.TimeoutExpired as e:
        print(cfg, "TIMEOUT", (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else e.stdout, flush=True)
