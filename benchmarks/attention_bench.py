"""Fused attention forward (csrc/attention.cu) vs the unfused composition and vs torch SDPA (flash / cuDNN), CUDA events."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from bigdl_b200.ops import attention as A


def timeit(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


for (B, L, heads, D, causal) in [(32, 256, 8, 64, False), (32, 256, 8, 64, True), (16, 1024, 16, 64, False),
                                 (16, 1024, 16, 64, True), (8, 2048, 8, 128, False), (4, 4096, 16, 64, True)]:
    W = heads * D
    q, k, v = (torch.randn(B, L, W, device="cuda").bfloat16() for _ in range(3))
    qh, kh, vh = (t.reshape(B, L, heads, D).transpose(1, 2) for t in (q, k, v))
    bias = torch.full((1, 1, L, L), -1e9, device="cuda").triu(1) if causal else None

    def unfused():
        s = torch.matmul(qh * D ** -0.5, kh.transpose(-1, -2))
        if bias is not None:
            s = s + bias.to(s.dtype)
        w = F.softmax(s.float(), dim=-1).to(q.dtype)
        return torch.matmul(w, vh).transpose(1, 2).reshape(B, L, W)

    ours = timeit(lambda: A.attention(q, k, v, None, heads, causal=causal))
    ours_bias = timeit(lambda: A.attention(q, k, v, bias, heads)) if causal else None
    sdpa = timeit(lambda: F.scaled_dot_product_attention(qh, kh, vh, is_causal=causal))
    unf = timeit(unfused)
    flops = 4.0 * B * heads * L * L * D * (0.5 if causal else 1.0)
    print(json.dumps({"B": B, "L": L, "heads": heads, "D": D, "causal": causal, "ours_ms": round(ours, 4),
                      "ours_with_bias_tensor_ms": None if ours_bias is None else round(ours_bias, 4),
                      "torch_sdpa_ms": round(sdpa, 4), "unfused_ms": round(unf, 4),
                      "ours_tflops": round(flops / ours / 1e9, 1), "vs_sdpa": round(sdpa / ours, 3),
                      "vs_unfused": round(unf / ours, 2)}), flush=True)
