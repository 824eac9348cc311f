"""One attention configuration, forward (and optionally backward), for ncu: python attention_one.py B L heads D causal bwd"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bigdl_b200.ops import attention as A

B, L, heads, D, causal, bwd = (int(x) for x in (sys.argv[1:7] + ["16", "1024", "16", "64", "0", "0"][len(sys.argv) - 1:]))
W = heads * D
q, k, v = (torch.randn(B, L, W, device="cuda").bfloat16().requires_grad_(bool(bwd)) for _ in range(3))
for _ in range(3):
    out = A.attention(q, k, v, None, heads, causal=bool(causal))
    if bwd:
        out.backward(torch.ones_like(out))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    out = A.attention(q, k, v, None, heads, causal=bool(causal))
    if bwd:
        out.backward(torch.ones_like(out))
e1.record()
torch.cuda.synchronize()
print("ms per iteration", e0.elapsed_time(e1) / 10)
