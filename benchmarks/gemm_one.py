"""One GEMM shape, a few launches of ours and of cuBLAS — the target of ``ncu`` captures (benchmarks/gemm_bench.py is the
timed sweep). Usage: python benchmarks/gemm_one.py M N K [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bigdl_b200.ops.gemm import gemm_bf16

M, N, K = (int(v) for v in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 3
a = torch.randn(M, K, device="cuda").bfloat16()
b = torch.randn(N, K, device="cuda").bfloat16()
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(iters):
    gemm_bf16(a, b, out=out)
    torch.matmul(a, b.t())
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
