"""Max-pooling and LRN forward/backward at the Inception-v1 shapes: device time (CUDA events, L2 flushed between
iterations) and achieved bytes/s against the minimum traffic of each kernel.
usage: python benchmarks/pool_lrn_bench.py [--batch 128] [--iters 10]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bigdl_b200.ops import functions as fn  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--iters", type=int, default=10)
a = ap.parse_args()
N = a.batch
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(f):
    ts = []
    for it in range(a.iters + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        f()
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


# (C, H, k, s, pad, ceil) — the stride-2 reductions and the stride-1 pools inside the inception modules
POOLS = [(64, 112, 3, 2, 0, True), (192, 56, 3, 2, 0, True), (192, 28, 3, 1, 1, False), (256, 28, 3, 1, 1, False),
         (480, 28, 3, 2, 0, True), (480, 14, 3, 1, 1, False), (512, 14, 3, 1, 1, False), (528, 14, 3, 1, 1, False),
         (832, 14, 3, 2, 0, True), (832, 7, 3, 1, 1, False)]
tot_f = tot_b = 0.0
for C, H, k, s, p, ceil in POOLS:
    Ho = (-(-(H + 2 * p - k) // s) if ceil else (H + 2 * p - k) // s) + 1
    x = torch.randn(N, H, H, C, device="cuda").bfloat16().permute(0, 3, 1, 2).requires_grad_()
    y = fn.MaxPoolFn.apply(x, k, k, s, s, p, p, Ho, Ho)
    gy = torch.randn(N, Ho, Ho, C, device="cuda").bfloat16().permute(0, 3, 1, 2)
    tf = timed(lambda: fn.MaxPoolFn.apply(x, k, k, s, s, p, p, Ho, Ho))
    tb = timed(lambda: torch.autograd.grad(y, x, gy, retain_graph=True))
    bin_, bout = N * H * H * C * 2, N * Ho * Ho * C * 2
    tot_f += tf
    tot_b += tb
    print(json.dumps({"op": "maxpool", "C": C, "H": H, "stride": s, "fwd_us": round(tf, 1), "bwd_us": round(tb, 1),
                      "fwd_GBps": round((bin_ + bout * 1.5) / tf / 1e3), "bwd_GBps": round((bin_ + bout * 1.5) / tb / 1e3)}))
print(json.dumps({"op": "maxpool_total", "fwd_us": round(tot_f, 1), "bwd_us": round(tot_b, 1)}))
for C, H in [(64, 56), (192, 56)]:
    x = torch.randn(N, H, H, C, device="cuda").bfloat16().permute(0, 3, 1, 2).requires_grad_()
    y = fn.LRNFn.apply(x, 5, 1e-4, 0.75, 1.0)
    gy = torch.randn(N, H, H, C, device="cuda").bfloat16().permute(0, 3, 1, 2)
    tf = timed(lambda: fn.LRNFn.apply(x, 5, 1e-4, 0.75, 1.0))
    tb = timed(lambda: torch.autograd.grad(y, x, gy, retain_graph=True))
    b = N * H * H * C * 2
    print(json.dumps({"op": "lrn", "C": C, "H": H, "fwd_us": round(tf, 1), "bwd_us": round(tb, 1),
                      "fwd_GBps": round(2 * b / tf / 1e3), "bwd_GBps": round(3 * b / tb / 1e3)}))
