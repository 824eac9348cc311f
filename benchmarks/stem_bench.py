"""Stem convolution (7x7/2, 3→64) forward+backward: space-to-depth implicit GEMM vs explicit im2col.
usage: python benchmarks/stem_bench.py [--batch 128]"""
import argparse
import json

import torch
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from bigdl_b200.ops import functions as fn

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--iters", type=int, default=20)
a = ap.parse_args()
N, C, H, W, O, k, p = a.batch, 3, 224, 224, 64, 7, 3
x = torch.randn(N, C, H, W, device="cuda")
geom = (k, k, 2, 2, p, p, p, p, 1, 1, 1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for use in (True, False):
    fn._STEM_S2D[0] = use
    w = (torch.randn(O, C, k, k, device="cuda") * 0.1).requires_grad_()
    b = torch.zeros(O, device="cuda").requires_grad_()
    w.grad, b.grad = torch.zeros_like(w), torch.zeros_like(b)
    w16 = w.detach().bfloat16().contiguous(memory_format=torch.channels_last)
    gy = torch.randn(N, 112, 112, O, device="cuda").bfloat16().permute(0, 3, 1, 2)
    tf = tb = 0.0
    for it in range(a.iters + 3):
        flush.zero_()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        y = fn.ConvFn.apply(x, w, b, w16, geom, True, 1.0, 1.0)
        e[1].record()
        y.backward(gy)
        e[2].record()
        torch.cuda.synchronize()
        if it >= 3:
            tf += e[0].elapsed_time(e[1])
            tb += e[1].elapsed_time(e[2])
    print(json.dumps({"path": "s2d" if use else "im2col", "batch": N, "fwd_us": round(tf / a.iters * 1e3, 1),
                      "bwd_us": round(tb / a.iters * 1e3, 1)}))
