"""One convolution layer forward+backward on the GPU path (for ncu captures and per-layer timing)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bigdl_b200.nn.layer import SpatialConvolution


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=64)
    ap.add_argument("--c", type=int, default=256)
    ap.add_argument("--o", type=int, default=256)
    ap.add_argument("--hw", type=int, default=56)
    ap.add_argument("--k", type=int, default=3)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    conv = SpatialConvolution(a.c, a.o, a.k, a.k, 1, 1, a.k // 2, a.k // 2).cuda()
    xs = [torch.randn(a.n, a.c, a.hw, a.hw, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
          .requires_grad_(True) for _ in range(3)]
    go = torch.randn(a.n, a.o, a.hw, a.hw, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    conv._refresh_compute_weights()
    def step(i):
        y = conv._call(xs[i % 3])
        y.backward(go)
    for i in range(2):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.iters):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    fl = 3 * 2.0 * a.n * a.hw * a.hw * a.o * a.c * a.k * a.k
    print(json.dumps({"shape": vars(a), "ms_fwd_bwd": round(ms, 4), "tflops": round(fl / ms / 1e9, 1)}))


if __name__ == "__main__":
    main()
