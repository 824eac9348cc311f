"""Micro-benchmark of the memory-bound training kernels against the HBM roofline (bytes moved / time).
    python benchmarks/eltwise_bench.py            # prints one JSON line per (kernel, shape)
Shapes are the Inception-v1 / ResNet-50 batch-128 activations that dominate the non-GEMM part of a training step."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3          # us


def main():
    from bigdl_b200.ops import elementwise as ew

    dev = torch.device("cuda")
    shapes = [(1605632, 64, 64), (401408, 192, 192), (100352, 128, 256), (100352, 32, 256), (25088, 208, 512),
              (25088, 48, 512), (6272, 832, 832), (6272, 128, 1024)]
    for M, C, pitch in shapes:
        big = torch.randn(M, pitch, device=dev).to(torch.bfloat16)
        ybig = torch.randn(M, pitch, device=dev).to(torch.bfloat16)
        dy, y = big[:, :C], ybig[:, :C]
        gb = torch.zeros(C, device=dev)
        # rotate over copies larger than L2 for the small shapes
        n_copy = max(1, (200 << 20) // (M * pitch * 2 * 2) + 1)
        dys = [big.clone()[:, :C] for _ in range(n_copy)]
        ys = [ybig.clone()[:, :C] for _ in range(n_copy)]
        k = [0]

        def run():
            i = k[0] % n_copy
            k[0] += 1
            ew.colsum(dys[i], gb, ys[i], dys[i])
        us = timeit(run)
        bytes_moved = M * C * 2 * 3
        print(json.dumps({"kernel": "colsum+relu_bwd", "M": M, "C": C, "pitch": pitch, "us": round(us, 1),
                          "GBps": round(bytes_moved / us / 1e3, 1), "bps": os.environ.get("BDL_COLSUM_BPS", "default")}),
              flush=True)


def bn_main():
    from bigdl_b200.ops import _lib

    lib = _lib.cuda_lib()
    dev = torch.device("cuda")
    for M, C in [(200704, 64), (200704, 256), (50176, 128), (50176, 512), (12544, 256), (12544, 1024), (3136, 512),
                 (3136, 2048)]:
        n_copy = max(1, (200 << 20) // (M * C * 2) + 1)
        xs = [torch.randn(M, C, device=dev).to(torch.bfloat16) for _ in range(n_copy)]
        dys = [torch.randn(M, C, device=dev).to(torch.bfloat16) for _ in range(min(n_copy, 4))]
        sums = torch.zeros(2 * C, device=dev)
        mean, inv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        k = [0]

        def fwd():
            i = k[0] % n_copy
            k[0] += 1
            _lib.check(lib.bdl_bn_stats(xs[i].data_ptr(), sums.data_ptr(), M, C, C, _lib.stream_ptr()), "bn_stats")

        def bwd():
            i = k[0] % n_copy
            k[0] += 1
            _lib.check(lib.bdl_bn_bwd_reduce(dys[i % len(dys)].data_ptr(), xs[i].data_ptr(), None, mean.data_ptr(),
                                             inv.data_ptr(), sums.data_ptr(), M, C, C, C, 0, _lib.stream_ptr()), "bn_bwd")
        for name, fn, passes in (("bn_stats", fwd, 1), ("bn_bwd_reduce", bwd, 2)):
            us = timeit(fn)
            print(json.dumps({"kernel": name, "M": M, "C": C, "us": round(us, 1),
                              "GBps": round(M * C * 2 * passes / us / 1e3, 1), "bps": os.environ.get("BDL_BN_BPS", "default")}),
                  flush=True)


if __name__ == "__main__":
    if "--bn" in sys.argv:
        bn_main()
    else:
        main()
