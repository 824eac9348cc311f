"""tcgen05 GEMM throughput vs cuBLAS (torch.matmul) on representative shapes; CUDA-event timed, L2 flushed between
iterations by rotating through operand sets larger than L2."""
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bigdl_b200.ops.gemm import gemm_bf16


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    shapes = [  # (M, N, K, a_mn, b_mn, label)
        (8192, 8192, 8192, False, False, "square 8192^3"),
        (4096, 4096, 4096, False, False, "square 4096^3"),
        (128 * 3136, 192, 576, False, False, "conv2/3x3 fprop"),
        (128 * 784, 128, 192, False, False, "3a/1x1 fprop"),
        (128 * 196, 208, 864, False, False, "4a/3x3 fprop"),
        (128 * 196, 864, 208, False, True, "4a/3x3 dgrad"),
        (208, 864, 128 * 196, True, True, "4a/3x3 wgrad (split-K)"),
        (128, 1000, 1024, False, False, "classifier fprop"),
        (128 * 12544, 64, 152, False, False, "stem 7x7 fprop"),
    ]
    rows = []
    for M, N, K, amn, bmn, label in shapes:
        nset = max(2, int(300e6 // ((M * K + N * K) * 2)) + 1)
        nset = min(nset, 8)
        As = [torch.randn((K, M) if amn else (M, K), device="cuda").bfloat16() for _ in range(nset)]
        Bs = [torch.randn((K, N) if bmn else (N, K), device="cuda").bfloat16() for _ in range(nset)]
        acc = amn and bmn
        out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if acc else torch.bfloat16)

        def ours(i):
            gemm_bf16(As[i % nset], Bs[i % nset], a_mn=amn, b_mn=bmn, out=out, accumulate=acc)

        def cublas(i):
            a = As[i % nset].t() if amn else As[i % nset]
            b = Bs[i % nset] if bmn else Bs[i % nset].t()
            torch.matmul(a, b)

        t_o, t_c = timeit(ours), timeit(cublas)
        fl = 2.0 * M * N * K
        rows.append({"shape": label, "M": M, "N": N, "K": K, "ours_ms": round(t_o, 4), "cublas_ms": round(t_c, 4),
                     "ours_tflops": round(fl / t_o / 1e9, 1), "cublas_tflops": round(fl / t_c / 1e9, 1)})
        print(json.dumps(rows[-1]), flush=True)


if __name__ == "__main__":
    main()
