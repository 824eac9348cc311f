"""Quantised-inference benchmark (BASELINE config #5): ResNet-50 ImageNet forward throughput, bf16 (BN folded) vs fp8
e4m3 (static scales, requantising epilogues), over a batch sweep. The reference's counterpart is the int8 whitepaper
table (docs/docs/whitepaper.md: int8 vs fp32 inference speed-up) and S/example/mkldnn/int8/ImageNetInference.scala.

    python benchmarks/quant_infer_bench.py --batches 1,8,32,128 --iters 30

Prints one JSON line per (batch, mode): img/s from CUDA events after warm-up; inputs are rotated over a pool larger
than L2 so no iteration re-reads a cached input. Random-init weights, synthetic data.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def build(model_name: str):
    if model_name == "resnet50":
        from bigdl_b200.models.resnet import ResNet, model_init

        m = ResNet(1000, depth=50, shortcut_type="B", dataset="imagenet")
        model_init(m)
        return m, (3, 224, 224)
    if model_name == "vgg16":
        from bigdl_b200.models.vgg import Vgg_16

        return Vgg_16(1000, has_dropout=False), (3, 224, 224)
    if model_name == "inception_v1":
        from bigdl_b200.models.inception import Inception_v1_NoAuxClassifier

        return Inception_v1_NoAuxClassifier(1000, has_dropout=False), (3, 224, 224)
    raise ValueError(model_name)


def time_forward(model, pool, iters, warmup, graph):
    from bigdl_b200.ops import _lib

    def run(x):
        with torch.no_grad():
            return model._call(x)

    for i in range(warmup):
        out = run(pool[i % len(pool)])
    torch.cuda.synchronize()
    n0 = _lib.LAUNCHES
    run(pool[0])
    launches = _lib.LAUNCHES - n0
    if graph:
        static_in = torch.empty_like(pool[0])
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            static_in.copy_(pool[0])
            run(static_in)
        torch.cuda.current_stream().wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            out = run(static_in)
        for i in range(3):
            static_in.copy_(pool[i % len(pool)])
            g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        if graph:
            static_in.copy_(pool[i % len(pool)], non_blocking=True)
            g.replay()
        else:
            out = run(pool[i % len(pool)])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters, launches, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batches", default="1,8,32,64,128,256")
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--modes", default="bf16,fp8")
    args = ap.parse_args()
    from bigdl_b200.nn.fusion import fold_bn, fuse
    from bigdl_b200.quantized import Quantizer
    from bigdl_b200.quantized.quantizer import _fuse_through_identities

    torch.manual_seed(0)
    fm, shape = build(args.model)
    fm.evaluate()
    q = fm.quantize().to("cuda")
    bf = fm.cloneModule().evaluate()
    fold_bn(bf)
    bf = bf.to("cuda")
    fuse(bf)
    _fuse_through_identities(bf)
    bf._refresh_compute_weights()
    dev = torch.device("cuda")
    calib = [torch.randn(16, *shape, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
             for _ in range(2)]
    Quantizer.calibrate(q, calib)
    links = sum(1 for l in Quantizer.quantized_layers(q) if l._emit_q is not None)
    for b in [int(v) for v in args.batches.split(",")]:
        per = b * shape[0] * shape[1] * shape[2] * 2
        npool = max(2, min(64, (160 << 20) // per + 1))       # > 126 MB L2 in total
        pool = [torch.randn(b, *shape, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
                for _ in range(npool)]
        res = {}
        for name, model in (("bf16", bf), ("fp8", q)):
            if name not in args.modes.split(","):
                continue
            ms, launches, out = time_forward(model, pool, args.iters, args.warmup, not args.no_graph)
            res[name] = (ms, out)
            print(json.dumps({"bench": "quant_infer", "model": args.model, "mode": name, "batch": b,
                              "ms": round(ms, 4), "img_per_s": round(b / ms * 1e3, 1), "launches": launches,
                              "cuda_graph": not args.no_graph, "fp8_chain_links": links, "input_pool": npool,
                              "data": "synthetic, random-init weights"}), flush=True)
        if len(res) < 2:
            continue
        o16, o8 = res["bf16"][1].float().reshape(b, -1), res["fp8"][1].float().reshape(b, -1)
        print(json.dumps({"bench": "quant_infer", "batch": b, "speedup_fp8_vs_bf16": round(res["bf16"][0] / res["fp8"][0], 3),
                          "rel_err_fp8_vs_bf16": round(float((o8 - o16).norm() / o16.norm()), 4),
                          "norm_bf16": float(o16.norm()), "norm_fp8": float(o8.norm()), "top1_agree": round(float((o8.argmax(1) == o16.argmax(1)).float().mean()), 4)}), flush=True)


if __name__ == "__main__":
    main()
