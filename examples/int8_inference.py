"""Quantised inference end to end (reference: S/example/mkldnn/int8/{GenerateInt8Scales,ImageNetInference}.scala and
docs/docs/ProgrammingGuide/quantization-support.md):

1. load or build a float model,  2. ``model.quantize()`` (weights → e4m3 with per-output-channel scales, conv+BN folded),
3. calibrate activation scales on a few batches (``calcScales`` in the reference),  4. save the quantised model,
5. compare outputs / top-1 agreement / throughput against the float model.

    python examples/int8_inference.py [--model resnet50|vgg16|inception_v1|lenet5] [-b 32] [--modelPath x.bigdl]
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="lenet5")
    ap.add_argument("--modelPath", default=None, help="a saved .bigdl float model (overrides --model)")
    ap.add_argument("-b", "--batchSize", type=int, default=32)
    ap.add_argument("--calibBatches", type=int, default=2)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("-o", "--output", default="/tmp/quantized.bigdl")
    a = ap.parse_args(argv)
    from bigdl_b200.cli._common import build_model
    from bigdl_b200.nn.module import Module
    from bigdl_b200.quantized import Quantizer
    from bigdl_b200.utils.engine import Engine

    dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    Engine.init(device=dev.type)
    if a.modelPath:
        model, shape = Module.loadModule(a.modelPath), None
    else:
        model, shape = build_model(a.model, 10 if a.model == "lenet5" else 1000)
    shape = shape or (3, 224, 224)
    model.evaluate()

    def batch():
        x = torch.randn(a.batchSize, *shape)
        if dev.type == "cuda":
            x = x.to(dev).to(torch.bfloat16)
            if x.dim() == 4:
                x = x.contiguous(memory_format=torch.channels_last)
        return x

    fmodel = model.to(dev)
    qmodel = model.quantize().to(dev)
    n_q = len(Quantizer.quantized_layers(qmodel))
    Quantizer.calibrate(qmodel, [batch() for _ in range(a.calibBatches)])
    links = sum(1 for l in Quantizer.quantized_layers(qmodel) if l._emit_q is not None)
    print(f"{n_q} layers quantised, {links} producer→consumer pairs hand e4m3 over without a bf16 round trip")
    x = batch()
    ref, out = fmodel.forward(x).float(), qmodel.forward(x).float()
    rel = float((out - ref).norm() / ref.norm())
    agree = float((out.reshape(a.batchSize, -1).argmax(1) == ref.reshape(a.batchSize, -1).argmax(1)).float().mean())
    print(f"relative output error {rel:.4f}, top-1 agreement with the float model {agree:.3f}")

    def rate(m):
        for _ in range(2):
            m.forward(x)
        if dev.type == "cuda":
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(a.iters):
            m.forward(x)
        if dev.type == "cuda":
            torch.cuda.synchronize()
        return a.batchSize * a.iters / (time.perf_counter() - t0)
    print(f"throughput (eager, batch {a.batchSize}): float {rate(fmodel):.1f} img/s, quantised {rate(qmodel):.1f} img/s "
          f"(see benchmarks/quant_infer_bench.py for the CUDA-graph batch sweep)")
    qmodel.to("cpu").saveModule(a.output, overWrite=True)
    back = Module.loadModule(a.output)
    print(f"saved quantised model to {a.output} ({os.path.getsize(a.output) / 1e6:.2f} MB) and reloaded "
          f"{len(Quantizer.quantized_layers(back))} quantised layers")
    return rel, agree


if __name__ == "__main__":
    main()
