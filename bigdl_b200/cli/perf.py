"""Layer-engine perf harness (S/nn/mkldnn/Perf.scala:40-140: ``--model resnet50|vgg16 --batchSize --iteration
--training true|false``): forward(+backward) of the model on a synthetic batch, device-timed, prints images/second.
``--int8`` (the reference's quantised Perf path) runs the fp8-quantised model for inference."""
from __future__ import annotations

import argparse
import json

import torch

from ._common import build_model


def main(argv=None):
    ap = argparse.ArgumentParser(description="Performance test of the native layer engine")
    ap.add_argument("-m", "--model", default="resnet50", choices=["resnet50", "vgg16", "vgg19", "inception_v1", "inception_v2"])
    ap.add_argument("-b", "--batchSize", type=int, default=64)
    ap.add_argument("-i", "--iteration", type=int, default=30)
    ap.add_argument("-t", "--training", default="true", choices=["true", "false"])
    ap.add_argument("--int8", action="store_true", help="quantised (fp8 e4m3) inference")
    a = ap.parse_args(argv)
    if not torch.cuda.is_available():
        raise SystemExit("bigdl_b200.cli.perf needs a GPU (use local_optimizer_perf --device cpu for CPU runs)")
    from ..nn.criterion import ClassNLLCriterion
    from ..utils.engine import Engine

    Engine.init()
    dev = torch.device("cuda")
    model, shape = build_model(a.model)
    training = a.training == "true" and not a.int8
    x = torch.rand(a.batchSize, *shape, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = torch.ones(a.batchSize, device=dev)
    if a.int8:
        from ..quantized import Quantizer

        model = model.evaluate().quantize().to(dev)
        Quantizer.calibrate(model, [x])
    else:
        model = model.to(dev)
        model.training() if training else model.evaluate()
    crit = ClassNLLCriterion()

    def step():
        if training:
            model.zeroGradParameters()
            out = model.forward(x)
            crit.forward(out, y)
            model.backward(x, crit.backward(out, y))
        else:
            model._refresh_compute_weights()
            with torch.no_grad():
                model._call(x)

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iteration):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iteration
    print(json.dumps({"tool": "Perf", "model": a.model, "training": training, "int8_as_fp8": a.int8,
                      "batch": a.batchSize, "ms_per_iteration": round(ms, 3),
                      "images_per_second": round(a.batchSize / ms * 1e3, 1)}))


if __name__ == "__main__":
    main()
