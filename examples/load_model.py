"""Load a model from another framework and validate / run it (reference: S/example/loadmodel/ModelValidator.scala —
caffe | torch | bigdl model types — plus S/example/tensorflow/loadandsave). With no arguments the script builds a small
CNN, exports it to every supported format and loads each one back, checking that all copies agree.

    python examples/load_model.py [-t caffe|torch|bigdl|tensorflow|onnx --modelPath ... --caffeDefPath ...]
"""
from __future__ import annotations

import argparse
import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def small_cnn():
    from bigdl_b200.nn import layer as L

    return (L.Sequential()
            .add(L.SpatialConvolution(3, 8, 3, 3, 1, 1, 1, 1).setName("conv1")).add(L.ReLU().setName("relu1"))
            .add(L.SpatialMaxPooling(2, 2, 2, 2).setName("pool1"))
            .add(L.View(8 * 8 * 8).setName("flat")).add(L.Linear(512, 10).setName("fc")).add(L.LogSoftMax().setName("prob")))


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("-t", "--modelType", default=None, choices=[None, "caffe", "torch", "bigdl", "tensorflow", "onnx"])
    ap.add_argument("--modelPath", default=None)
    ap.add_argument("--caffeDefPath", default=None)
    ap.add_argument("--tfInputs", default="input")
    ap.add_argument("--tfOutputs", default="prob")
    a = ap.parse_args(argv)
    from bigdl_b200.nn.module import Module
    from bigdl_b200.utils.engine import Engine

    Engine.init(device="cpu")
    if a.modelType:
        if a.modelType == "caffe":
            m = Module.loadCaffeModel(a.caffeDefPath, a.modelPath)
        elif a.modelType == "torch":
            m = Module.loadTorch(a.modelPath)
        elif a.modelType == "tensorflow":
            m = Module.loadTF(a.modelPath, a.tfInputs.split(","), a.tfOutputs.split(","))
        elif a.modelType == "onnx":
            from bigdl_b200.interop.onnx_loader import load

            m = load(a.modelPath)
        else:
            m = Module.loadModule(a.modelPath)
        print(m)
        return m
    model = small_cnn().evaluate()
    x = torch.randn(2, 3, 16, 16)
    ref = model.forward(x).clone()
    d = tempfile.mkdtemp()
    model.saveModule(os.path.join(d, "m.bigdl"), overWrite=True)
    model.saveTorch(os.path.join(d, "m.t7"), overWrite=True)
    model.saveCaffe(os.path.join(d, "m.prototxt"), os.path.join(d, "m.caffemodel"), overwrite=True)
    model.saveTF([("input", [None, 3, 16, 16])], os.path.join(d, "m.pb"))
    copies = {
        "bigdl": Module.loadModule(os.path.join(d, "m.bigdl")),
        "torch": Module.loadTorch(os.path.join(d, "m.t7")),
        "caffe": Module.loadCaffeModel(os.path.join(d, "m.prototxt"), os.path.join(d, "m.caffemodel")),
        "tensorflow": Module.loadTF(os.path.join(d, "m.pb"), ["input"], ["prob"]),
    }
    worst = 0.0
    for name, m in copies.items():
        inp = x.permute(0, 2, 3, 1).contiguous() if name == "tensorflow" else x
        out = m.evaluate().forward(inp)
        if name == "caffe":
            out = torch.log(out)           # caffe has Softmax, not LogSoftmax
        err = float((out - ref).abs().max())
        worst = max(worst, err)
        print(f"{name:10s} round trip: max |diff| = {err:.2e}")
    return worst


if __name__ == "__main__":
    main()
