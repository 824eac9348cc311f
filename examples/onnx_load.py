"""Load an ONNX model and run it (reference: pyspark/bigdl/examples/onnx/load_onnx_resnet.py — `load("resnet.onnx")`,
forward a random batch). Without --model the script writes a small residual CNN as an .onnx file itself (protobuf wire
format, `interop/onnx_pb.py`; no onnx package needed), loads it back and checks the output against PyTorch.

    python examples/onnx_load.py [--model resnet50.onnx --shape 1,3,224,224]
"""
from __future__ import annotations

import argparse
import os
import sys
import tempfile

import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def write_small_resnet(path: str, seed: int = 0):
    """conv-bn-relu -> residual block (conv-bn-relu-conv-bn, + shortcut, relu) -> global average pool -> fc -> softmax"""
    from bigdl_b200.interop import onnx_pb as OP

    rs = np.random.RandomState(seed)
    P = {}

    def conv(name, cout, cin, k):
        P[name + "_w"] = (rs.randn(cout, cin, k, k) / np.sqrt(cin * k * k)).astype(np.float32)

    def bn(name, c):
        P[name + "_s"], P[name + "_b"] = (rs.rand(c) + 0.5).astype(np.float32), rs.randn(c).astype(np.float32) * 0.1
        P[name + "_m"], P[name + "_v"] = rs.randn(c).astype(np.float32) * 0.1, (rs.rand(c) + 0.5).astype(np.float32)
    conv("c0", 8, 3, 3); bn("b0", 8); conv("c1", 8, 8, 3); bn("b1", 8); conv("c2", 8, 8, 3); bn("b2", 8)
    P["fc_w"], P["fc_b"] = rs.randn(10, 8).astype(np.float32), rs.randn(10).astype(np.float32)
    m = OP.ModelProto()
    m.ir_version = 7
    m.opset_import.add().version = 11
    g = m.graph
    g.input.add().name = "data"
    for name, arr in P.items():
        g.initializer.append(OP.from_numpy(arr, name))

    def node(op_type, ins, outs, **attrs):
        n = g.node.add()
        n.op_type = op_type
        n.input.extend(ins)
        n.output.extend(outs)
        for k, v in attrs.items():
            a = n.attribute.add()
            a.name = k
            if isinstance(v, list):
                a.ints.extend(v)
            elif isinstance(v, float):
                a.f = v
            else:
                a.i = v

    def cbr(x, c, b, out, relu=True):
        node("Conv", [x, c + "_w"], [out + "_c"], kernel_shape=[3, 3], pads=[1, 1, 1, 1], strides=[1, 1])
        node("BatchNormalization", [out + "_c", b + "_s", b + "_b", b + "_m", b + "_v"], [out + "_n"], epsilon=1e-5)
        if relu:
            node("Relu", [out + "_n"], [out])
            return out
        return out + "_n"
    x = cbr("data", "c0", "b0", "stem")
    y = cbr(x, "c1", "b1", "blk1")
    y = cbr(y, "c2", "b2", "blk2", relu=False)
    node("Sum", [y, x], ["res"])
    node("Relu", ["res"], ["res_r"])
    node("GlobalAveragePool", ["res_r"], ["gap"])
    node("Flatten", ["gap"], ["flat"], axis=1)
    node("Gemm", ["flat", "fc_w", "fc_b"], ["logits"], transB=1)
    node("Softmax", ["logits"], ["prob"], axis=1)
    g.output.add().name = "prob"
    with open(path, "wb") as f:
        f.write(m.SerializeToString())
    return P


def torch_reference(P, x):
    t = {k: torch.from_numpy(v) for k, v in P.items()}

    def cbr(h, c, b, relu=True):
        h = F.conv2d(h, t[c + "_w"], None, 1, 1)
        h = F.batch_norm(h, t[b + "_m"], t[b + "_v"], t[b + "_s"], t[b + "_b"], False, 0.0, 1e-5)
        return torch.relu(h) if relu else h
    s = cbr(x, "c0", "b0")
    y = cbr(cbr(s, "c1", "b1"), "c2", "b2", relu=False)
    h = torch.relu(y + s).mean((2, 3))
    return torch.softmax(h @ t["fc_w"].t() + t["fc_b"], 1)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default=None, help="path of an .onnx file")
    ap.add_argument("--shape", default="2,3,16,16")
    a = ap.parse_args(argv)
    from bigdl_b200.interop.onnx_loader import load
    from bigdl_b200.utils.engine import Engine

    Engine.init(device="cpu")
    shape = [int(v) for v in a.shape.split(",")]
    x = torch.randn(*shape)
    if a.model:
        model = load(a.model).evaluate()
        y = model.forward(x)
        print("output", tuple(y.shape) if hasattr(y, "shape") else y)
        return 0.0
    path = os.path.join(tempfile.mkdtemp(), "small_resnet.onnx")
    params = write_small_resnet(path)
    model = load(path).evaluate()
    y = model.forward(x)
    err = float((y - torch_reference(params, x)).abs().max())
    print(f"loaded {os.path.getsize(path)} bytes of ONNX, output {tuple(y.shape)}, max |ours - torch| = {err:.2e}")
    return err


if __name__ == "__main__":
    main()
