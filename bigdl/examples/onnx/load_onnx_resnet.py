"""pyspark/bigdl/examples/onnx/load_onnx_resnet.py: load an ONNX ResNet and forward one random batch."""
import sys


def load_onnx_resnet(path="resnet50.onnx", shape=(1, 3, 224, 224)):
    import torch
    from bigdl.contrib.onnx import load
    model = load(path).evaluate()
    return model.forward(torch.randn(*shape))


if __name__ == "__main__":
    print(load_onnx_resnet(*(sys.argv[1:2] or ["resnet50.onnx"])))
