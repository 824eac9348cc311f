"""Model format converter (S/utils/ConvertModel.scala:30-133):

    python -m bigdl_b200.cli.convert_model --from caffe --to bigdl --input net.caffemodel --prototxt net.prototxt \\
        --output net.bigdl [--quantize true]

from ∈ {bigdl, caffe, torch, tensorflow, onnx, keras}; to ∈ {bigdl, caffe, torch, tensorflow}. TensorFlow needs
``--tf_inputs a,b --tf_outputs c`` on import and ``--tf_inputs name:N,C,H,W`` on export."""
from __future__ import annotations

import argparse


def load_any(fmt, a):
    from ..nn.module import Module

    if fmt == "bigdl":
        return Module.loadModule(a.input)
    if fmt == "caffe":
        if not a.prototxt:
            raise SystemExit("caffe input needs --prototxt")
        return Module.loadCaffeModel(a.prototxt, a.input)
    if fmt == "torch":
        return Module.loadTorch(a.input)
    if fmt == "tensorflow":
        if not (a.tf_inputs and a.tf_outputs):
            raise SystemExit("tensorflow input needs --tf_inputs and --tf_outputs")
        return Module.loadTF(a.input, a.tf_inputs.split(","), a.tf_outputs.split(","))
    if fmt == "onnx":
        from ..interop.onnx_loader import load

        return load(a.input)
    if fmt == "keras":
        return Module.load_keras(a.input, a.keras_weights)
    raise SystemExit(f"unsupported source format {fmt}")


def main(argv=None):
    ap = argparse.ArgumentParser(description="Convert models between different dl frameworks")
    ap.add_argument("--from", dest="src", required=True, choices=["bigdl", "caffe", "torch", "tensorflow", "onnx", "keras"])
    ap.add_argument("--to", dest="dst", required=True, choices=["bigdl", "caffe", "torch", "tensorflow"])
    ap.add_argument("--input", required=True)
    ap.add_argument("--output", required=True)
    ap.add_argument("--prototxt", default=None)
    ap.add_argument("--keras_weights", default=None)
    ap.add_argument("--tf_inputs", default=None)
    ap.add_argument("--tf_outputs", default=None)
    ap.add_argument("--quantize", default="false", choices=["true", "false"])
    a = ap.parse_args(argv)
    model = load_any(a.src, a)
    if a.quantize == "true":
        if a.dst != "bigdl":
            raise SystemExit("only support quantizing models saved in the bigdl format")
        model = model.quantize()
    if a.dst == "bigdl":
        model.saveModule(a.output, overWrite=True)
    elif a.dst == "torch":
        model.saveTorch(a.output, overWrite=True)
    elif a.dst == "caffe":
        if not a.prototxt:
            raise SystemExit("caffe output needs --prototxt (path of the definition file to write)")
        model.saveCaffe(a.prototxt, a.output, overwrite=True)
    elif a.dst == "tensorflow":
        if not a.tf_inputs:
            raise SystemExit("tensorflow output needs --tf_inputs name:N,C,H,W")
        ins = []
        for spec in a.tf_inputs.split(";"):
            name, _, shp = spec.partition(":")
            ins.append((name, [None if s in ("", "-1", "None") else int(s) for s in shp.split(",")]))
        model.saveTF(ins, a.output)
    print(f"converted {a.src} model {a.input} → {a.dst} {a.output}")


class ConvertModel:
    """``ConvertModel.main(args)`` (S/utils/ConvertModel.scala)."""
    main = staticmethod(main)


if __name__ == "__main__":
    main()
