"""Regenerate the API lists under docs/APIGuide from the code, so they cannot drift:
    python docs/gen_api_lists.py
(reference counterparts: docs/docs/APIGuide/{Layers/*,Losses.md,tensorflow_ops_list.md,caffe_layer_list.md,ONNX.md})"""
import inspect
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "docs", "APIGuide")


def first_line(obj):
    d = inspect.getdoc(obj) or ""
    return d.split("\n")[0][:140]


def sig(cls):
    try:
        return str(inspect.signature(cls.__init__)).replace("(self, ", "(").replace("(self)", "()")
    except (TypeError, ValueError):
        return "(...)"


def main():
    from bigdl_b200.nn import criterion as C
    from bigdl_b200.nn import keras as K
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn import ops as O
    from bigdl_b200.nn.module import Module
    from bigdl_b200.interop import caffe, onnx_loader, tensorflow
    from bigdl_b200.optim import optim_method as OM
    from bigdl_b200.optim import validation as V

    def classes(mod, base):
        return sorted((n, c) for n, c in vars(mod).items() if inspect.isclass(c) and issubclass(c, base)
                      and not n.startswith("_"))

    with open(os.path.join(OUT, "Layers.md"), "w") as f:
        f.write("# Layers (`bigdl_b200.nn.layer`)\n\n| layer | constructor | summary |\n|---|---|---|\n")
        for n, c in classes(L, Module):
            f.write(f"| `{n}` | `{sig(c)[:160]}` | {first_line(c)} |\n")
    with open(os.path.join(OUT, "Losses.md"), "w") as f:
        f.write("# Criterions (`bigdl_b200.nn.criterion`)\n\n| criterion | constructor | summary |\n|---|---|---|\n")
        for n, c in classes(C, C.Criterion):
            f.write(f"| `{n}` | `{sig(c)[:160]}` | {first_line(c)} |\n")
    with open(os.path.join(OUT, "KerasLayers.md"), "w") as f:
        f.write("# Keras-style layers (`bigdl_b200.nn.keras`)\n\n| layer | constructor |\n|---|---|\n")
        for n, c in classes(K.layer, K.KerasLayer):
            f.write(f"| `{n}` | `{sig(c)[:200]}` |\n")
    with open(os.path.join(OUT, "Operations.md"), "w") as f:
        f.write("# Forward-only operations (`bigdl_b200.nn.ops`)\n\n" + ", ".join(f"`{n}`" for n, _ in classes(O, O.Operation)) + "\n")
    with open(os.path.join(OUT, "Optimizers.md"), "w") as f:
        f.write("# Optim methods and learning-rate schedules (`bigdl_b200.optim.optim_method`)\n\n")
        f.write("## Optim methods\n\n" + "\n".join(f"* `{n}{sig(c)[:200]}`" for n, c in classes(OM, OM.OptimMethod)) + "\n\n")
        f.write("## Schedules\n\n" + "\n".join(f"* `{n}{sig(c)}`" for n, c in classes(OM, OM.LearningRateSchedule)) + "\n\n")
        f.write("## Validation methods (`bigdl_b200.optim.validation`)\n\n" +
                "\n".join(f"* `{n}`" for n, c in classes(V, V.ValidationMethod)) + "\n")
    with open(os.path.join(OUT, "tensorflow_ops_list.md"), "w") as f:
        f.write(f"# TensorFlow ops understood by `Module.loadTF` ({len(tensorflow.SUPPORTED_OPS)})\n\n" +
                ", ".join(f"`{o}`" for o in tensorflow.SUPPORTED_OPS) + "\n")
    with open(os.path.join(OUT, "caffe_layer_list.md"), "w") as f:
        types = sorted(n[1:] for n in dir(caffe._Converter) if n.startswith("_") and not n.startswith("__")
                       and n not in ("_copy",))
        f.write("# Caffe layer types understood by `Module.loadCaffeModel`\n\n" + ", ".join(f"`{t}`" for t in types) +
                "\n\nplus data layers (`Data`, `Input`, `DummyData`, `MemoryData`, `ImageData`, `HDF5Data`), `Slice`, loss "
                "layers (`SoftmaxWithLoss`, `EuclideanLoss`) mapped to criterions; `Accuracy`/`Silence` are dropped.\n")
    with open(os.path.join(OUT, "ONNX.md"), "w") as f:
        f.write(f"# ONNX operators understood by `bigdl_b200.interop.onnx_loader` ({len(onnx_loader.SUPPORTED_OPS)})\n\n" +
                ", ".join(f"`{o}`" for o in onnx_loader.SUPPORTED_OPS) + "\n")
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
