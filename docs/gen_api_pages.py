"""Generate the per-topic API reference pages from the code (constructor signatures + docstrings), so the pages cannot
drift from the implementation:      python docs/gen_api_pages.py
Writes docs/APIGuide/Layers/*.md, Optimizers/*.md, Module.md, Tensor.md, Data.md, Transformer.md, Triggers.md, Metrics.md,
Regularizers.md, Initializers.md, Engine.md and docs/KerasStyleAPIGuide/*.md — the page set of the reference's
docs/docs/APIGuide and KerasStyleAPIGuide trees."""
import inspect
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
DOCS = os.path.join(ROOT, "docs")


def sig(obj):
    try:
        s = str(inspect.signature(obj.__init__ if inspect.isclass(obj) else obj))
    except (TypeError, ValueError):
        return "(...)"
    return s.replace("(self, ", "(").replace("(self)", "()")


def doc(obj):
    return (inspect.getdoc(obj) or "").strip()


def own_classes(mod, base=None):
    out = []
    for n, c in vars(mod).items():
        if inspect.isclass(c) and not n.startswith("_") and getattr(c, "__module__", None) == mod.__name__:
            if base is None or issubclass(c, base):
                out.append((n, c))
    return sorted(out)


def class_section(n, c, methods=False):
    lines = [f"## {n}\n", f"```python\n{n}{sig(c)}\n```\n"]
    d = doc(c)
    if d:
        lines.append(d + "\n")
    if methods:
        for mn, m in sorted(vars(c).items()):
            if mn.startswith("_") or not callable(m):
                continue
            md = doc(m).split("\n\n")[0]
            lines.append(f"* `{mn}{sig(m)}`" + (f" — {md}" if md else ""))
        lines.append("")
    return "\n".join(lines)


def write(path, title, intro, sections):
    path = os.path.join(DOCS, path)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        f.write(f"# {title}\n\n{intro}\n\n" + "\n".join(sections) + "\n")
    return path


def main():
    import importlib
    from bigdl_b200.nn.module import Module
    sys.path.insert(0, os.path.join(ROOT, "docs"))
    import gen_api_lists
    gen_api_lists.main()                 # the flat name lists (Layers.md table, op lists); pages below refine them
    written = []
    # ---------------------------------------------------------------- layers, grouped like the reference's pages
    groups = {
        "Activations": (["layers_activation"], "Element-wise activations. `ReLU` runs on the native kernels (and fuses into the producing convolution / Linear / BatchNorm); the others are torch element-wise ops."),
        "Convolution-Layers": (["layers_conv_extra"], "Convolution family beyond `SpatialConvolution` (which lives in `layers_core`, on the tcgen05 implicit-GEMM kernels): dilated, full (transposed), separable, shared, map, volumetric, temporal, locally-connected, up-sampling, cropping."),
        "Simple-Layers": (["layers_core", "layers_linear"], "The core layers: `Linear`, `SpatialConvolution`, `SpatialBatchNormalization`, pooling, `SpatialCrossMapLRN`, `Dropout`, `SoftMax` / `LogSoftMax`, reshapes — the ones with native sm_100a kernels — and the Linear family (`Bilinear`, `Cosine`, `Euclidean`, `LookupTable`, `Maxout`, `Highway`, sparse variants)."),
        "Recurrent-Layers": (["layers_recurrent"], "`Recurrent` / `BiRecurrent` / `RecurrentDecoder` containers and the cells. `Recurrent(LSTM)` runs as ONE persistent kernel per sequence (`csrc/lstm_seq.cu`, `bigdl.rnn.persistent`)."),
        "MergeSplit-Layers": (["layers_table"], "Containers over tables (`ConcatTable`, `ParallelTable`, `MapTable`, `Bottle`), table arithmetic (`CAddTable`, `CMulTable`, ...), selection / narrowing / joining and the shape utilities."),
        "Attention-Layers": (["layers_attention"], "`Attention`, `FeedForwardNetwork`, `Transformer`, `SequenceBeamSearch`, `TransformerOperation`. `Attention` runs the fused tcgen05 kernel (`csrc/attention.cu`) for head sizes 64 / 128."),
        "Detection-Layers": (["layers_detection"], "Object-detection layers: anchors, proposals, `RoiAlign` / `RoiPooling`, `BoxHead`, `MaskHead`, `RegionProposal`, `FPN`, `DetectionOutputSSD` / `DetectionOutputFrcnn`, `Nms`, `PriorBox`."),
        "TensorFlow-Layers": (["tf"], "Control-flow and data ops that loaded TensorFlow graphs need (`Switch`, `Merge`, `Enter`, `Exit`, `NextIteration`, `TensorArray*`, `Stack*`, `ParseExample`, `DecodeImage`, ...)."),
        "ONNX-Layers": (["onnx"], "Thin operator layers with ONNX attribute conventions (`Gemm`, `Reshape`, `Shape`, ...)."),
        "Quantized-Layers": ([], ""),
    }
    for page, (mods, intro) in groups.items():
        if not mods:
            continue
        secs = []
        for m in mods:
            mod = importlib.import_module(f"bigdl_b200.nn.{m}")
            secs += [class_section(n, c) for n, c in own_classes(mod, Module)]
        written.append(write(f"APIGuide/Layers/{page}.md", page.replace("-", " "), intro + f"\n\nSource: " + ", ".join(f"`bigdl_b200/nn/{m}.py`" for m in mods) + ".", secs))
    # ---- the reference's topic pages that cut ACROSS source files: classes picked by name from nn.layer
    from bigdl_b200.nn import layer as NL
    topics = {
        "Distance-Layers": ("Layers that compare two inputs or an input with learned centres.",
                            ["Cosine", "CosineDistance", "Euclidean", "PairwiseDistance", "DotProduct", "CrossProduct"]),
        "Dropout-Layers": ("Noise layers; all are identity in evaluate mode. `Dropout` uses the Philox-counter kernel so "
                           "a captured CUDA graph draws fresh masks on every replay.",
                           ["Dropout", "GaussianDropout", "GaussianNoise", "SpatialDropout1D", "SpatialDropout2D",
                            "SpatialDropout3D", "GaussianSampler"]),
        "Embedding-Layers": ("Index → vector look-ups (1-based indices).", ["LookupTable", "LookupTableSparse"]),
        "Math-Layers": ("Element-wise and reduction arithmetic as layers.",
                        ["Scale", "Min", "Max", "Add", "CAdd", "CMul", "Mul", "MulConstant", "AddConstant", "Bilinear",
                         "Clamp", "Square", "Sqrt", "Power", "Abs", "Log", "Exp", "Mean", "Sum", "MM", "MV", "Negative",
                         "CAveTable", "CDivTable", "CMaxTable", "CMinTable", "CSubTable", "Threshold", "BinaryThreshold"]),
        "Normalization-Layers": ("Batch / layer / local-response normalisation. `SpatialBatchNormalization` runs the "
                                 "fused kernels of `csrc/batchnorm.cu` (statistics from the producing convolution's "
                                 "epilogue, residual add + ReLU fused forward and backward).",
                                 ["BatchNormalization", "SpatialBatchNormalization", "VolumetricBatchNormalization"
                                  if hasattr(NL, "VolumetricBatchNormalization") else "LayerNormalization",
                                  "LayerNormalization", "SpatialCrossMapLRN", "SpatialWithinChannelLRN",
                                  "SpatialSubtractiveNormalization", "SpatialDivisiveNormalization",
                                  "SpatialContrastiveNormalization", "Normalize", "NormalizeScale"]),
        "Padding-Layers": ("Zero / constant padding.", ["Padding", "SpatialZeroPadding"]),
        "Pooling-Layers": ("Max / average pooling in 1, 2 and 3 dimensions, RoI pooling and the up-sampling inverses.",
                           ["SpatialMaxPooling", "SpatialAveragePooling", "VolumetricMaxPooling",
                            "VolumetricAveragePooling", "TemporalMaxPooling", "RoiPooling", "RoiAlign", "Pooler",
                            "UpSampling1D", "UpSampling2D", "UpSampling3D", "ResizeBilinear"]),
        "Recursive-Layers": ("Tree-structured networks: a `TensorTree` describes each tree, `BinaryTreeLSTM` runs the "
                             "constituency Tree-LSTM over it.", ["TensorTree", "TreeLSTM", "BinaryTreeLSTM"]),
        "Sparse-Layers": ("Layers over `SparseTensor` inputs (wide & deep style models).",
                          ["SparseLinear", "SparseJoinTable", "DenseToSparse", "LookupTableSparse"]),
        "Utilities": ("Plumbing layers: placeholders, debugging, gradient control, shape utilities.",
                      ["Input", "Identity", "Echo", "Contiguous", "GradientReversal", "Masking", "Reverse", "Replicate",
                       "Tile", "ExpandSize", "Squeeze", "Unsqueeze", "Transpose", "Select", "Narrow", "Index",
                       "InferReshape", "Reshape", "View", "Pack", "ConstInitMethod" if hasattr(NL, "ConstInitMethod") else "Pack"]),
    }
    for page, (intro, names) in topics.items():
        secs, seen = [], set()
        for n in names:
            c = getattr(NL, n, None)
            if c is None or n in seen:
                continue
            seen.add(n)
            secs.append(class_section(n, c) if inspect.isclass(c) else f"## {n}\n\n```python\n{n}{sig(c)}\n```\n\n{doc(c)}\n")
        written.append(write(f"APIGuide/Layers/{page}.md", page.replace("-", " "), intro, secs))
    from bigdl_b200.quantized import layers as QL
    written.append(write("APIGuide/Layers/Quantized-Layers.md", "Quantized layers",
                         "fp8 (e4m3) inference layers produced by `Quantizer.quantize(model)`; per-tensor x per-channel scales or MX block scales (`block_scale=True`). Source: `bigdl_b200/quantized/layers.py`.",
                         [class_section(n, c) for n, c in own_classes(QL)]))
    from bigdl_b200.nn import module as M
    written.append(write("APIGuide/Layers/Containers.md", "Containers", "`Sequential`, `Concat`, `Graph` (static), `DynamicGraph`; source `bigdl_b200/nn/module.py`, `nn/dynamic_graph.py`.",
                         [class_section(n, c) for n, c in own_classes(M, Module) if n != "Module"]))
    written.append(write("APIGuide/Module.md", "Module (AbstractModule)", "Every layer and container derives from `Module` (S/nn/abstractnn/AbstractModule.scala).",
                         [class_section("Module", M.Module, methods=True)]))
    # ---------------------------------------------------------------- criterions, optimizers
    from bigdl_b200.nn import criterion as C
    written.append(write("APIGuide/Losses.md", "Criterions (`bigdl_b200.nn.criterion`)", "Loss functions; `ClassNLLCriterion` / `CrossEntropyCriterion` use the fused xent kernels on the GPU.",
                         [class_section(n, c) for n, c in own_classes(C, C.Criterion)]))
    from bigdl_b200.optim import optim_method as OM
    written.append(write("APIGuide/Optimizers/Optim-Methods.md", "Optim methods", "Update rules; on the GPU all of them run in the flat multi-tensor kernel (`csrc/optim.cu`) or inside the fused reduce-scatter + update kernel (`csrc/comm_fused.cu`).",
                         [class_section(n, c) for n, c in own_classes(OM, OM.OptimMethod)]))
    written.append(write("APIGuide/Optimizers/OptimMethod.md", "OptimMethod",
                         "The base class of every update rule: `optimize(feval, x)` for closures, `update(x, g)` for "
                         "flat vectors, `state` (epoch, neval, learning-rate bookkeeping), `save` / `load`, "
                         "`getLearningRate`, `clearHistory`, hyper-parameter tables for per-layer rates.",
                         [class_section("OptimMethod", OM.OptimMethod, methods=True), class_section("LBFGS", OM.LBFGS, True),
                          class_section("LineSearch", OM.LineSearch), class_section("WolfeLineSearch", OM.WolfeLineSearch)]))
    written.append(write("APIGuide/Optimizers/Learningrate-Scheduler.md", "Learning-rate schedules", "Attached with `SGD(learningrate_schedule=...)`; evaluated on the host each iteration and written into the device-resident hyper-parameter block, so CUDA-graph replays pick them up.",
                         [class_section(n, c) for n, c in own_classes(OM, OM.LearningRateSchedule)]))
    from bigdl_b200.optim import optimizer as OPT
    written.append(write("APIGuide/Optimizers/Optimizer.md", "Optimizer", "`Optimizer(model, dataset, criterion)` factory, `LocalOptimizer`, `DistriOptimizer`; setters mirror S/optim/Optimizer.scala.",
                         [class_section(n, c, methods=True) for n, c in own_classes(OPT)]))
    from bigdl_b200.optim import trigger as T, validation as V, regularizer as R
    written.append(write("APIGuide/Triggers.md", "Triggers", "When to validate / checkpoint / stop (S/optim/Trigger.scala).", [class_section(n, c, True) for n, c in own_classes(T)]))
    written.append(write("APIGuide/Metrics.md", "Validation methods", "Metrics for `setValidation` / `Evaluator` (S/optim/ValidationMethod.scala).", [class_section(n, c) for n, c in own_classes(V)]))
    written.append(write("APIGuide/Regularizers.md", "Regularizers", "Per-layer weight / bias regularizers (S/optim/Regularizer.scala).", [class_section(n, c) for n, c in own_classes(R)]))
    from bigdl_b200.nn import initialization_method as IM
    written.append(write("APIGuide/Initializers.md", "Initialization methods", "`layer.setInitMethod(weightInitMethod, biasInitMethod)` (S/nn/InitializationMethod.scala).", [class_section(n, c) for n, c in own_classes(IM)]))
    # ---------------------------------------------------------------- data, tensor, engine
    from bigdl_b200.dataset import base as DB, image as DI, text as DT
    written.append(write("APIGuide/Data.md", "Data", "`DataSet`, `Sample`, `MiniBatch`, `Transformer` chains and the image / text pipelines.",
                         [class_section(n, c) for mod in (DB, DI, DT) for n, c in own_classes(mod)]))
    from bigdl_b200.transform.vision import image as VI
    written.append(write("APIGuide/Transformer.md", "Vision transformers", "`ImageFeature` / `ImageFrame` and the augmentation steps (S/transform/vision/image).", [class_section(n, c) for n, c in own_classes(VI)]))
    from bigdl_b200.tensor import tensor as TT
    written.append(write("APIGuide/Tensor.md", "Tensor", "The reference's 1-based `Tensor` API over torch storage; fp32 / fp64 BLAS and vector math go through the native entry points (`tensor/mkl.py`).",
                         [class_section("Tensor", TT.Tensor, methods=True)]))
    from bigdl_b200.utils import engine as E
    written.append(write("APIGuide/Engine.md", "Engine", "Process / device topology (`Engine.init`, node and core numbers, `checkSingleton`).", [class_section("Engine", E._Engine, True)]))
    # ---------------------------------------------------------------- keras-style
    from bigdl_b200.nn import keras as K
    kl = own_classes(K.layers, K.KerasLayer) if hasattr(K, "layers") else own_classes(K.layer, K.KerasLayer)
    buckets = {"core": [], "convolutional": [], "pooling": [], "recurrent": [], "normalization": [], "embedding": [],
               "advanced-activation": [], "dropout": [], "wrappers": []}
    for n, c in kl:
        ln = n.lower()
        key = ("convolutional" if any(k in ln for k in ("conv", "cropping", "upsampling", "zeropadding", "locallyconnected")) else
               "pooling" if "pooling" in ln else
               "recurrent" if any(k in ln for k in ("lstm", "gru", "rnn")) else
               "normalization" if "normalization" in ln else
               "embedding" if "embedding" in ln else
               "advanced-activation" if any(k in ln for k in ("elu", "relu", "threshold", "softplus")) and ln != "activation" else
               "dropout" if any(k in ln for k in ("dropout", "gaussian")) else
               "wrappers" if any(k in ln for k in ("timedistributed", "bidirectional")) else "core")
        buckets[key].append(class_section(n, c))
    for k, secs in buckets.items():
        written.append(write(f"KerasStyleAPIGuide/Layers/{k}.md", f"Keras-style layers: {k}", "Keras-1.2.2-style layers with shape inference (`bigdl_b200.nn.keras`).", secs))
    from bigdl_b200.nn.keras import layers as KL
    acts = ["tanh", "sigmoid", "hard_sigmoid", "relu", "softplus", "softsign", "softmax", "linear"]
    written.append(write("KerasStyleAPIGuide/Layers/activation.md", "Keras-style activations",
                         "`Activation(name)` as a layer, or `activation=name` on `Dense`, the convolutions and the "
                         "recurrent layers. Names: " + ", ".join(f"`{a}`" for a in acts) + ".",
                         [class_section("Activation", KL.Activation)] + [
                             f"## {a}\n\n→ `{type(KL.activation_module(a)).__name__}`\n" for a in acts]))
    inits = ["glorot_uniform", "one", "zero", "uniform", "normal"]
    written.append(write("KerasStyleAPIGuide/Layers/initialization.md", "Keras-style initialisers",
                         "`init=name` on the parameterised layers.",
                         [f"## {i}\n\n→ `{type(KL._init_method(i)).__name__}`"
                          + (f"({', '.join(f'{k}={v}' for k, v in vars(KL._init_method(i)).items())})\n" if vars(KL._init_method(i)) else "\n")
                          for i in inits]))
    from bigdl_b200.nn.keras import topology as KT
    from bigdl_b200.interop.keras_helpers import OptimConverter as _OC
    losses = ["categorical_crossentropy", "mse", "mean_squared_error", "binary_crossentropy", "mae", "mean_absolute_error",
              "hinge", "mape", "mean_absolute_percentage_error", "msle", "mean_squared_logarithmic_error",
              "squared_hinge", "sparse_categorical_crossentropy", "kld", "kullback_leibler_divergence", "cosine_proximity",
              "poisson"]
    rows = []
    for name in losses:
        try:
            rows.append(f"| `{name}` | `{type(_OC.to_bigdl_criterion(name)).__name__}` |")
        except Exception:
            pass
    written.append(write("KerasStyleAPIGuide/Optimization/loss.md", "Losses for `compile`",
                         "`model.compile(optimizer, loss, metrics)` accepts a criterion object or one of these names.",
                         ["| name | criterion |", "|---|---|"] + rows))
    orow = []
    for name in ["sgd", "adam", "adamax", "adagrad", "adadelta", "rmsprop"]:
        try:
            orow.append(f"| `{name}` | `{type(_OC.to_bigdl_optim_method(name)).__name__}` |")
        except Exception:
            pass
    written.append(write("KerasStyleAPIGuide/Optimization/optimizer.md", "Optimizers for `compile`",
                         "An `OptimMethod` object or one of these names (default hyper-parameters of Keras 1.2.2).",
                         ["| name | optim method |", "|---|---|"] + orow))
    written.append(write("KerasStyleAPIGuide/Optimization/training.md", "compile / fit / evaluate / predict", "Keras-style `Sequential` and `Model` (S/nn/keras/Topology.scala).", [class_section(n, c, True) for n, c in own_classes(KT)]))
    # ---------------------------------------------------------------- configuration keys, from the registry source
    import re
    rows, pending = [], []
    for line in open(os.path.join(ROOT, "bigdl_b200", "utils", "config.py")):
        m = re.match(r'\s*"(bigdl[^"]+)":\s*(.+?),\s*(?:#\s*(.*))?$', line)
        if m:
            note = " ".join(pending + ([m.group(3)] if m.group(3) else []))
            rows.append(f"| `{m.group(1)}` | `{m.group(2)}` | {note.replace(chr(124), chr(47))} |")
            pending = []
        else:
            c = re.match(r"\s*#\s*(.*)$", line)
            pending = pending + [c.group(1)] if c and rows is not None and "_DEFAULTS" not in line else []
    written.append(write("ProgrammingGuide/configuration.md", "Configuration",
                         "Every flag lives in `bigdl_b200/utils/config.py`. Priority: `config.set(key, value)` > environment "
                         "variable `BIGDL_<KEY>` (dots -> underscores, upper case; e.g. `BIGDL_PARAMETER_BACKEND=nccl`) > "
                         "`conf/bigdl-b200.conf` > the default. The reference's JVM property names are accepted verbatim "
                         "(S/utils/Engine.scala, SURVEY §5.6). `bench.py --set key=value` overrides a key for one run.",
                         ["| key | default | meaning |", "|---|---|---|"] + rows))
    print(len(written), "pages")
    for p in written:
        print(" ", os.path.relpath(p, ROOT), os.path.getsize(p))


if __name__ == "__main__":
    main()
