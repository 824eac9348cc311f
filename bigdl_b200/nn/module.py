"""AbstractModule / Container / Sequential / Graph — the Torch7-style module API on a PyTorch substrate.

Reference: S/nn/abstractnn/AbstractModule.scala (forward:255, backward:282, getParameters:988 →
S/nn/Module.scala:76-107 flatten, freeze:204, setScaleW:95, getTimes:168, saveModule:548, predict:637,
evaluate:856, quantize:920), S/nn/Container.scala, S/nn/Sequential.scala:35, S/nn/Graph.scala / StaticGraph.scala:56-152,
S/nn/Input.scala, S/utils/DirectedGraph.scala.

Design (not a translation): a layer only defines ``updateOutput`` with differentiable tensor ops (torch ops on CPU,
hand-written sm_100a kernels wrapped in ``torch.autograd.Function`` on GPU). ``backward`` is derived by autograd,
so ``updateGradInput``/``accGradParameters`` exist for API parity but are computed from the same graph.
Parameters live as views into ONE flat fp32 buffer (+ one flat fp32 gradient buffer, + on GPU one flat bf16
compute copy) after ``getParameters()`` — the layout the sharded optimizer update and the NVLink reduce-scatter
operate on directly.
"""
from __future__ import annotations

import copy
import itertools
import time
from collections import OrderedDict
from typing import Callable, Iterable

import numpy as np
import torch

from ..utils.table import Table, flatten_activity, map_activity, to_table

_uid = itertools.count(1)


def _as_activity(x, device=None):
    """numpy / python scalars / lists → Tensor | Table."""
    if isinstance(x, torch.Tensor):
        return x if device is None or x.device == device else x.to(device)
    if isinstance(x, np.ndarray):
        if x.dtype == object or x.dtype.kind in "US":
            return x                       # string / object columns (feature-column ops) stay numpy
        t = torch.from_numpy(np.ascontiguousarray(x))
        if t.dtype == torch.float64:
            t = t.float()
        return t if device is None else t.to(device)
    if isinstance(x, Table):
        out = Table()
        for k, v in x.items():               # keeps non-positional keys ("bboxes", "masks", …)
            out[k] = _as_activity(v, device)
        return out
    if isinstance(x, (list, tuple)):
        return Table(*[_as_activity(v, device) for v in x])
    if isinstance(x, (int, float)):
        t = torch.tensor(float(x))
        return t if device is None else t.to(device)
    return x        # None, strings, sparse tensors and other opaque activities pass through


def _detach(x):
    return map_activity(lambda t: t.detach(), x)


def _plain_tensor(x) -> torch.Tensor:
    """torch.Tensor from a torch tensor, a ``tensor.Tensor`` wrapper (its ``.t``), an ndarray or a nested list."""
    if isinstance(x, torch.Tensor):
        return x
    inner = getattr(x, "t", None)
    if isinstance(inner, torch.Tensor):
        return inner
    return torch.as_tensor(x)


def _capture_ctor(cls):
    """Wrap ``cls.__init__`` so every instance remembers its constructor arguments (``_ctor_args``) and whether the
    constructor itself created sub-modules (``_children_from_ctor``) — the Python stand-in for the constructor
    reflection the reference's serializer relies on (S/utils/ReflectionUtils.scala, ModuleSerializable.scala:150-200)."""
    import functools
    import inspect

    init = cls.__dict__.get("__init__")
    if init is None or getattr(init, "_captures", False):
        return
    try:
        sig = inspect.signature(init)
    except (TypeError, ValueError):
        return

    @functools.wraps(init)
    def wrapped(self, *a, **k):
        outermost = "_ctor_args" not in self.__dict__
        if outermost:
            try:
                bound = sig.bind(self, *a, **k)
                bound.apply_defaults()
                self.__dict__["_ctor_args"] = dict(list(bound.arguments.items())[1:])
                self.__dict__["_ctor_cls"] = cls
            except TypeError:
                self.__dict__["_ctor_args"] = {}
                self.__dict__["_ctor_cls"] = cls
        init(self, *a, **k)
        if outermost:
            self.__dict__["_children_from_ctor"] = len(self.__dict__.get("modules", [])) > 0

    wrapped._captures = True
    cls.__init__ = wrapped


_NVTX = [False]       # set by utils.profiling.enable_nvtx

class Module:
    """Base class of every layer (reference: AbstractModule[A, B, T])."""

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        _capture_ctor(cls)

    def __init__(self):
        self._uid = next(_uid)
        self._name: str | None = None
        self.output = None
        self.gradInput = None
        self.train = True
        self._parameters: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        self._buffers: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        self._compute: dict[str, torch.Tensor] = {}   # bf16 views of the parameters (GPU compute copy)
        self.modules: list[Module] = []
        self.scaleW = 1.0
        self.scaleB = 1.0
        self.forwardTime = 0
        self.backwardTime = 0
        self.wRegularizer = None
        self.bRegularizer = None
        self._init_methods = {}
        self._graph_in = None
        self._graph_out = None
        self._frozen = False

    # ------------------------------------------------------------------ naming
    def setName(self, name: str):
        self._name = name
        return self

    set_name = setName

    def getName(self) -> str:
        return self._name if self._name is not None else f"{type(self).__name__}{self._uid}"

    get_name = getName
    name = getName

    def hasName(self) -> bool:
        return self._name is not None

    def __repr__(self):
        return f"{type(self).__name__}[{self.getName()}]"

    # ------------------------------------------------------------------ parameters
    def register_parameter(self, name: str, value: torch.Tensor):
        p = value.detach().clone().float() if not value.is_floating_point() else value.detach().clone()
        p.requires_grad_(True)
        p.grad = torch.zeros_like(p)
        self._parameters[name] = p
        return p

    def register_buffer(self, name: str, value: torch.Tensor):
        self._buffers[name] = value
        return value

    def __getattr__(self, item):
        d = self.__dict__
        if "_parameters" in d and item in d["_parameters"]:
            return d["_parameters"][item]
        if "_buffers" in d and item in d["_buffers"]:
            return d["_buffers"][item]
        if item.startswith("grad") and "_parameters" in d:
            base = item[4:5].lower() + item[5:]
            if base in d["_parameters"]:
                return d["_parameters"][base].grad
        raise AttributeError(f"{type(self).__name__} has no attribute {item!r}")

    def own_parameters(self) -> list[torch.Tensor]:
        return list(self._parameters.values())

    def all_modules(self) -> Iterable["Module"]:
        """Depth-first, parents before children, each module once."""
        seen = set()

        def rec(m):
            if id(m) in seen:
                return
            seen.add(id(m))
            yield m
            for c in m.modules:
                yield from rec(c)

        yield from rec(self)

    def leaf_parameters(self) -> list[tuple["Module", str, torch.Tensor]]:
        out, seen = [], set()
        for m in self.all_modules():
            for n, p in m._parameters.items():
                if id(p) not in seen:
                    seen.add(id(p))
                    out.append((m, n, p))
        return out

    def parameters(self):
        """(weights, gradWeights) lists in traversal order (reference: AbstractModule.parameters())."""
        ps = [p for _, _, p in self.leaf_parameters()]
        return [p.data for p in ps], [p.grad for p in ps]

    def torch_parameters(self) -> list[torch.Tensor]:
        return [p for _, _, p in self.leaf_parameters()]

    def getParametersTable(self) -> Table:
        t = Table()
        for m in self.all_modules():
            if m._parameters:
                sub = Table()
                for n, p in m._parameters.items():
                    sub[n] = p.data
                    sub["grad" + n[0].upper() + n[1:]] = p.grad
                t[m.getName()] = sub
        return t

    get_parameters_table = getParametersTable

    def getParameters(self):
        """Flatten every parameter into ONE contiguous fp32 vector (and one gradient vector) and re-point the
        layers at views of it. Returns (flat_weights, flat_grads). Idempotent. (S/nn/Module.scala:76-107)"""
        from .flat import flatten_parameters

        return flatten_parameters(self)

    get_parameters = getParameters

    def getWeights(self):
        return [p.detach().float().cpu().contiguous().numpy() for p in self.torch_parameters()]

    get_weights = getWeights

    def setWeights(self, weights):
        ps = self.torch_parameters()
        if len(weights) != len(ps):
            raise ValueError(f"expected {len(ps)} weight arrays, got {len(weights)}")
        with torch.no_grad():
            for p, w in zip(ps, weights):
                w = torch.as_tensor(np.asarray(w), dtype=p.dtype).reshape(p.shape)
                p.copy_(w.to(p.device))
        self._mark_weights_dirty()
        return self

    set_weights = setWeights

    def _mark_weights_dirty(self):
        root = self
        fl = getattr(root, "_flat", None)
        if fl is not None:
            fl.dirty = True
        for m in self.all_modules():
            f = getattr(m, "_flat", None)
            if f is not None:
                f.dirty = True

    def release(self):
        """Drop what this module holds beyond its parameters — the last activations and their gradients — so the
        device memory returns to the caching allocator (the reference frees the layers' native MKL-DNN resources
        here, AbstractModule.release)."""
        for m in self.all_modules():
            m.output = None
            m.gradInput = None
        return self

    def zeroGradParameters(self):
        fl = getattr(self, "_flat", None)
        if fl is not None:
            fl.grads.zero_()
            return self
        for p in self.torch_parameters():
            if p.grad is not None:
                p.grad.zero_()
        return self

    zero_grad_parameters = zeroGradParameters

    def updateParameters(self, learningRate: float):
        with torch.no_grad():
            for p in self.torch_parameters():
                if p.grad is not None and p.requires_grad:
                    p.add_(p.grad, alpha=-learningRate)
        self._mark_weights_dirty()
        return self

    update_parameters = updateParameters

    def reset(self):
        for m in self.all_modules():
            if m is not self:
                m.reset()
        return self

    def setInitMethod(self, weightInitMethod=None, biasInitMethod=None):
        if weightInitMethod is not None:
            self._init_methods["weight"] = weightInitMethod
        if biasInitMethod is not None:
            self._init_methods["bias"] = biasInitMethod
        self.reset()
        return self

    set_init_method = setInitMethod

    # ------------------------------------------------------------------ gradient scaling / freezing
    def setScaleW(self, w: float):
        self.scaleW = w
        for m in self.modules:
            m.setScaleW(w)
        return self

    def setScaleB(self, b: float):
        self.scaleB = b
        for m in self.modules:
            m.setScaleB(b)
        return self

    def getScaleW(self):
        return self.scaleW

    def getScaleB(self):
        return self.scaleB

    def freeze(self, *names):
        """Stop updating this module (or the named sub-modules). (AbstractModule.scala:204)"""
        if not names:
            for m in self.all_modules():
                m._frozen = True
                for p in m._parameters.values():
                    p.requires_grad_(False)
        else:
            for n in names:
                m = self.find(n)
                if m is None:
                    raise ValueError(f"cannot find module named {n}")
                m.freeze()
        return self

    def unFreeze(self, *names):
        if not names:
            for m in self.all_modules():
                m._frozen = False
                for p in m._parameters.values():
                    p.requires_grad_(True)
        else:
            for n in names:
                m = self.find(n)
                if m is None:
                    raise ValueError(f"cannot find module named {n}")
                m.unFreeze()
        return self

    unfreeze = unFreeze

    def find(self, name: str):
        for m in self.all_modules():
            if m.getName() == name:
                return m
        return None

    @staticmethod
    def of(jvalue, bigdl_type="float"):
        """Python API ``Layer.of(jvalue)``: there is no JVM peer, the module is its own value."""
        return jvalue

    def apply(self, name: str):
        """The (sub-)module called ``name``, or None — ``model("conv1")`` in the reference (AbstractModule.scala:840,
        Container.scala:159)."""
        return self.find(name)

    def findModules(self, moduleType: str):
        """Every (sub-)module whose class is called ``moduleType`` (Container.scala:123-139)."""
        return [m for m in self.all_modules() if type(m).__name__ == moduleType]

    # ---- weights by position / by layer name (AbstractModule.scala:692-771)
    def getWeightsBias(self):
        ps = self.parameters()
        return None if ps is None or not ps[0] else list(ps[0])

    def setWeightsBias(self, newWeights):
        ws = self.getWeightsBias()
        if ws is None:
            raise ValueError("this layer does not have weight/bias")
        if len(ws) != len(newWeights):
            raise ValueError("the number of input weight/bias is not consistant with number of weight/bias of this "
                             f"layer, number of input {len(ws)}, number of output {len(newWeights)}")
        with torch.no_grad():
            for w, n in zip(ws, newWeights):
                w.copy_(_plain_tensor(n).reshape(w.shape).to(w.dtype))
        self._mark_weights_dirty()
        return self

    def getExtraParameter(self):
        """State that is neither weight nor gradient — batch-norm running statistics — as a list of tensors."""
        out = []
        for m in self.all_modules():
            for n in ("runningMean", "runningVar"):
                b = m._buffers.get(n)
                if isinstance(b, torch.Tensor):
                    out.append(b)
        return out or None

    def setExtraParameter(self, extraParam):
        cur = self.getExtraParameter()
        if extraParam is None and cur is None:
            return self
        if extraParam is None or cur is None:
            raise ValueError("module's extraParameter is null or the given extraParameter is null")
        if len(cur) != len(extraParam):
            raise ValueError(f"state's length doesn't match, excepted:{len(cur)}, but got  {len(extraParam)}")
        with torch.no_grad():
            for c, e in zip(cur, extraParam):
                c.copy_(_plain_tensor(e).reshape(c.shape).to(c.dtype))
        return self

    @staticmethod
    def _copy_weights(target, src, matchAll: bool):
        with torch.no_grad():
            for name, tp in target.items():
                if name in src:
                    for k in ("weight", "bias"):
                        if k in src[name] and src[name][k] is not None and k in tp and tp[k] is not None:
                            tp[k].copy_(torch.as_tensor(src[name][k]).reshape(tp[k].shape).to(tp[k].dtype))
                elif matchAll:
                    raise KeyError(f"module {name} cannot find corresponding weight bias")

    def loadModelWeights(self, srcModel, matchAll: bool = True):
        """Copy ``weight`` / ``bias`` from the layers of ``srcModel`` that carry the same names."""
        self._copy_weights(self.getParametersTable(), srcModel.getParametersTable(), matchAll)
        self._mark_weights_dirty()
        return self

    def saveWeights(self, path: str, overWrite: bool = False):
        """{layer name: {weight, bias}} as one file (``loadWeights`` reads it back)."""
        from ..utils.file import File

        table = {name: {k: v.detach().cpu().clone() for k, v in params.items() if k in ("weight", "bias") and v is not None}
                 for name, params in self.getParametersTable().items()}
        File.save(table, path, overWrite)

    def loadWeights(self, weightPath: str, matchAll: bool = True):
        from ..utils.file import File

        self._copy_weights(self.getParametersTable(), File.load(weightPath), matchAll)
        self._mark_weights_dirty()
        return self

    def saveDefinition(self, path: str, overWrite: bool = False):
        """The weight-free text form of the model structure."""
        from ..serialization.serializer import ModulePersister

        self.clearState()
        ModulePersister.saveModelDefinitionToFile(path, self, overWrite)
        return self

    def getNumericType(self):
        from ..tensor.types import TensorDataType

        for p in self.torch_parameters():
            return TensorDataType.of(p.dtype)
        return TensorDataType.of(torch.float32)

    def setLine(self, line: str):
        """Separator printed between the children in ``str(container)``."""
        self.line = line
        return self

    def getHashCode(self) -> int:
        return hash((type(self).__name__, self.getName()))

    def evaluateImage(self, imageFrame, vMethods, batchSize: int | None = None):
        """Validate on the ``sample`` stored in every valid ImageFeature of the frame (AbstractModule.scala:886-900)."""
        from ..dataset.base import DataSet

        samples = []
        for f in imageFrame.features:
            if f.isValid:
                if "sample" not in f:
                    raise ValueError("ImageFeature must have sample")
                samples.append(f["sample"])
        return self.evaluate(DataSet.array(samples, getattr(imageFrame, "isDistributed", lambda: False)()), vMethods,
                             batchSize)

    def getEndNodes(self, startNodes):
        """End nodes of this module hooked after ``startNodes`` (used when a Sequential is unrolled into a graph)."""
        return [self.inputs(*startNodes)]

    def checkEngineType(self):
        """The reference refuses to mix mkl-dnn and blas layers in one container; every layer here runs on either
        engine, so there is nothing to reject."""
        return self

    apply_name = find

    # ------------------------------------------------------------------ mode
    def training(self):
        for m in self.all_modules():
            m.train = True
        return self

    def evaluate(self, *args, **kw):
        """No arguments: switch to inference mode. With (dataset, batch_size, vMethods): run validation."""
        if not args and not kw:
            for m in self.all_modules():
                m.train = False
            return self
        from ..optim.evaluator import Evaluator

        return Evaluator(self).test(*args, **kw)

    def isTraining(self) -> bool:
        return self.train

    is_training = isTraining

    # ------------------------------------------------------------------ device
    def device(self) -> torch.device:
        for p in self.torch_parameters():
            return p.device
        for m in self.all_modules():
            for b in m._buffers.values():
                return b.device
        return torch.device("cpu")

    def to(self, device):
        device = torch.device(device)
        from .flat import unflatten_parameters

        unflatten_parameters(self)
        for m in self.all_modules():
            for n, p in list(m._parameters.items()):
                if p.device != device:
                    q = p.detach().to(device)
                    q.requires_grad_(p.requires_grad)
                    q.grad = p.grad.to(device) if p.grad is not None else torch.zeros_like(q)
                    m._parameters[n] = q
            for n, b in list(m._buffers.items()):
                if isinstance(b, torch.Tensor) and b.device != device:
                    m._buffers[n] = b.to(device)
            m._compute = {}
            m._on_device_change(device)
        if device.type == "cuda":
            self.getParameters()
            from .fusion import fuse

            fuse(self)
        return self

    def cuda(self, index: int | None = None):
        return self.to(torch.device("cuda", torch.cuda.current_device() if index is None else index))

    def cpu(self):
        return self.to("cpu")

    def _on_device_change(self, device):
        pass

    # ------------------------------------------------------------------ compute
    def updateOutput(self, input):  # pragma: no cover - abstract
        raise NotImplementedError(type(self).__name__)

    def _call(self, input):
        """Graph-building call used by containers and training loops (no detach, no host sync)."""
        pre = self.__dict__.get("_pre_call_hook")      # ParallelOptimizer bucket boundary / lazy parameter update
        if pre is not None:
            input = pre(self, input)
        if _NVTX[0]:                                   # utils/profiling.enable_nvtx: one range per layer call
            torch.cuda.nvtx.range_push(f"{type(self).__name__}:{self.getName()}")
            try:
                out = self.updateOutput(input)
            finally:
                torch.cuda.nvtx.range_pop()
        else:
            out = self.updateOutput(input)
        self.output = out
        return out

    def weight_for_compute(self, name: str) -> torch.Tensor:
        """bf16 compute copy of a parameter when one exists (GPU), else the fp32 parameter itself."""
        c = self._compute.get(name)
        return c if c is not None else self._parameters[name]

    def _refresh_compute_weights(self):
        fl = getattr(self, "_flat", None)
        if fl is not None:
            fl.refresh_compute()

    def forward(self, input):
        """Run the layer; returns the output detached from the autograd graph (numpy in → numpy out)."""
        t0 = time.perf_counter_ns()
        def _num(v):
            return isinstance(v, np.ndarray) and v.dtype != object and v.dtype.kind not in "US"

        want_numpy = _num(input) or (
            isinstance(input, (list, tuple)) and len(input) > 0 and all(_num(v) for v in input))
        x = _as_activity(input, self.device() if self.torch_parameters() else None)

        def leaf(t):
            t = t.detach()
            if t.is_floating_point():
                t.requires_grad_(True)
            return t

        self._graph_in = map_activity(leaf, x)
        self._refresh_compute_weights()
        with torch.enable_grad():
            self._graph_out = self._call(self._graph_in)
        self.output = _detach(self._graph_out)
        self.forwardTime += time.perf_counter_ns() - t0
        if want_numpy:
            return _to_numpy(self.output)
        return self.output

    def backward(self, input, gradOutput):
        """gradInput = d(loss)/d(input) given gradOutput; parameter gradients are accumulated (scaled by
        scaleW/scaleB) into the gradient buffers. Must follow a ``forward`` on the same input."""
        t0 = time.perf_counter_ns()
        want_numpy = isinstance(gradOutput, np.ndarray)
        if self._graph_out is None:
            self.forward(input)
        outs = flatten_activity(self._graph_out)
        gos = flatten_activity(_as_activity(gradOutput, outs[0].device if outs else None))
        ins = [t for t in flatten_activity(self._graph_in) if t.requires_grad]
        pairs = [(o, g.to(o.dtype)) for o, g in zip(outs, gos) if o.requires_grad]
        if pairs:
            self._backward_scaled(pairs, ins)
        if torch.cuda.is_available():
            from ..ops.functions import join_side_streams

            join_side_streams()
        self.gradInput = map_activity(
            lambda t: (t.grad if t.grad is not None else torch.zeros_like(t)) if t.is_floating_point()
            else torch.zeros_like(t, dtype=torch.float32), self._graph_in)
        self._graph_out = None
        self._apply_regularizers()
        self.backwardTime += time.perf_counter_ns() - t0
        if want_numpy:
            return _to_numpy(self.gradInput)
        return self.gradInput

    def _backward_scaled(self, pairs, ins):
        scaled = [(m, n, p) for m, n, p in self.leaf_parameters()
                  if p.requires_grad and ((m.scaleW != 1.0 and n != "bias") or (m.scaleB != 1.0 and n == "bias"))]
        if not scaled:
            torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
            return
        # gradient scaling: divert the scaled parameters' gradients, scale, then accumulate
        saved = [(p, p.grad) for _, _, p in scaled]
        for p, _ in saved:
            p.grad = None
        torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
        for (m, n, p), (_, old) in zip(scaled, saved):
            s = m.scaleB if n == "bias" else m.scaleW
            new = p.grad
            p.grad = old
            if new is not None and old is not None:
                old.add_(new, alpha=s)

    def updateGradInput(self, input, gradOutput):
        if self._graph_out is None:
            self.forward(input)
        outs = flatten_activity(self._graph_out)
        gos = flatten_activity(_as_activity(gradOutput, outs[0].device))
        ins = [t for t in flatten_activity(self._graph_in) if t.requires_grad]
        grads = torch.autograd.grad(outs, ins, [g.to(o.dtype) for o, g in zip(outs, gos)], retain_graph=True,
                                    allow_unused=True)
        it = iter(grads)
        self.gradInput = map_activity(
            lambda t: (lambda g: g if g is not None else torch.zeros_like(t))(next(it)) if t.requires_grad
            else torch.zeros_like(t, dtype=torch.float32), self._graph_in)
        return self.gradInput

    def accGradParameters(self, input, gradOutput):
        if self._graph_out is None:
            self.forward(input)
        ps = [(m, n, p) for m, n, p in self.leaf_parameters() if p.requires_grad]
        if not ps:
            return
        outs = flatten_activity(self._graph_out)
        gos = flatten_activity(_as_activity(gradOutput, outs[0].device))
        grads = torch.autograd.grad(outs, [p for _, _, p in ps], [g.to(o.dtype) for o, g in zip(outs, gos)],
                                    retain_graph=True, allow_unused=True)
        with torch.no_grad():
            for (m, n, p), g in zip(ps, grads):
                if g is not None:
                    p.grad.add_(g, alpha=m.scaleB if n == "bias" else m.scaleW)
        self._apply_regularizers()

    def _apply_regularizers(self):
        for m in self.all_modules():
            if m.wRegularizer is not None and "weight" in m._parameters and m._parameters["weight"].requires_grad:
                m.wRegularizer.accRegularization(m._parameters["weight"], m._parameters["weight"].grad, m.scaleW)
            if m.bRegularizer is not None and "bias" in m._parameters and m._parameters["bias"].requires_grad:
                m.bRegularizer.accRegularization(m._parameters["bias"], m._parameters["bias"].grad, m.scaleB)

    # ------------------------------------------------------------------ timing (AbstractModule.scala:168-190)
    def getTimes(self):
        return [(m, m.forwardTime, m.backwardTime) for m in self.all_modules()]

    get_times = getTimes

    def getTimesGroupByModuleType(self):
        agg: dict[str, list[int]] = {}
        for m, f, b in self.getTimes():
            a = agg.setdefault(type(m).__name__, [0, 0])
            a[0] += f
            a[1] += b
        return [(k, v[0], v[1]) for k, v in agg.items()]

    def resetTimes(self):
        for m in self.all_modules():
            m.forwardTime = 0
            m.backwardTime = 0
        return self

    reset_times = resetTimes

    # ------------------------------------------------------------------ cloning / state
    def clearState(self):
        for m in self.all_modules():
            m.output = None
            m.gradInput = None
            m._graph_in = None
            m._graph_out = None
        return self

    clear_state = clearState

    def cloneModule(self):
        """Deep copy with its own parameter storage. The source keeps its flat parameter buffer: optimizers, captured
        CUDA graphs and the parameter synchroniser hold pointers into it, so it must never be rebuilt behind their
        back (``__deepcopy__`` clones every parameter out of the buffer individually)."""
        self.clearState()
        was_flat = getattr(self, "_flat", None) is not None
        dev = self.device()
        c = copy.deepcopy(self)
        for m in c.all_modules():
            m._uid = next(_uid)
        if was_flat and dev.type == "cuda":
            c.getParameters()
        return c

    clone = cloneModule

    def __deepcopy__(self, memo):
        cls = self.__class__
        new = object.__new__(cls)      # Input.__new__ returns a Node, so bypass class-level __new__
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            if k in ("_flat", "_compute"):
                new.__dict__[k] = None if k == "_flat" else {}
            elif k == "_parameters":
                od = OrderedDict()
                for n, p in v.items():
                    if id(p) in memo:
                        od[n] = memo[id(p)]
                        continue
                    q = p.detach().clone().contiguous() if not p.is_contiguous(memory_format=torch.channels_last) or p.dim() != 4 \
                        else p.detach().clone(memory_format=torch.channels_last)
                    q.requires_grad_(p.requires_grad)
                    q.grad = torch.zeros_like(q) if p.grad is None else p.grad.detach().clone()
                    memo[id(p)] = q
                    od[n] = q
                new.__dict__[k] = od
            elif k in ("_train_summary", "_val_summary", "_pre_call_hook", "_gather_plan"):
                continue              # run-time attachments (event writers own threads and locks): not part of a copy
            else:
                new.__dict__[k] = copy.deepcopy(v, memo)
        return new

    # ------------------------------------------------------------------ graph-node sugar: layer(node, ...) / .inputs()
    def inputs(self, *nodes):
        flat = []
        for n in nodes:
            if isinstance(n, (list, tuple)):
                flat.extend(n)
            else:
                flat.append(n)
        return Node(self, flat)

    def __call__(self, *nodes):
        if len(nodes) == 1 and isinstance(nodes[0], (list, tuple)) and nodes[0] and isinstance(nodes[0][0], Node):
            nodes = tuple(nodes[0])
        if nodes and all(isinstance(n, Node) for n in nodes):
            return self.inputs(*nodes)
        if len(nodes) == 1:
            return self.forward(nodes[0])   # tensor / ndarray / table: behave like torch's module(x)
        if not nodes:
            return self.inputs()
        return self.forward(list(nodes))

    # ------------------------------------------------------------------ persistence / inference helpers
    def saveModule(self, path: str, weightPath: str | None = None, overWrite: bool = False):
        from ..serialization.module_io import save_module

        save_module(self, path, weightPath, overWrite)
        return self

    saveModel = saveModule
    save_model = saveModule

    def save(self, path: str, overWrite: bool = False):
        from ..serialization.snapshot import save_snapshot

        save_snapshot(self, path, overWrite)
        return self

    @staticmethod
    def load(path: str):
        from ..serialization.snapshot import load_snapshot

        return load_snapshot(path)

    @staticmethod
    def loadModule(path: str, weightPath: str | None = None):
        from ..serialization.module_io import load_module

        return load_module(path, weightPath)

    load_model = loadModule

    @staticmethod
    def loadTorch(path: str):
        from ..interop.torchfile import load_torch_module

        return load_torch_module(path)

    load_torch = loadTorch

    @staticmethod
    def loadCaffe(model, defPath: str, modelPath: str, matchAll: bool = True, match_all: bool | None = None):
        """Copy weights from a caffemodel into an existing model by layer name (CaffeLoader.load); ``match_all`` is the
        Python API's spelling of ``matchAll``."""
        from ..interop.caffe import CaffeLoader

        if match_all is not None:
            matchAll = match_all

        return CaffeLoader.load(model, defPath, modelPath, matchAll)

    @staticmethod
    def loadCaffeModel(defPath: str, modelPath: str):
        from ..interop.caffe import CaffeLoader

        return CaffeLoader.loadCaffe(defPath, modelPath)[0]

    load_caffe, load_caffe_model = loadCaffe, loadCaffeModel

    @staticmethod
    def loadTF(path: str, inputs, outputs, byteOrder: str = "little_endian", binFile: str | None = None):
        from ..interop.tensorflow import load_tf

        return load_tf(path, inputs, outputs, byteOrder, binFile)

    load_tensorflow = loadTF

    @staticmethod
    def load_keras(json_path=None, hdf5_path=None, by_name=False):
        from ..interop.keras_converter import load_keras

        return load_keras(json_path, hdf5_path, by_name)

    def saveTorch(self, path: str, overWrite: bool = False):
        from ..interop.torchfile import save_torch_module

        save_torch_module(self, path, overWrite)
        return self

    def saveCaffe(self, prototxtPath: str, modelPath: str, useV2: bool = True, overwrite: bool = False):
        from ..interop.caffe import CaffePersister

        CaffePersister.persist(prototxtPath, modelPath, self, useV2, overwrite)
        return self

    save_caffe = saveCaffe

    def saveTF(self, inputs, path: str, byteOrder="little_endian", dataFormat="nhwc"):
        from ..interop.tensorflow import save_tf

        save_tf(self, inputs, path, byteOrder, dataFormat)
        return self

    save_tensorflow = saveTF

    def predict(self, data, batch_size: int = -1, **kw):
        from ..optim.predictor import Predictor

        return Predictor(self, batch_size=batch_size, **kw).predict(data)

    def predictClass(self, data, batch_size: int = -1):
        from ..optim.predictor import Predictor

        return Predictor(self, batch_size=batch_size).predictClass(data)

    predict_class = predictClass

    def predictImage(self, imageFrame, outputLayer=None, shareBuffer=False, batchPerPartition=4, predictKey="predict"):
        from ..optim.predictor import Predictor

        return Predictor(self, batch_size=batchPerPartition).predictImage(imageFrame, outputLayer, predictKey)

    predict_image = predictImage

    def quantize(self):
        from ..quantized.quantizer import Quantizer

        return Quantizer.quantize(self)

    def toGraph(self, *start_nodes):
        from .graph_convert import to_graph

        return to_graph(self, *start_nodes)

    to_graph = toGraph

    def checkDuplicate(self):
        """No layer instance may appear twice in a container tree (S/optim/Optimizer.scala:80)."""
        seen = set()

        def rec(m):
            if id(m) in seen:
                raise ValueError(f"Some module is duplicate in the model: {m}")
            seen.add(id(m))
            for c in m.modules:
                rec(c)

        rec(self)
        return self

    # ------------------------------------------------------------------ parameter synchroniser hooks (ParallelOptimizer)
    def setParameterSynchronizer(self, sync, priority: int = 0):
        self._param_sync = (sync, priority)
        return self


def _to_numpy(x):
    if isinstance(x, torch.Tensor):
        return x.detach().float().cpu().numpy() if x.dtype == torch.bfloat16 else x.detach().cpu().numpy()
    if isinstance(x, Table):
        return [_to_numpy(v) for v in x.to_list()]
    return x


TensorModule = Module
AbstractModule = Module


# ====================================================================== containers
class Container(Module):
    """Module holding an ordered list of sub-modules (S/nn/Container.scala)."""

    def __init__(self, *modules):
        super().__init__()
        for m in modules:
            self.add(m)

    def add(self, module: Module):
        if not isinstance(module, Module):
            raise TypeError(f"Container.add expects a Module, got {type(module)}")
        self.modules.append(module)
        return self

    def get(self, index: int) -> Module:
        """1-based, like the reference's ``apply(index)``."""
        return self.modules[index - 1]

    def __getitem__(self, index: int) -> Module:
        return self.modules[index]

    def __len__(self):
        return len(self.modules)

    def size(self):
        return len(self.modules)

    @property
    def layers(self):
        return self.modules

    def flattened_layers(self, include_container: bool = False):
        out = []
        for m in self.all_modules():
            if m is self:
                continue
            if m.modules and not include_container:
                continue
            out.append(m)
        return out

    def __repr__(self):
        inner = "\n".join("  " + l for m in self.modules for l in repr(m).split("\n"))
        return f"{type(self).__name__}[{self.getName()}]{{\n{inner}\n}}"


class Sequential(Container):
    """Feed-forward chain (S/nn/Sequential.scala:35-45)."""

    def updateOutput(self, input):
        x = input
        for m in self.modules:
            x = m._call(x)
        return x

    def getEndNodes(self, startNodes):
        cur = list(startNodes)
        for m in self.modules:
            cur = m.getEndNodes(cur)
        return cur


class BaseModule(Module):
    """A layer defined by the model its ``buildModel()`` returns (S/nn/BaseModule.scala:28-120; Transformer and the
    Mask R-CNN heads are built this way in the reference): forward, parameters, training mode and state all
    delegate to that inner model."""

    def __init__(self):
        super().__init__()
        self.model = self.buildModel()
        self.modules = [self.model]

    def buildModel(self) -> Module:  # pragma: no cover - interface
        raise NotImplementedError

    def updateOutput(self, input):
        return self.model._call(input)


class DynamicContainer(Container):
    pass


# ====================================================================== graph
class Node:
    """A vertex of the model DAG: a module plus its predecessor nodes (S/utils/DirectedGraph.scala Node[T])."""

    def __init__(self, element: Module, prev: list["Node"] | None = None):
        self.element = element
        self.prev: list[Node] = list(prev or [])
        self.next: list[Node] = []
        self.edge_index: list[int | None] = [None] * len(self.prev)   # from_index (1-based) into the producer's Table
        for p in self.prev:
            p.next.append(self)

    # reference sugar: node -> node
    def add(self, other: "Node", from_index: int | None = None):
        other.prev.append(self)
        other.edge_index.append(from_index)
        self.next.append(other)
        return other

    def __rshift__(self, other):
        return self.add(other)

    def removePrevEdges(self):
        for p in self.prev:
            p.next = [n for n in p.next if n is not self]
        self.prev, self.edge_index = [], []
        return self

    remove_prev_edges = remove_pre_edges = removePreEdges = removePrevEdges

    def removeNextEdges(self):
        for n in self.next:
            keep = [i for i, p in enumerate(n.prev) if p is not self]
            n.prev = [n.prev[i] for i in keep]
            n.edge_index = [n.edge_index[i] for i in keep]
        self.next = []
        return self

    remove_next_edges = removeNextEdges

    @classmethod
    def of(cls, jvalue, bigdl_type="float"):
        """Python-API constructor from an existing value: a Node is returned as is, a module is wrapped."""
        return jvalue if isinstance(jvalue, Node) else Node(jvalue)

    def __repr__(self):
        return f"Node({self.element})"


class Input(Module):
    """Placeholder source node (S/nn/Input.scala). ``Input()`` returns a Node, as in the reference."""

    def __new__(cls, name: str | None = None):
        m = object.__new__(cls)
        Module.__init__(m)
        if name:
            m.setName(name)
        return Node(m)

    def updateOutput(self, input):
        return input


class Identity(Module):
    def updateOutput(self, input):
        return input


class Echo(Module):
    """Identity that prints activation shapes (debugging aid, S/nn/Echo.scala)."""

    def updateOutput(self, input):
        shapes = [tuple(t.shape) for t in flatten_activity(input)]
        print(f"{self.getName()} : Activation size is {shapes}")
        return input


def topo_sort(outputs: list[Node]) -> list[Node]:
    order, state = [], {}

    def visit(n: Node):
        st = state.get(id(n), 0)
        if st == 2:
            return
        if st == 1:
            raise ValueError("the graph has a cycle")
        state[id(n)] = 1
        for p in n.prev:
            visit(p)
        state[id(n)] = 2
        order.append(n)

    import sys

    old = sys.getrecursionlimit()
    sys.setrecursionlimit(max(old, 100000))
    try:
        for o in outputs:
            visit(o)
    finally:
        sys.setrecursionlimit(old)
    return order


class Graph(Container):
    """Static DAG of modules executed in topological order (S/nn/StaticGraph.scala:56-68; backward comes from autograd
    instead of the reference's explicit reverse traversal :128-152).

    ``Graph(input_nodes, output_nodes)``; ``Model`` is the Python-API alias (P/nn/layer.py:704).
    """

    def __init__(self, inputs, outputs, variables=None):
        super().__init__()
        self.input_nodes: list[Node] = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
        self.output_nodes: list[Node] = list(outputs) if isinstance(outputs, (list, tuple)) else [outputs]
        self._stop_gradient: set[str] = set()
        self._build()

    def _build(self):
        self.exec_order = topo_sort(self.output_nodes)
        names = set()
        self.modules = []
        for n in self.exec_order:
            if id(n.element) not in names:
                names.add(id(n.element))
                self.modules.append(n.element)
        for n in self.input_nodes:
            if n not in self.exec_order:
                raise ValueError("an input node is not connected to the outputs")

    def node(self, name: str) -> Node:
        for n in self.exec_order:
            if n.element.getName() == name:
                return n
        raise KeyError(name)

    def getForwardExecutions(self):
        """Nodes in execution order (Graph.scala:286)."""
        return list(self.exec_order)

    def getSortedForwardExecutions(self):
        """Execution order without the input placeholders at the front moved around — the same topological order."""
        return list(self.exec_order)

    def getStopGradientLayers(self):
        return set(self._stop_gradient)

    def resetModules(self):
        """Rebuild ``modules`` from the nodes after their elements were replaced (quantisation, conversion)."""
        self._build()
        return self

    def buildBackwardGraph(self):
        """The backward pass is derived by autograd from the forward execution; kept for API parity."""
        return self

    def findInput(self, node, input):
        """The activity ``node`` receives given the graph input (first element for single-input graphs)."""
        if node in self.input_nodes:
            if len(self.input_nodes) == 1:
                return input
            return to_table(input)[self.input_nodes.index(node) + 1]
        return None

    def toIRgraph(self):
        from ..utils.intermediate import BlasToIR

        return BlasToIR.convert(self)

    def setInputFormats(self, formats):
        self.inputFormats = list(formats)
        return self

    def setOutputFormats(self, formats):
        self.outputFormats = list(formats)
        return self

    def stopGradient(self, names):
        self._stop_gradient.update(names)
        return self

    stop_gradient = stopGradient

    def updateOutput(self, input):
        vals: dict[int, object] = {}
        if len(self.input_nodes) == 1:
            ins = [input]
        else:
            ins = to_table(input).to_list()
            if len(ins) != len(self.input_nodes):
                raise ValueError(f"graph expects {len(self.input_nodes)} inputs, got {len(ins)}")
        for n, v in zip(self.input_nodes, ins):
            vals[id(n)] = n.element._call(v)
        for n in self.exec_order:
            if id(n) in vals:
                continue
            args = []
            for p, idx in zip(n.prev, n.edge_index):
                v = vals[id(p)]
                if idx is not None:
                    v = to_table(v)[idx]
                args.append(v)
            x = args[0] if len(args) == 1 else (Table(*args) if args else None)
            y = n.element._call(x)
            if n.element.getName() in self._stop_gradient:
                y = _detach(y)
            vals[id(n)] = y
        outs = [vals[id(n)] for n in self.output_nodes]
        return outs[0] if len(outs) == 1 else Table(*outs)

    def saveGraphTopology(self, log_path: str, backward: bool = False):
        from ..visualization.summary import write_graph_topology

        write_graph_topology(self, log_path)
        return self

    save_graph_topology = saveGraphTopology


StaticGraph = Graph
Model = Graph
