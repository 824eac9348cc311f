"""TensorFlow-specific helper ops used by graphs imported from ``.pb`` files (S/nn/tf/*.scala, 18 files): constants
and shape ops, bias add / strided slice, control flow (Switch / Merge / Enter / Exit / NextIteration / LoopCond),
control dependencies, variables, tensor arrays and stacks (DataFlowOps.scala), example parsing and image decoding.
Control-flow ops are executed by ``DynamicGraph`` (nn/dynamic_graph.py), which implements TensorFlow's dead-tensor
dataflow semantics."""
from __future__ import annotations

import numpy as np
import torch

from ..utils.table import Table
from .module import Module
from .ops import Operation, _ins


class _Dead:
    """Marker travelling along untaken Switch branches."""

    def __repr__(self):
        return "<dead>"


DEAD = _Dead()


# ------------------------------------------------------------------------------------------------ basic ops
class Const(Operation):
    def __init__(self, value):
        super().__init__()
        self.value = value if isinstance(value, (torch.Tensor, np.ndarray, str, bytes)) else torch.as_tensor(value)

    def updateOutput(self, input):
        return self.value


class Fill(Operation):
    def updateOutput(self, input):
        shape, v = _ins(input, 2)
        return torch.full([int(s) for s in torch.as_tensor(shape).reshape(-1).tolist()], torch.as_tensor(v).item(),
                          dtype=torch.as_tensor(v).dtype)


class Shape(Operation):
    def updateOutput(self, input):
        return torch.tensor(list(input.shape), dtype=torch.int32)


class NoOp(Operation):
    def updateOutput(self, input):
        return input


class ControlDependency(Operation):
    """Pure ordering edge: forwards nothing (ControlDependency.scala)."""

    def updateOutput(self, input):
        return None


class Assert(Operation):
    def updateOutput(self, input):
        items = _ins(input)
        if not bool(torch.as_tensor(items[0]).all()):
            msg = items[1] if len(items) > 1 else "assertion failed"
            raise AssertionError(msg.decode() if isinstance(msg, bytes) else str(msg))
        return None


class BiasAdd(Module):
    """value + bias on the last dimension (trainable inputs: gradients flow to both)."""

    def updateOutput(self, input):
        v, b = _ins(input, 2)
        return v + b


class Log1p(Module):
    def updateOutput(self, input):
        return torch.log1p(input)


class StridedSlice(Module):
    """``specs``: [(dim (1-based), begin (1-based), end (exclusive), stride)] as the loader computes them from the
    begin/end/strides tensors and masks (StridedSlice.scala)."""

    def __init__(self, specs):
        super().__init__()
        self.specs = [tuple(s) for s in specs]

    def updateOutput(self, input):
        idx = [slice(None)] * input.dim()
        for dim, b, e, s in self.specs:
            idx[dim - 1] = slice(b - 1, e - 1, s)
        return input[tuple(idx)]


class SplitAndSelect(Module):
    """Split ``dimension`` into ``num_split`` equal parts and return part ``index`` (both 1-based)."""

    def __init__(self, dimension: int, index: int, num_split: int):
        super().__init__()
        self.dimension, self.index, self.num_split = dimension, index, num_split

    def updateOutput(self, input):
        d = self.dimension - 1 if self.dimension > 0 else input.dim() + self.dimension
        n = input.shape[d] // self.num_split
        return input.narrow(d, (self.index - 1) * n, n)


class InvertPermutation(Operation):
    def updateOutput(self, input):
        p = input.long()
        out = torch.empty_like(p)
        out[p] = torch.arange(p.numel())
        return out.to(input.dtype)


class ConcatOffset(Operation):
    def updateOutput(self, input):
        items = _ins(input)
        dim = int(torch.as_tensor(items[0]).item())
        off, outs = 0, []
        for s in items[1:]:
            o = torch.zeros_like(torch.as_tensor(s))
            o[dim] = off
            off += int(s[dim])
            outs.append(o)
        return Table(*outs)


class BroadcastGradientArgs(Operation):
    """Reduction axes needed to un-broadcast gradients of a binary op (MathOps / ArrayOps in the reference)."""

    def updateOutput(self, input):
        a, b = [torch.as_tensor(v).reshape(-1).tolist() for v in _ins(input, 2)]
        n = max(len(a), len(b))
        a, b = [1] * (n - len(a)) + a, [1] * (n - len(b)) + b
        # TensorFlow's BCast also lists a dimension where BOTH sides are 1 (reducing over it is a no-op)
        ra = [i for i in range(n) if a[i] == 1]
        rb = [i for i in range(n) if b[i] == 1]
        return Table(torch.tensor(ra, dtype=torch.int32), torch.tensor(rb, dtype=torch.int32))


# ------------------------------------------------------------------------------------------------ state
class Variable(Module):
    """A trainable tensor living in the graph (StateOps.scala ``Variable``)."""

    def __init__(self, initial_value, trainable: bool = True):
        super().__init__()
        self.register_parameter("weight", torch.as_tensor(initial_value, dtype=torch.float32).clone())
        self.trainable = trainable

    def updateOutput(self, input):
        return self.weight if self.trainable else self.weight.detach()


class Assign(Operation):
    def updateOutput(self, input):
        ref, value = _ins(input, 2)
        with torch.no_grad():
            ref.copy_(value)
        return ref


# ------------------------------------------------------------------------------------------------ control flow
class ControlOps(Operation):
    pass


class SwitchOps(ControlOps):
    """(data, pred) → Table(false_output, true_output); the untaken side is DEAD (ControlOps.scala:65-85)."""

    def updateOutput(self, input):
        data, pred = _ins(input, 2)
        p = bool(torch.as_tensor(pred).reshape(-1)[0])
        return Table(DEAD, data) if p else Table(data, DEAD)


class MergeOps(ControlOps):
    """Forwards whichever input is alive."""

    def updateOutput(self, input):
        for v in _ins(input):
            if v is not DEAD and v is not None:
                return v
        return DEAD


class Enter(ControlOps):
    def __init__(self, frame: str = "", is_constant: bool = False):
        super().__init__()
        self.frame, self.is_constant = frame, is_constant

    def updateOutput(self, input):
        return input


class Exit(ControlOps):
    def updateOutput(self, input):
        return input


class NextIteration(ControlOps):
    def updateOutput(self, input):
        return input


class LoopCondition(ControlOps):
    def updateOutput(self, input):
        return input


class ControlNodes:
    """Graph-building helpers (ControlOps.scala:240-330)."""

    @staticmethod
    def switch(data, condition):
        from .module import Node

        return Node(SwitchOps(), [data, condition])

    @staticmethod
    def false_edge(switch_node):
        return (switch_node, 1)

    @staticmethod
    def true_edge(switch_node):
        return (switch_node, 2)

    @staticmethod
    def merge(*nodes):
        from .module import Node

        n = Node(MergeOps(), [])
        for item in nodes:
            src, idx = item if isinstance(item, tuple) else (item, None)
            src.add(n, idx)
        return n

    @staticmethod
    def whileLoop(condition_fn, body_fn, loop_vars, frame: str = "while"):
        """Build ``while cond(*vars): vars = body(*vars)``; ``loop_vars`` are nodes, ``condition_fn(merge_nodes)`` →
        node producing a boolean, ``body_fn(true-branch nodes)`` → list of nodes. Returns the Exit nodes."""
        from .module import Identity, Node

        enters = [Node(Enter(frame), [v]) for v in loop_vars]
        merges = [Node(MergeOps(), [e]) for e in enters]
        cond = Node(LoopCondition(), [condition_fn(merges)])
        switches = [Node(SwitchOps(), [m, cond]) for m in merges]
        body_in = []
        for s in switches:
            idn = Node(Identity(), [])
            s.add(idn, 2)
            body_in.append(idn)
        body_out = body_fn(body_in)
        for m, o in zip(merges, body_out):
            Node(NextIteration(), [o]).add(m)
        exits = []
        for s in switches:
            ex = Node(Exit(), [])
            s.add(ex, 1)
            exits.append(ex)
        return exits


# ------------------------------------------------------------------------------------------------ tensor arrays / stacks
class _TensorArrayStore:
    """The resource behind a TensorArray handle (DataFlowOps.scala:45-140). A gradient array (``grad()``) has the
    locked size of its source and SUMS repeated writes to one index, which is how TF accumulates the gradients of a
    value read several times."""

    def __init__(self, size=0, dynamic=True, aggregate=False):
        self.items: list = [None] * size
        self.dynamic = dynamic
        self.aggregate = aggregate
        self.grads: dict = {}

    def write(self, i, v):
        while i >= len(self.items):
            if not self.dynamic:
                raise IndexError("TensorArray index out of range")
            self.items.append(None)
        if self.aggregate and self.items[i] is not None:
            v = self.items[i] + v
        self.items[i] = v

    def grad(self, source: str = ""):
        self.dynamic = False
        if source not in self.grads:
            self.grads[source] = _TensorArrayStore(len(self.items), False, True)
        return self.grads[source]

    def release(self):
        self.items = []
        self.grads = {}


TensorArray = _TensorArrayStore


class TensorArrayCreator(Operation):
    def __init__(self, dynamic_size: bool = True):
        super().__init__()
        self.dynamic = dynamic_size

    def updateOutput(self, input):
        n = int(torch.as_tensor(input).reshape(-1)[0]) if input is not None else 0
        return Table(_TensorArrayStore(n, self.dynamic), torch.tensor(0.0))        # (handle, flow)


class TensorArrayWrite(Operation):
    def updateOutput(self, input):
        h, idx, v, flow = _ins(input, 4)
        h.write(int(torch.as_tensor(idx).item()), v)
        return flow


class TensorArrayRead(Operation):
    def updateOutput(self, input):
        h, idx, _ = _ins(input, 3)
        return h.items[int(torch.as_tensor(idx).item())]


class TensorArrayGather(Operation):
    def updateOutput(self, input):
        h, idx, _ = _ins(input, 3)
        return torch.stack([h.items[int(i)] for i in torch.as_tensor(idx).reshape(-1).tolist()])


class TensorArrayScatter(Operation):
    def updateOutput(self, input):
        h, idx, v, flow = _ins(input, 4)
        for k, i in enumerate(torch.as_tensor(idx).reshape(-1).tolist()):
            h.write(int(i), v[k])
        return flow


class TensorArrayConcat(Operation):
    def updateOutput(self, input):
        h, _ = _ins(input, 2)
        return Table(torch.cat(h.items, 0), torch.tensor([t.shape[0] for t in h.items], dtype=torch.int64))


class TensorArraySplit(Operation):
    def updateOutput(self, input):
        h, v, lens, flow = _ins(input, 4)
        pos = 0
        for i, n in enumerate(torch.as_tensor(lens).reshape(-1).tolist()):
            h.write(i, v[pos:pos + int(n)])
            pos += int(n)
        return flow


class TensorArraySize(Operation):
    def updateOutput(self, input):
        h, _ = _ins(input, 2)
        return torch.tensor(len(h.items), dtype=torch.int32)


class TensorArrayGrad(Operation):
    """(handle, flow) → (handle of the gradient array registered under ``source``, flow)
    (DataFlowOps.scala:230-248)."""

    def __init__(self, source: str = ""):
        super().__init__()
        self.source = source

    def updateOutput(self, input):
        h = _ins(input)[0]
        return Table(h.grad(self.source), torch.tensor(0.0))


class TensorArrayClose(Operation):
    """Frees the array behind the handle (DataFlowOps.scala:562-577)."""

    def updateOutput(self, input):
        h = input if isinstance(input, _TensorArrayStore) else _ins(input)[0]
        h.release()
        return torch.tensor(0.0)


class _StackStore:
    """Handle of a TF stack resource (a plain list would be flattened into a Table when passed between nodes)."""

    def __init__(self):
        self.items: list = []


Stack = _StackStore


class StackCreator(Operation):
    def updateOutput(self, input):
        return _StackStore()


class StackPush(Operation):
    def updateOutput(self, input):
        h, v = _ins(input, 2)
        h.items.append(v)
        return v


class StackPop(Operation):
    def updateOutput(self, input):
        h = input if isinstance(input, _StackStore) else _ins(input)[0]
        return h.items.pop()


# ------------------------------------------------------------------------------------------------ parsing / images
def _parse_feature(buf):
    from ..visualization.proto import parse
    import struct

    for f, wt, v in parse(buf):
        vals = []
        if f == 1:        # BytesList
            vals = [x for k, _, x in parse(v) if k == 1]
            return "bytes", vals
        if f == 2:        # FloatList (packed or not)
            for k, w2, x in parse(v):
                if k == 1:
                    vals += list(struct.unpack(f"<{len(x) // 4}f", x)) if w2 == 2 else [x]
            return "float", vals
        if f == 3:        # Int64List
            from ..visualization.proto import _read_varint

            for k, w2, x in parse(v):
                if k == 1:
                    if w2 == 2:
                        i = 0
                        while i < len(x):
                            n, i = _read_varint(x, i)
                            vals.append(n - (1 << 64) if n >= 1 << 63 else n)
                    else:
                        vals.append(x - (1 << 64) if x >= 1 << 63 else x)
            return "int64", vals
    return "bytes", []


def parse_example(serialized: bytes) -> dict:
    """Decode a ``tf.train.Example`` into {name: (kind, values)} (used by ParseExample, ParsingOps.scala)."""
    from ..visualization.proto import parse

    out = {}
    for f, _, features in parse(serialized):
        if f != 1:
            continue
        for f2, _, entry in parse(features):
            if f2 != 1:
                continue
            key, val = None, None
            for f3, _, x in parse(entry):
                if f3 == 1:
                    key = x.decode()
                elif f3 == 2:
                    val = _parse_feature(x)
            out[key] = val
    return out


class ParseExample(Operation):
    """(serialized batch [N] bytes, names, dense_keys…, defaults…) → Table of dense tensors, one per key."""

    def __init__(self, n_dense: int, dense_types, dense_shapes):
        super().__init__()
        self.n, self.types, self.shapes = n_dense, list(dense_types), [list(s) for s in dense_shapes]

    def updateOutput(self, input):
        items = _ins(input)
        serialized = np.asarray(items[0], dtype=object).reshape(-1)
        keys = [k.decode() if isinstance(k, bytes) else str(k) for k in items[2:2 + self.n]]
        cols = [[] for _ in keys]
        for rec in serialized:
            ex = parse_example(rec)
            for i, k in enumerate(keys):
                kind, vals = ex[k]
                cols[i].append(vals)
        outs = []
        for i, c in enumerate(cols):
            if self.types[i] == "string":
                outs.append(np.array(c, dtype=object).reshape([len(c)] + self.shapes[i]))
            else:
                dt = torch.float32 if self.types[i] == "float" else torch.int64
                outs.append(torch.tensor(c, dtype=dt).reshape([len(c)] + self.shapes[i]))
        return Table(*outs)


class ParseSingleExample(Operation):
    """ONE serialized ``tf.train.Example`` → Table of its dense features, in the order of ``dense_keys`` (the keys are
    attributes of the op, not inputs) with the given types and shapes; a key the record does not carry is skipped
    (ParsingOps.scala:93-140)."""

    def __init__(self, dense_types, dense_keys, dense_shapes):
        super().__init__()
        self.types = list(dense_types)
        self.keys = [k.decode() if isinstance(k, bytes) else str(k) for k in dense_keys]
        self.shapes = [list(s) for s in dense_shapes]
        while len(self.shapes) < len(self.keys):
            self.shapes.append([])

    def updateOutput(self, input):
        first = _ins(input)[0] if isinstance(input, (Table, list, tuple)) else input
        rec = first if isinstance(first, (bytes, bytearray)) else np.asarray(first, dtype=object).reshape(-1)[0]
        ex = parse_example(rec)
        outs = []
        for key, kind, shape in zip(self.keys, self.types, self.shapes):
            if key not in ex:
                continue
            found, vals = ex[key]
            kind = kind or {"bytes": "string", "float": "float"}.get(found, "int64")
            if kind == "string":
                outs.append(np.array(vals, dtype=object).reshape(shape if shape else [len(vals)]))
            else:
                dt = torch.float32 if kind == "float" else torch.int64
                t = torch.tensor(vals, dtype=dt)
                outs.append(t.reshape(shape) if shape else t)
        return Table(*outs)


class DecodeImage(Operation):
    """JPEG / PNG / BMP / GIF bytes → uint8 HWC (RGB, like TensorFlow) (ImageOps.scala)."""

    def __init__(self, channels: int = 0):
        super().__init__()
        self.channels = channels

    def updateOutput(self, input):
        import cv2

        data = input if isinstance(input, (bytes, bytearray)) else np.asarray(input, dtype=object).reshape(-1)[0]
        flag = cv2.IMREAD_GRAYSCALE if self.channels == 1 else cv2.IMREAD_COLOR
        img = cv2.imdecode(np.frombuffer(data, dtype=np.uint8), flag)
        if img is None:
            raise ValueError("cannot decode image")
        if img.ndim == 2:
            img = img[..., None]
        else:
            img = img[..., ::-1]
        return torch.from_numpy(np.ascontiguousarray(img))


DecodeJpeg = DecodePng = DecodeBmp = DecodeGif = DecodeImage


class DecodeRaw(Operation):
    def __init__(self, out_type=torch.uint8, little_endian: bool = True):
        super().__init__()
        self.out_type, self.le = out_type, little_endian

    def updateOutput(self, input):
        data = input if isinstance(input, (bytes, bytearray)) else np.asarray(input, dtype=object).reshape(-1)[0]
        np_dt = {torch.uint8: "u1", torch.int8: "i1", torch.int16: "i2", torch.int32: "i4", torch.int64: "i8",
                 torch.float32: "f4", torch.float64: "f8"}[self.out_type]
        arr = np.frombuffer(data, dtype=np.dtype(("<" if self.le else ">") + np_dt if np_dt[1] != "1" else np_dt))
        return torch.from_numpy(arr.astype(arr.dtype.newbyteorder("=")).copy())


def __getattr__(name):
    """The backward-half / TF-layout operations live in ``nn.tf_grad`` (they are built on the op loaders, which import
    this module); ``nn.tf.ReluGrad`` etc. resolve there lazily."""
    from . import tf_grad

    try:
        return getattr(tf_grad, name)
    except AttributeError:
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}") from None
