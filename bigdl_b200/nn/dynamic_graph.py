"""DynamicGraph: a model DAG executed by a data-flow scheduler instead of a fixed topological order, so that
TensorFlow control flow (Switch / Merge conditionals, Enter / NextIteration / Exit while-loops) can run.

Reference: S/nn/DynamicGraph.scala:30-144, S/nn/Scheduler.scala (ready queue; a node fires when its inputs arrived,
Merge fires on the first live input), S/nn/FrameManager.scala (loop frames). Semantics are TensorFlow's dead-tensor
rules: a Switch marks its untaken output DEAD, a node with a DEAD input produces DEAD without running, Merge forwards the
live input, a DEAD NextIteration is dropped (that is how a loop ends), and loop-invariant ``Enter`` values stay available
to every iteration. Each firing calls the module through autograd, so — unlike the reference, which refuses to build a
backward graph when control ops are present — gradients flow through whatever path actually executed."""
from __future__ import annotations

from collections import deque

import torch

from ..utils.table import Table
from .module import Graph, Node
from .tf import DEAD, ControlOps, Enter, Exit, MergeOps, NextIteration, SwitchOps


class DynamicGraph(Graph):
    def __init__(self, inputs, outputs, variables=None, generate_backward: bool = True):
        self._dynamic = True
        super().__init__(inputs, outputs, variables)
        self.generate_backward = generate_backward

    def _build(self):
        """Collect nodes reachable backwards from the outputs; cycles (NextIteration → Merge) are allowed."""
        seen, order, stack = set(), [], list(self.output_nodes)
        while stack:
            n = stack.pop()
            if id(n) in seen:
                continue
            seen.add(id(n))
            order.append(n)
            stack.extend(n.prev)
        self.exec_order = order[::-1]
        self.modules = [n.element for n in self.exec_order]
        self._has_control = any(isinstance(m, ControlOps) for m in self.modules)
        if not self._has_control:
            super()._build()
        self._by_name = {n.element.getName(): n for n in self.exec_order}

    def updateOutput(self, input):
        if not self._has_control:
            return super().updateOutput(input)
        return self._run_dataflow(input)

    # ------------------------------------------------------------------ scheduler
    def _run_dataflow(self, input, max_firings: int = 10_000_000):
        ins = [input] if len(self.input_nodes) == 1 else (input.to_list() if isinstance(input, Table) else list(input))
        if len(ins) != len(self.input_nodes):
            raise ValueError(f"graph expects {len(self.input_nodes)} inputs, got {len(ins)}")
        member = {id(n) for n in self.exec_order}
        consumers: dict[int, list] = {id(n): [] for n in self.exec_order}
        for n in self.exec_order:
            for slot, (p, ei) in enumerate(zip(n.prev, n.edge_index)):
                if id(p) in member:
                    consumers[id(p)].append((n, slot, ei))
        received: dict[int, dict[int, object]] = {id(n): {} for n in self.exec_order}
        sticky: dict[int, set] = {id(n): set() for n in self.exec_order}
        results: dict[int, object] = {}
        queue = deque()
        for n, v in zip(self.input_nodes, ins):
            queue.append((n, v, True))
        feed_ids = {id(n) for n in self.input_nodes}
        for n in self.exec_order:
            if not n.prev and id(n) not in feed_ids:
                queue.append((n, None, True))
        firings = 0
        while queue:
            node, arg, live = queue.popleft()
            firings += 1
            if firings > max_firings:
                raise RuntimeError("DynamicGraph: firing budget exhausted (non-terminating loop?)")
            m = node.element
            out = m._call(arg) if live else DEAD
            if isinstance(m, NextIteration) and out is DEAD:
                continue                                   # the loop is over: nothing re-enters the Merge
            if out is not DEAD or id(node) not in results:
                results[id(node)] = out
            for c, slot, ei in consumers[id(node)]:
                v = out
                if v is not DEAD and ei is not None:
                    v = v[ei] if isinstance(v, Table) else (v if ei == 1 else DEAD)
                cm = c.element
                if isinstance(cm, MergeOps):
                    rec = received[id(c)]
                    if v is not DEAD:
                        received[id(c)] = {}
                        queue.append((c, Table(v), True))
                    else:
                        rec[slot] = DEAD
                        if len(rec) == len(c.prev):
                            received[id(c)] = {}
                            queue.append((c, None, False))
                    continue
                rec = received[id(c)]
                rec[slot] = v
                if isinstance(m, Enter) and m.is_constant or (isinstance(m, Enter) and not isinstance(cm, MergeOps)):
                    sticky[id(c)].add(slot)
                if len(rec) == len(c.prev):
                    vals = [rec[i] for i in range(len(c.prev))]
                    received[id(c)] = {s: rec[s] for s in sticky[id(c)]}
                    alive = all(x is not DEAD for x in vals)
                    arg2 = (vals[0] if len(vals) == 1 else Table(*vals)) if alive else None
                    queue.append((c, arg2, alive))
        outs = []
        for n in self.output_nodes:
            v = results.get(id(n), DEAD)
            if v is DEAD:
                raise RuntimeError(f"output node {n.element.getName()} never produced a value")
            outs.append(v)
        return outs[0] if len(outs) == 1 else Table(*outs)


class Scheduler:
    """Names the reference gives the two halves of its dynamic execution (S/nn/Scheduler.scala, FrameManager.scala): a
    ready-queue of nodes and the bookkeeping of loop frames. Here one data-flow loop (``DynamicGraph._run_dataflow``)
    does both — tokens carry a live / dead flag, ``Enter`` / ``Exit`` / ``NextIteration`` re-queue their consumers —
    so these two classes are small views over a graph for code that asks for them by name."""

    def __init__(self, graph: DynamicGraph):
        self.graph = graph

    def nodes(self):
        return list(self.graph.exec_order)

    def run(self, input):
        return self.graph._run_dataflow(input)


class FrameManager:
    """Loop frames of a graph: {frame name: [Enter nodes]} (FrameManager.scala)."""

    def __init__(self, graph: DynamicGraph):
        self.frames: dict = {}
        for n in graph.exec_order:
            m = n.element
            if type(m).__name__ == "Enter":
                self.frames.setdefault(getattr(m, "frame", getattr(m, "frame_name", "")), []).append(n)

    def names(self):
        return list(self.frames)
