"""``model.toGraph()``: turn a container tree into an equivalent static Graph that shares the layers (and weights) of the
source model (reference: S/nn/abstractnn/AbstractModule.scala:1060-1100 ``toGraph``, S/nn/Sequential.scala ``getEndNodes``,
S/nn/Concat*.scala ``getEndNodes``). Sequential chains are unrolled into nodes; ConcatTable / ParallelTable / Concat
become fan-out + JoinTable structures; every other module is one node."""
from __future__ import annotations

from .module import Graph, Input, Module, Node, Sequential


def _end_nodes(m: Module, starts: list[Node]) -> list[Node]:
    from . import layer as L

    if isinstance(m, Graph):
        return [Node(m, starts)]
    if isinstance(m, Sequential):
        cur = starts
        for c in m.modules:
            cur = _end_nodes(c, cur)
        return cur
    if type(m) is L.ConcatTable:
        outs = []
        for c in m.modules:
            outs.extend(_end_nodes(c, starts))
        return outs                                  # several end nodes: the consumer takes them as a table
    if type(m) is L.Concat:
        outs = []
        for c in m.modules:
            outs.extend(_end_nodes(c, starts))
        join = L.JoinTable(m.dimension)
        if m.hasName():
            join.setName(m.getName())
        return [Node(join, outs)]
    return [Node(m, starts)]


def to_graph(model: Module, *start_nodes) -> Graph:
    if isinstance(model, Graph):
        return model
    starts = list(start_nodes) if start_nodes else [Input()]
    ends = _end_nodes(model, starts)
    g = Graph(starts, ends)
    if model.hasName():
        g.setName(model.getName())
    return g
