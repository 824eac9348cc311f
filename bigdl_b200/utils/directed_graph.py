"""Generic DAG utilities (S/utils/DirectedGraph.scala:24-370): topological sort, DFS, BFS, reverse, clone. ``Graph`` in
nn/module.py keeps its own light ``Node``; this is the general container the loaders (Caffe/TF/ONNX) build first."""
from __future__ import annotations

from collections import deque


class Edge:
    def __init__(self, fromIndex: int | None = None):
        self.fromIndex = fromIndex


class Node:
    def __init__(self, element):
        self.element = element
        self.nexts: list[tuple["Node", Edge]] = []
        self.prevs: list[tuple["Node", Edge]] = []

    def add(self, node: "Node", edge: Edge | None = None) -> "Node":
        e = edge or Edge()
        if not any(n is node for n, _ in self.nexts):
            self.nexts.append((node, e))
        if not any(n is self for n, _ in node.prevs):
            node.prevs.append((self, e))
        return node

    __rshift__ = add

    def delete(self, node: "Node"):
        self.nexts = [(n, e) for n, e in self.nexts if n is not node]
        node.prevs = [(n, e) for n, e in node.prevs if n is not self]
        return self

    def nextNodes(self):
        return [n for n, _ in self.nexts]

    def prevNodes(self):
        return [n for n, _ in self.prevs]

    def removePrevEdges(self):
        for p, _ in list(self.prevs):
            p.nexts = [(n, e) for n, e in p.nexts if n is not self]
        self.prevs = []
        return self

    def graph(self, reverse: bool = False) -> "DirectedGraph":
        return DirectedGraph(self, reverse)

    def __repr__(self):
        return f"(Node {self.element})"


class DirectedGraph:
    def __init__(self, source: Node, reverse: bool = False):
        self.source, self.reverse = source, reverse

    def _next(self, n: Node):
        return n.prevNodes() if self.reverse else n.nextNodes()

    def _prev(self, n: Node):
        return n.nextNodes() if self.reverse else n.prevNodes()

    def size(self) -> int:
        return sum(1 for _ in self.BFS())

    def edges(self) -> int:
        return sum(len(self._next(n)) for n in self.BFS())

    def BFS(self):
        seen, q = {id(self.source)}, deque([self.source])
        while q:
            n = q.popleft()
            yield n
            for m in self._next(n):
                if id(m) not in seen:
                    seen.add(id(m))
                    q.append(m)

    def DFS(self):
        seen, st = set(), [self.source]
        while st:
            n = st.pop()
            if id(n) in seen:
                continue
            seen.add(id(n))
            yield n
            st.extend(m for m in self._next(n) if id(m) not in seen)

    def topologySort(self) -> list[Node]:
        """Kahn's algorithm over the nodes reachable from the source; raises on a cycle."""
        nodes = list(self.BFS())
        ids = {id(n) for n in nodes}
        indeg = {id(n): sum(1 for p in self._prev(n) if id(p) in ids) for n in nodes}
        ready = [n for n in nodes if indeg[id(n)] == 0]
        out = []
        while ready:
            n = ready.pop()
            out.append(n)
            for m in self._next(n):
                indeg[id(m)] -= 1
                if indeg[id(m)] == 0:
                    ready.append(m)
        if len(out) != len(nodes):
            raise ValueError("There's a cycle in the graph")
        return out

    def cloneGraph(self, reverseEdge: bool = False) -> "DirectedGraph":
        mapping = {id(n): Node(n.element) for n in self.BFS()}
        for n in self.BFS():
            for m, e in n.nexts:
                if id(m) in mapping:
                    if reverseEdge:
                        mapping[id(m)].add(mapping[id(n)], Edge(e.fromIndex))
                    else:
                        mapping[id(n)].add(mapping[id(m)], Edge(e.fromIndex))
        return DirectedGraph(mapping[id(self.source)], self.reverse)
