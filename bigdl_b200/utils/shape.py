"""Shape descriptors for Keras-style shape inference (S/utils/Shape.scala:24-129): ``SingleShape`` wraps one list of
ints (−1 = unknown batch), ``MultiShape`` a list of shapes (table-valued activities)."""
from __future__ import annotations


class Shape:
    def __new__(cls, *a):
        if cls is Shape:
            if len(a) == 1 and isinstance(a[0], (list, tuple)) and a[0] and isinstance(a[0][0], (Shape, list, tuple)):
                return MultiShape([s if isinstance(s, Shape) else SingleShape(list(s)) for s in a[0]])
            vals = a[0] if len(a) == 1 and isinstance(a[0], (list, tuple)) else a
            return SingleShape(list(vals))
        return object.__new__(cls)

    def toSingle(self) -> list[int]:
        raise ValueError("Cannot cast to single shape")

    def toMulti(self) -> list["Shape"]:
        raise ValueError("Cannot cast to multi shape")

    to_single, to_multi = lambda self: self.toSingle(), lambda self: self.toMulti()


class SingleShape(Shape):
    def __init__(self, *value):
        if len(value) == 1 and isinstance(value[0], (list, tuple)):
            value = value[0]
        self.value = [int(v) for v in value]

    def toSingle(self):
        return list(self.value)

    def copyAndUpdate(self, dim: int, v: int) -> "SingleShape":
        out = list(self.value)
        out[dim if dim >= 0 else len(out) + dim] = v
        return SingleShape(out)

    def __eq__(self, o):
        return isinstance(o, SingleShape) and o.value == self.value

    def __hash__(self):
        return hash(tuple(self.value))

    def __repr__(self):
        return f"SingleShape({self.value})"


class MultiShape(Shape):
    def __init__(self, value):
        self.value = [s if isinstance(s, Shape) else SingleShape(list(s)) for s in value]

    def toMulti(self):
        return list(self.value)

    def __eq__(self, o):
        return isinstance(o, MultiShape) and o.value == self.value

    def __hash__(self):
        return hash(tuple(self.value))

    def __repr__(self):
        return f"MultiShape({self.value})"
