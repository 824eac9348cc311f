"""Lua-table-like heterogeneous container used for Activities, optimizer state and configs.

Reference: S/utils/Table.scala (``T()`` builder, 1-based integer keys for array-style use, arbitrary keys for
state maps). Python lists/tuples are accepted wherever a Table is expected and are converted with ``to_table``.
"""
from __future__ import annotations

from typing import Any, Iterable

import torch


class Table:
    __slots__ = ("_d", "_top")

    def __init__(self, *items, **kw):
        self._d: dict = {}
        self._top = 0
        for it in items:
            self.insert(it)
        for k, v in kw.items():
            self._d[k] = v

    # ---- array-style access (1-based, like the reference)
    def insert(self, value, index: int | None = None):
        if index is None:
            self._top += 1
            self._d[self._top] = value
        else:
            n = self.length()
            for i in range(n, index - 1, -1):
                self._d[i + 1] = self._d[i]
            self._d[index] = value
            self._top = max(self._top, n + 1)
        return self

    def remove(self, index: int | None = None):
        n = self.length()
        if n == 0:
            return None
        if index is None:
            index = n
        v = self._d.pop(index)
        for i in range(index + 1, n + 1):
            self._d[i - 1] = self._d.pop(i)
        self._top = n - 1
        return v

    def length(self) -> int:
        n = 0
        while (n + 1) in self._d:
            n += 1
        return n

    def __len__(self):
        return self.length() if all(isinstance(k, int) for k in self._d) else len(self._d)

    def __getitem__(self, key):
        return self._d[key]

    def __setitem__(self, key, value):
        self._d[key] = value
        if isinstance(key, int) and key > self._top:
            self._top = key

    def __contains__(self, key):
        return key in self._d

    def __delitem__(self, key):
        del self._d[key]

    def delete(self, key):
        self._d.pop(key, None)
        return self

    def get(self, key, default=None):
        return self._d.get(key, default)

    def getOrElse(self, key, default):
        return self._d.get(key, default)

    get_or_else = getOrElse

    def update(self, key, value):
        self[key] = value
        return self

    def keys(self):
        return self._d.keys()

    def items(self):
        return self._d.items()

    def values(self):
        return self._d.values()

    def to_list(self) -> list:
        return [self._d[i] for i in range(1, self.length() + 1)]

    def __iter__(self):
        return iter(self.to_list())

    def clear(self):
        self._d.clear()
        self._top = 0
        return self

    def clone(self):
        t = Table()
        for k, v in self._d.items():
            t._d[k] = v.clone() if isinstance(v, (torch.Tensor, Table)) else v
        t._top = self._top
        return t

    def __eq__(self, other):
        if not isinstance(other, Table) or self._d.keys() != other._d.keys():
            return False
        for k, v in self._d.items():
            o = other._d[k]
            if isinstance(v, torch.Tensor):
                if not (isinstance(o, torch.Tensor) and v.shape == o.shape and torch.equal(v, o)):
                    return False
            elif v != o:
                return False
        return True

    def __repr__(self):
        return "T{" + ", ".join(f"{k}: {_short(v)}" for k, v in self._d.items()) + "}"


def _short(v):
    if isinstance(v, torch.Tensor):
        return f"Tensor{tuple(v.shape)}"
    return repr(v)


def T(*items, **kw) -> Table:
    """Builder mirroring the reference's ``T(a, b, ...)`` / ``T("key" -> v)``."""
    return Table(*items, **kw)


def to_table(x) -> Table:
    if isinstance(x, Table):
        return x
    if isinstance(x, (list, tuple)):
        return Table(*x)
    raise TypeError(f"cannot interpret {type(x)} as a Table")


def is_table(x) -> bool:
    return isinstance(x, (Table, list, tuple))


def flatten_activity(x) -> list:
    """Depth-first list of the tensors inside a Tensor|Table activity."""
    if isinstance(x, torch.Tensor):
        return [x]
    if x is None:
        return []
    out = []
    for v in (x.to_list() if isinstance(x, Table) else x):
        out.extend(flatten_activity(v))
    return out


def map_activity(fn, x):
    if isinstance(x, torch.Tensor):
        return fn(x)
    if x is None:
        return None
    if isinstance(x, Table):
        out = Table()
        for k, v in x.items():          # keeps non-positional keys ("bboxes", "masks", …)
            out[k] = map_activity(fn, v)
        return out
    if isinstance(x, (list, tuple)):
        if all(isinstance(v, (torch.Tensor, Table)) for v in x):
            return Table(*[map_activity(fn, v) for v in x])
        return x
    return x            # opaque leaf (string column, sparse tensor, …) travels unchanged


def unflatten_like(template, flat: Iterable[Any]):
    it = iter(flat)

    def rec(t):
        if isinstance(t, torch.Tensor):
            return next(it)
        if t is None:
            return None
        return Table(*[rec(v) for v in (t.to_list() if isinstance(t, Table) else t)])

    return rec(template)
