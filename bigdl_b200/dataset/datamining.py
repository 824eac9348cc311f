"""Tabular rows → Samples (S/dataset/datamining/RowTransformer.scala:30-326): a ``RowTransformer`` is a set of
``RowTransformSchema``s, each selecting columns (by name or index) of a row and writing them into one tensor of a Table.
Rows are dicts / pandas rows / sequences (Spark SQL ``Row`` in the reference)."""
from __future__ import annotations

from typing import Sequence

import numpy as np
import torch

from ..utils.table import Table
from .base import Transformer


class RowTransformSchema:
    def __init__(self, schema_key: str, field_names: Sequence[str] = (), indices: Sequence[int] = ()):
        self.schemaKey, self.fieldNames, self.indices = schema_key, list(field_names), list(indices)

    def transform(self, values: list) -> torch.Tensor:
        return torch.tensor([float(v) for v in values], dtype=torch.float32)


class ColToTensor(RowTransformSchema):
    """One column → one size-1 tensor keyed by the column name."""

    def __init__(self, schema_key: str, field_name: str):
        super().__init__(schema_key, [field_name])


class ColsToNumeric(RowTransformSchema):
    """Several numeric columns → one 1-D tensor."""


class RowTransformer(Transformer):
    def __init__(self, schemas: Sequence[RowTransformSchema], row_size: int | None = None):
        keys = [s.schemaKey for s in schemas]
        if len(set(keys)) != len(keys):
            raise ValueError("RowTransformer: found replicated schemaKey")
        self.schemas, self.row_size = list(schemas), row_size

    @staticmethod
    def atomic(field_names: Sequence[str]):
        return RowTransformer([ColToTensor(n, n) for n in field_names])

    @staticmethod
    def numeric(schema_key: str = "all", field_names: Sequence[str] = ()):
        return RowTransformer([ColsToNumeric(schema_key, field_names)])

    def _get(self, row, schema):
        if hasattr(row, "_asdict"):
            row = row._asdict()
        if isinstance(row, dict) or hasattr(row, "keys"):
            names = list(row.keys())
            if schema.fieldNames:
                return [row[n] for n in schema.fieldNames]
            if schema.indices:
                return [row[names[i]] for i in schema.indices]
            return [row[n] for n in names]
        seq = list(row)
        if schema.indices:
            return [seq[i] for i in schema.indices]
        return seq

    def __call__(self, it):
        for row in it:
            t = Table()
            for s in self.schemas:
                t[s.schemaKey] = s.transform(self._get(row, s))
            yield t
