"""COO sparse 2-D tensor (S/tensor/SparseTensor.scala, SparseTensorBLAS.scala, SparseTensorMath.scala) used by
SparseLinear, LookupTableSparse and SparseJoinTable. Indices are stored 0-based internally (torch sparse COO) and
exposed 1-based through ``indices()`` like the reference's ``Tensor.sparse(indices, values, shape)`` factory."""
from __future__ import annotations

import numpy as np
import torch


class SparseTensor:
    def __init__(self, indices, values, shape, one_based: bool = False):
        idx = torch.as_tensor(np.asarray(indices), dtype=torch.long)
        if idx.dim() == 2 and idx.shape[0] != len(shape):
            idx = idx.t()
        if one_based:
            idx = idx - 1
        val = torch.as_tensor(np.asarray(values) if not isinstance(values, torch.Tensor) else values)
        if val.dtype == torch.float64:
            val = val.float()             # the reference's default numeric type is Float
        import warnings

        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            self.sp = torch.sparse_coo_tensor(idx, val, tuple(shape), check_invariants=False).coalesce()

    @staticmethod
    def from_dense(t) -> "SparseTensor":
        from .tensor import _raw

        d = _raw(t)
        idx = d.nonzero().t()
        return SparseTensor(idx, d[tuple(idx)], d.shape)

    def size(self, d: int | None = None):
        return list(self.sp.shape) if d is None else self.sp.shape[d - 1]

    def nDimension(self) -> int:
        return self.sp.dim()

    def nElement(self) -> int:
        return self.sp._nnz()

    def indices(self, one_based: bool = True) -> torch.Tensor:
        return self.sp.indices() + (1 if one_based else 0)

    def values(self) -> torch.Tensor:
        return self.sp.values()

    def to_dense(self):
        from .tensor import Tensor

        return Tensor(self.sp.to_dense())

    toDense = to_dense

    def numNonZeroByRow(self) -> list[int]:
        rows = self.sp.indices()[0]
        return torch.bincount(rows, minlength=self.sp.shape[0]).tolist()

    def narrow(self, dim: int, index: int, size: int) -> "SparseTensor":
        idx, val = self.sp.indices(), self.sp.values()
        d = dim - 1
        keep = (idx[d] >= index - 1) & (idx[d] < index - 1 + size)
        nidx = idx[:, keep].clone()
        nidx[d] -= index - 1
        shape = list(self.sp.shape)
        shape[d] = size
        return SparseTensor(nidx, val[keep], shape)

    # ---- SparseTensorBLAS
    def mm(self, dense) -> torch.Tensor:
        """sparse [m,k] @ dense [k,n] (coomm)."""
        from .tensor import _raw

        return torch.sparse.mm(self.sp, _raw(dense))

    def mv(self, vec) -> torch.Tensor:
        from .tensor import _raw

        return torch.sparse.mm(self.sp, _raw(vec).reshape(-1, 1)).reshape(-1)

    def t(self) -> "SparseTensor":
        idx = self.sp.indices().flip(0)
        return SparseTensor(idx, self.sp.values(), self.sp.shape[::-1])

    @staticmethod
    def concat(dim: int, tensors) -> "SparseTensor":
        """SparseJoinTable: concatenate along ``dim`` (1-based)."""
        d = dim - 1
        idxs, vals, off = [], [], 0
        shape = list(tensors[0].sp.shape)
        for t in tensors:
            i = t.sp.indices().clone()
            i[d] += off
            off += t.sp.shape[d]
            idxs.append(i)
            vals.append(t.sp.values())
        shape[d] = off
        return SparseTensor(torch.cat(idxs, 1), torch.cat(vals), shape)

    def __repr__(self):
        return f"SparseTensor(shape={list(self.sp.shape)}, nnz={self.sp._nnz()})"
