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

    # ---- the rest of the SparseTensor surface the reference implements (most TensorMath calls are unsupported there too)
    def apply(self, indexes):
        """Value at 1-based ``indexes`` (0 when no entry is stored)."""
        idx = torch.tensor([i - 1 for i in indexes], dtype=torch.long).reshape(-1, 1)
        hit = (self.sp.indices() == idx).all(0).nonzero()
        return float(self.sp.values()[hit[0, 0]]) if hit.numel() else 0.0

    def getTensorType(self) -> str:
        return "SparseType"

    def dim(self) -> int:
        return self.sp.dim()

    def sum(self) -> float:
        return float(self.sp.values().sum())

    def clone(self) -> "SparseTensor":
        return SparseTensor(self.sp.indices().clone(), self.sp.values().clone(), self.sp.shape)

    def copy(self, other: "SparseTensor") -> "SparseTensor":
        self.sp = other.sp.clone()
        return self

    set = copy

    def resize(self, sizes, nElement: int | None = None) -> "SparseTensor":
        """New logical shape; stored entries that fall outside are dropped (the reference re-uses the storage)."""
        idx, val = self.sp.indices(), self.sp.values()
        sizes = list(sizes)
        if len(sizes) != idx.shape[0]:
            idx, val = torch.zeros((len(sizes), 0), dtype=torch.long), val[:0]
        keep = torch.ones(idx.shape[1], dtype=torch.bool)
        for d, n in enumerate(sizes):
            keep &= idx[d] < n
        self.sp = torch.sparse_coo_tensor(idx[:, keep], val[keep], tuple(sizes)).coalesce()
        return self

    def cast(self, dtype: torch.dtype) -> "SparseTensor":
        return SparseTensor(self.sp.indices(), self.sp.values().to(dtype), self.sp.shape)

    def dot(self, dense) -> float:
        return SparseTensorBLAS.vdot(dense, self)

    def __repr__(self):
        return f"SparseTensor(shape={list(self.sp.shape)}, nnz={self.sp._nnz()})"


def _dense(t) -> torch.Tensor:
    from .tensor import _raw
    return _raw(t)


class SparseTensorBLAS:
    """COO kernels (S/tensor/SparseTensorBLAS.scala:26-460). On the GPU the products run as torch sparse ops (cuSPARSE);
    the forms below keep the reference's alpha / beta / transpose conventions."""

    @staticmethod
    def vdot(vec1, vec2: SparseTensor) -> float:
        """dense 1-D . sparse 1-D"""
        d = _dense(vec1).reshape(-1)
        idx = vec2.sp.indices()[0]
        return float((d[idx] * vec2.sp.values().to(d.dtype)).sum())

    @staticmethod
    def coomv(alpha, A: SparseTensor, x, beta, y):
        """y = alpha * A x + beta * y (A sparse [m,n], x dense [n], y dense [m], updated in place)."""
        yd, xd = _dense(y), _dense(x)
        prod = torch.sparse.mm(A.sp.to(xd.dtype), xd.reshape(-1, 1)).reshape(-1)
        yd.mul_(beta).add_(prod, alpha=alpha)
        return y

    @staticmethod
    def coomm(alpha, A, B, beta, C):
        """C = alpha * A B + beta * C where exactly one of A, B is a SparseTensor (2-D), C dense, updated in place."""
        Cd = _dense(C)
        if isinstance(A, SparseTensor) and not isinstance(B, SparseTensor):
            Bd = _dense(B)
            prod = torch.sparse.mm(A.sp.to(Bd.dtype), Bd)
        elif isinstance(B, SparseTensor) and not isinstance(A, SparseTensor):
            Ad = _dense(A)
            prod = torch.sparse.mm(B.sp.to(Ad.dtype).t(), Ad.t()).t()      # (B^T A^T)^T
        else:
            raise ValueError("coomm: exactly one operand must be sparse")
        Cd.mul_(beta).add_(prod, alpha=alpha)
        return C


class SparseTensorMath:
    """``addmv`` / ``addmm`` / ``vdot`` entry points DenseTensor math dispatches to when an operand is sparse
    (S/tensor/SparseTensorMath.scala:23-95)."""

    vdot = staticmethod(SparseTensorBLAS.vdot)

    @staticmethod
    def addmv(r, beta, t, alpha, mat: SparseTensor, vec):
        """r = beta * t + alpha * mat vec"""
        rd, td = _dense(r), _dense(t)
        if rd.data_ptr() != td.data_ptr():
            rd.resize_(td.shape).copy_(td)
        return SparseTensorBLAS.coomv(alpha, mat, vec, beta, r)

    @staticmethod
    def addmm(r, beta, t, alpha, mat1, mat2):
        """r = beta * t + alpha * mat1 mat2 (one of the matrices sparse)"""
        rd, td = _dense(r), _dense(t)
        if rd.data_ptr() != td.data_ptr():
            rd.resize_(td.shape).copy_(td)
        return SparseTensorBLAS.coomm(alpha, mat1, mat2, beta, r)
