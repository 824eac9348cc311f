"""Flat parameter storage: every weight a view into ONE fp32 buffer, every gradient a view into ONE fp32 buffer,
and (on GPU) every compute weight a view into ONE bf16 buffer.

Reference: S/nn/Module.scala:76-107 (``Module.flatten``) — the layout ``AllReduceParameter`` slices
(S/parameters/AllReduceParameter.scala:103-104). Partition r of N owns
``[r*(P//N) + min(r, P%N), +P//N + (r < P%N))``.
"""
from __future__ import annotations

import torch


class FlatParams:
    def __init__(self, weights: torch.Tensor, grads: torch.Tensor, entries):
        self.weights = weights          # fp32 [P]
        self.grads = grads              # fp32 [P]
        self.entries = entries          # [(module, name, offset, numel)]
        self.compute: torch.Tensor | None = None   # bf16 [P] on GPU
        self.dirty = True               # compute copy stale w.r.t. weights

    @property
    def numel(self) -> int:
        return self.weights.numel()

    def refresh_compute(self, force: bool = False):
        if self.compute is None or not (self.dirty or force):
            return
        from ..ops import elementwise as ew

        ew.cast_f32_to_bf16(self.weights, self.compute)
        self.dirty = False

    def offsets_of(self, module) -> tuple[int, int]:
        """(offset, length) of the contiguous range holding all parameters of ``module`` and its children."""
        ids = {id(m) for m in module.all_modules()}
        rng = [(o, n) for m, _, o, n in self.entries if id(m) in ids]
        if not rng:
            return (0, 0)
        start = min(o for o, _ in rng)
        end = max(o + n for o, n in rng)
        return start, end - start


    def module_ranges(self):
        """[(leaf module, offset, padded length)] in flat (= forward registration) order, one entry per module."""
        out: list = []
        for m, _, o, n in self.entries:
            end = o + _align(n)
            if out and out[-1][0] is m:
                out[-1] = (m, out[-1][1], end - out[-1][1])
            else:
                out.append((m, o, end - o))
        if out:                          # the last range stops at the end of the buffer
            m, o, _ = out[-1]
            out[-1] = (m, o, self.weights.numel() - o)
        return out


def partition_range(total: int, parts: int, rank: int) -> tuple[int, int]:
    base, extra = divmod(total, parts)
    start = rank * base + min(rank, extra)
    return start, base + (1 if rank < extra else 0)


def _physical_view(flat: torch.Tensor, offset: int, like: torch.Tensor) -> torch.Tensor:
    """View of ``flat[offset:offset+numel]`` with the logical shape of ``like`` and its (dense) memory order."""
    n = like.numel()
    seg = flat[offset:offset + n]
    if like.dim() == 4 and like.is_contiguous(memory_format=torch.channels_last) and not like.is_contiguous():
        o, i, h, w = like.shape
        return seg.view(o, h, w, i).permute(0, 3, 1, 2)
    return seg.view(like.shape)


def _align(n: int, a: int = 8) -> int:
    return (n + a - 1) // a * a


def flatten_parameters(root):
    fl = getattr(root, "_flat", None)
    if fl is not None:
        return fl.weights, fl.grads
    leaves = root.leaf_parameters()
    if not leaves:
        root._flat = None
        z = torch.zeros(0)
        return z, z.clone()
    dev = leaves[0][2].device
    # Offsets are padded to 8 elements so every bf16 view is 16-byte aligned (TMA / vector loads).
    offsets, total = [], 0
    for _, _, p in leaves:
        offsets.append(total)
        total += _align(p.numel())
    weights = torch.zeros(total, dtype=torch.float32, device=dev)
    grads = torch.zeros(total, dtype=torch.float32, device=dev)
    compute = torch.zeros(total, dtype=torch.bfloat16, device=dev) if dev.type == "cuda" else None
    entries = []
    with torch.no_grad():
        for (m, n, p), off in zip(leaves, offsets):
            wv = _physical_view(weights, off, p)
            gv = _physical_view(grads, off, p)
            wv.copy_(p.detach())
            if p.grad is not None:
                gv.copy_(p.grad)
            req = p.requires_grad
            q = wv.detach()
            q.requires_grad_(req)
            q.grad = gv
            for mm in root.all_modules():   # shared parameters: re-point every holder
                for nn_, pp in list(mm._parameters.items()):
                    if pp is p:
                        mm._parameters[nn_] = q
                        if compute is not None:
                            mm._compute[nn_] = _physical_view(compute, off, p)
            entries.append((m, n, off, p.numel()))
    fl = FlatParams(weights, grads, entries)
    fl.compute = compute
    fl.dirty = True
    root._flat = fl
    return weights, grads


def unflatten_parameters(root):
    """Give every parameter its own storage again (needed before deepcopy / device moves)."""
    fl = getattr(root, "_flat", None)
    if fl is None:
        return
    seen = {}
    for m in root.all_modules():
        for n, p in list(m._parameters.items()):
            if id(p) in seen:
                m._parameters[n] = seen[id(p)]
                continue
            mf = torch.channels_last if (p.dim() == 4 and not p.is_contiguous()) else torch.contiguous_format
            q = p.detach().clone(memory_format=mf)
            q.requires_grad_(p.requires_grad)
            q.grad = p.grad.detach().clone(memory_format=mf) if p.grad is not None else torch.zeros_like(q)
            seen[id(p)] = q
            m._parameters[n] = q
        m._compute = {}
    root._flat = None
