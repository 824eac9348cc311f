"""Gradient post-processing hooks run between the reduce-scatter and the optimizer step
(S/parameters/ParameterOperations.scala:34-135): ``collectGlobalData`` gathers what needs a cross-rank reduction,
``processParameters`` rewrites the local gradient slice. ``LarsProcessor`` (S/optim/LarsSGD.scala:289-357) computes the
global per-layer ‖w‖, ‖g‖ and hands every LarsSGD instance its trust ratio.

On GPU the constant clip and the L2-norm clip are folded into the fused update kernel (csrc/comm_fused.cu, modes SUMSQ
and the clip fields of OptimHP); these classes are the explicit form used by the CPU/gloo backend, by
per-submodule optim methods and by user code."""
from __future__ import annotations

import math

import torch
import torch.distributed as dist


def _world():
    return dist.get_world_size() if dist.is_initialized() else 1


class ParameterProcessor:
    def collectGlobalData(self, grad_slice: torch.Tensor, state: dict):
        pass

    def processParameters(self, grad_slice: torch.Tensor, state: dict):
        pass


class ConstantClippingProcessor(ParameterProcessor):
    def __init__(self, min: float, max: float):
        self.min, self.max = float(min), float(max)

    def processParameters(self, grad_slice, state):
        grad_slice.clamp_(self.min, self.max)


class L2NormClippingProcessor(ParameterProcessor):
    def __init__(self, l2_norm_threshold: float):
        self.threshold = float(l2_norm_threshold)

    def collectGlobalData(self, grad_slice, state):
        from ..ops.elementwise import sumsq

        sq = sumsq(grad_slice.contiguous()).double().cpu() if grad_slice.is_cuda else \
            grad_slice.double().pow(2).sum().reshape(1)
        if _world() > 1:
            t = sq.to(grad_slice.device) if dist.get_backend() == "nccl" else sq
            dist.all_reduce(t)
            sq = t.cpu()
        state["l2Norm"] = math.sqrt(float(sq))

    def processParameters(self, grad_slice, state):
        nrm = state.get("l2Norm", 0.0)
        if nrm > self.threshold:
            grad_slice.mul_(self.threshold / nrm)


class LarsProcessor(ParameterProcessor):
    """``ranges``: [(offset, length, LarsSGD)] over the flat vector; ``slice_start/slice_len`` describe which part of
    the vector this rank owns. Norm partial sums are all-reduced so every rank derives the same trust ratio."""

    def __init__(self, ranges, slice_start: int = 0, slice_len: int | None = None, scale: float = 1.0):
        self.ranges, self.slice_start, self.slice_len, self.scale = ranges, slice_start, slice_len, scale

    def collectGlobalData(self, grad_slice, state, weight_slice: torch.Tensor | None = None):
        assert weight_slice is not None
        end = self.slice_start + (self.slice_len if self.slice_len is not None else grad_slice.numel())
        sums = torch.zeros(len(self.ranges), 2, dtype=torch.float64)
        for i, (off, ln, _) in enumerate(self.ranges):
            a, b = max(off, self.slice_start), min(off + ln, end)
            if a < b:
                g = grad_slice[a - self.slice_start:b - self.slice_start].double()
                w = weight_slice[a - self.slice_start:b - self.slice_start].double()
                sums[i, 0], sums[i, 1] = float(w.pow(2).sum()), float(g.pow(2).sum())
        if _world() > 1:
            t = sums.to(grad_slice.device) if dist.get_backend() == "nccl" else sums
            dist.all_reduce(t)
            sums = t.cpu()
        state["larsNorms"] = sums

    def processParameters(self, grad_slice, state):
        sums = state["larsNorms"]
        for i, (_, _, om) in enumerate(self.ranges):
            nw, ng = math.sqrt(float(sums[i, 0])), math.sqrt(float(sums[i, 1])) * self.scale
            om.setGradientScale((ng + om.weightDecay * nw) / nw if nw != 0 else float("inf"))
