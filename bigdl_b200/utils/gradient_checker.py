"""Finite-difference gradient checker (reference: the ``GradientChecker`` used by the layer specs,
spark/dl/src/test/scala/.../nn/GradientChecker.scala: perturb every input element by ±eps, compare the numerical
derivative of sum(output * gradOutput) with ``backward``; ``checkWeight`` does the same for the parameters)."""
from __future__ import annotations

import torch

from .table import Table


def _flat(act):
    if isinstance(act, torch.Tensor):
        return [act]
    if isinstance(act, Table):
        out = []
        for v in act.to_list():
            out.extend(_flat(v))
        return out
    return []


class GradientChecker:
    def __init__(self, step_size: float = 1e-3, threshold: float = 1e-2):
        self.eps, self.threshold = step_size, threshold

    def _objective(self, layer, x, gout):
        ys = _flat(layer.forward(x))
        return sum(float((y.double() * g.double()).sum()) for y, g in zip(ys, gout))

    def _compare(self, numeric, analytic) -> bool:
        scale = max(1.0, float(analytic.abs().max()), float(numeric.abs().max()))
        return float((numeric - analytic).abs().max()) <= self.threshold * scale

    def checkLayer(self, layer, input, max_elements: int = 64) -> bool:
        """d/d(input): numeric vs ``layer.backward`` on (at most ``max_elements`` of) every floating input tensor."""
        layer.evaluate() if not getattr(layer, "_check_in_training", False) else layer.training()
        xs = _flat(input)
        ys = _flat(layer.forward(input))
        g = torch.Generator().manual_seed(0)
        gout = [torch.randn(y.shape, generator=g, dtype=y.dtype) for y in ys]
        gin = layer.backward(input, gout[0] if len(gout) == 1 and not isinstance(layer.output, Table) else Table(*gout))
        gins = _flat(gin)
        ok = True
        for x, gi in zip(xs, gins):
            if not x.is_floating_point():
                continue
            flat = x.reshape(-1)
            idx = torch.randperm(flat.numel(), generator=g)[:max_elements]
            num = torch.zeros(len(idx), dtype=torch.float64)
            for k, i in enumerate(idx):
                old = float(flat[i])
                flat[i] = old + self.eps
                up = self._objective(layer, input, gout)
                flat[i] = old - self.eps
                dn = self._objective(layer, input, gout)
                flat[i] = old
                num[k] = (up - dn) / (2 * self.eps)
            ok = ok and self._compare(num, gi.reshape(-1)[idx].double())
        return ok

    def checkWeight(self, layer, input, max_elements: int = 48) -> bool:
        """d/d(parameters): numeric vs the accumulated gradWeight / gradBias after one backward."""
        layer.evaluate() if not getattr(layer, "_check_in_training", False) else layer.training()
        ys = _flat(layer.forward(input))
        g = torch.Generator().manual_seed(1)
        gout = [torch.randn(y.shape, generator=g, dtype=y.dtype) for y in ys]
        layer.zeroGradParameters()
        layer.forward(input)
        layer.backward(input, gout[0] if len(gout) == 1 and not isinstance(layer.output, Table) else Table(*gout))
        ws, gs = layer.parameters()
        ok = True
        for w, gw in zip(ws, gs):
            # index logically (unravel): parameters may be stored in a permuted physical order (channels-last conv
            # weights), where a flat reshape would silently copy
            idx = torch.randperm(w.numel(), generator=g)[:max_elements]
            num = torch.zeros(len(idx), dtype=torch.float64)
            ana = torch.zeros(len(idx), dtype=torch.float64)
            for k, i in enumerate(idx):
                pos = tuple(int(v) for v in torch.unravel_index(i, w.shape))
                old = float(w.data[pos])
                ana[k] = float(gw[pos])
                w.data[pos] = old + self.eps
                if hasattr(layer, "_mark_weights_dirty"):
                    layer._mark_weights_dirty()
                up = self._objective(layer, input, gout)
                w.data[pos] = old - self.eps
                dn = self._objective(layer, input, gout)
                w.data[pos] = old
                num[k] = (up - dn) / (2 * self.eps)
            ok = ok and self._compare(num, ana)
        return ok
