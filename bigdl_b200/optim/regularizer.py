"""L1/L2 regularizers applied inside accGradParameters (S/optim/Regularizer.scala:30-193)."""
from __future__ import annotations

import torch


class Regularizer:
    is_regularized = True

    def accRegularization(self, parameter: torch.Tensor, gradParameter: torch.Tensor, scale: float):  # pragma: no cover
        raise NotImplementedError


class L1L2Regularizer(Regularizer):
    def __init__(self, l1: float, l2: float):
        self.l1, self.l2 = l1, l2

    def accRegularization(self, parameter, gradParameter, scale: float = 1.0):
        if gradParameter is None or not self.is_regularized:
            return
        with torch.no_grad():
            if self.l1 != 0:
                gradParameter.add_(torch.sign(parameter.detach()), alpha=self.l1 * scale)
            if self.l2 != 0:
                gradParameter.add_(parameter.detach(), alpha=self.l2 * scale)

    def enable(self):
        self.is_regularized = True
        return self

    def disable(self):
        self.is_regularized = False
        return self


class L1Regularizer(L1L2Regularizer):
    def __init__(self, l1: float):
        super().__init__(l1, 0.0)


class L2Regularizer(L1L2Regularizer):
    def __init__(self, l2: float):
        super().__init__(0.0, l2)
