"""Loss functions (reference: S/nn/abstractnn/AbstractCriterion.scala and the 37 criterion files under S/nn/;
Python surface P/nn/criterion.py).

``forward(input, target)`` returns the loss as a Python float; ``backward(input, target)`` returns d(loss)/d(input).
Training loops call ``_call`` to obtain a differentiable scalar instead. Class targets are **1-based**.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

from ..utils.table import Table, flatten_activity, map_activity, to_table
from .module import _as_activity, _to_numpy


class Criterion:
    def __init__(self):
        self.output = 0.0
        self.gradInput = None
        self._graph = None

    # subclasses implement this with differentiable ops and return a 0-dim tensor
    def updateOutput(self, input, target):  # pragma: no cover
        raise NotImplementedError

    def _call(self, input, target):
        return self.updateOutput(input, target)

    def forward(self, input, target):
        x = _as_activity(input)
        dev = flatten_activity(x)[0].device
        t = _as_activity(target, dev)

        def leaf(v):
            v = v.detach()
            if v.is_floating_point():
                v.requires_grad_(True)
            return v

        xin = map_activity(leaf, x)
        with torch.enable_grad():
            loss = self._call(xin, t)
        self._graph = (xin, loss)
        self.output = float(loss.detach())
        return self.output

    def backward(self, input, target):
        want_numpy = isinstance(input, np.ndarray)
        if self._graph is None:
            self.forward(input, target)
        xin, loss = self._graph
        leaves = [v for v in flatten_activity(xin) if v.requires_grad]
        grads = torch.autograd.grad(loss, leaves, allow_unused=True)
        it = iter(grads)
        self.gradInput = map_activity(
            lambda v: (lambda g: g if g is not None else torch.zeros_like(v))(next(it)) if v.requires_grad
            else torch.zeros_like(v, dtype=torch.float32), xin)
        self._graph = None
        return _to_numpy(self.gradInput) if want_numpy else self.gradInput

    updateGradInput = backward

    def cloneCriterion(self):
        import copy

        self._graph = None
        return copy.deepcopy(self)

    def __call__(self, input, target):
        return self.forward(input, target)


AbstractCriterion = Criterion
TensorCriterion = Criterion


def _class_index(target: torch.Tensor) -> torch.Tensor:
    return target.reshape(-1).long() - 1


def fuse_final_logsoftmax(model, criterion, on: bool = True) -> bool:
    """Mark (or unmark) the model's final LogSoftMax so that, while training, a plain ClassNLLCriterion consumes the
    logits directly through the fused cross-entropy kernels. Returns whether the pattern matched. Only the LAST module
    of a (nested) Sequential qualifies: its output goes to the criterion and nowhere else."""
    from .layers_core import LogSoftMax
    from .module import Sequential

    m = model
    while type(m) is Sequential and m.modules:
        m = m.modules[-1]
    ok = (isinstance(m, LogSoftMax) and type(criterion) is ClassNLLCriterion and criterion.logProbAsInput)
    if isinstance(m, LogSoftMax):
        if on and ok:
            m.__dict__["_defer_to_criterion"] = True
        else:
            m.__dict__.pop("_defer_to_criterion", None)
    return bool(ok and on)


class ClassNLLCriterion(Criterion):
    """Negative log-likelihood over log-probabilities (or probabilities when ``logProbAsInput`` is false, clipped to
    [1e-8, 1-1e-8]); 1-based targets; rows whose target equals ``paddingValue`` are ignored; optional class weights;
    ``sizeAverage`` divides by the total weight (S/nn/ClassNLLCriterion.scala:50-242)."""

    def __init__(self, weights=None, size_average: bool = True, logProbAsInput: bool = True, paddingValue: int = -1):
        super().__init__()
        self.weights = None if weights is None else torch.as_tensor(np.asarray(weights), dtype=torch.float32)
        self.size_average, self.logProbAsInput, self.paddingValue = size_average, logProbAsInput, paddingValue

    def updateOutput(self, input, target):
        if getattr(input, "_bdl_logits", False) and self.logProbAsInput:
            # the model's final LogSoftMax deferred to us (see LogSoftMax): fused log-softmax + NLL on the logits
            from ..ops.functions import XentFn

            if self.weights is not None and (self.__dict__.get("_w_dev") is None or self._w_dev.device != input.device):
                self._w_dev = self.weights.to(input.device).contiguous()
            return XentFn.apply(input, target, self.__dict__.get("_w_dev"), self.size_average, self.paddingValue)
        x = input if input.dim() > 1 else input.unsqueeze(0)
        x = x.float()
        if not self.logProbAsInput:
            x = torch.log(torch.clamp(x, 1e-8, 1 - 1e-8))
        t = target.reshape(-1)
        valid = t != self.paddingValue
        idx = torch.where(valid, t.long() - 1, torch.zeros_like(t, dtype=torch.long))
        if not x.is_cuda and bool(((idx < 0) | (idx >= x.shape[1])).any()):   # no host sync on the GPU path
            raise ValueError(f"curTarget out of range 1 to {x.shape[1]}")
        w = torch.ones(x.shape[1], device=x.device) if self.weights is None else self.weights.to(x.device)
        wt = w[idx] * valid.to(x.dtype)
        picked = x.gather(1, idx.unsqueeze(1)).squeeze(1)
        total = -(picked * wt).sum()
        if self.size_average:
            denom = wt.sum()
            total = total / torch.where(denom == 0, torch.ones_like(denom), denom)
        return total


class CrossEntropyCriterion(Criterion):
    """LogSoftMax followed by ClassNLL (S/nn/CrossEntropyCriterion.scala:35-53); on CUDA the two are one fused kernel
    pair (csrc/loss_softmax.cu)."""

    def __init__(self, weights=None, size_average: bool = True):
        super().__init__()
        self.weights = None if weights is None else torch.as_tensor(np.asarray(weights), dtype=torch.float32)
        self.size_average = size_average
        self._nll = ClassNLLCriterion(weights, size_average)
        self._w_dev = None

    def updateOutput(self, input, target):
        from .layers_core import native

        x = input if input.dim() > 1 else input.unsqueeze(0)
        if native(x) and x.dtype in (torch.bfloat16, torch.float32):
            from ..ops.functions import XentFn

            if self.weights is not None and (self._w_dev is None or self._w_dev.device != x.device):
                self._w_dev = self.weights.to(x.device).contiguous()
            return XentFn.apply(x, target, self._w_dev, self.size_average, -1)
        return self._nll.updateOutput(F.log_softmax(x.float(), dim=1), target)


class MSECriterion(Criterion):
    def __init__(self, size_average: bool = True):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target):
        # the reference compares the two buffers element by element and only requires equal element counts
        # (S/nn/MSECriterion.scala:38-45), e.g. a [N, 784] reconstruction against [N, 28, 28] images
        if target.shape != input.shape and target.numel() == input.numel():
            target = target.reshape(input.shape)
        d = (input.float() - target.float()) ** 2
        return d.mean() if self.size_average else d.sum()


class AbsCriterion(Criterion):
    def __init__(self, size_average: bool = True):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target):
        d = (input.float() - target.float()).abs()
        return d.mean() if self.size_average else d.sum()


class BCECriterion(Criterion):
    """Binary cross entropy with eps=1e-12 inside the logs and optional per-element weights (S/nn/BCECriterion.scala)."""

    def __init__(self, weights=None, size_average: bool = True):
        super().__init__()
        self.weights = None if weights is None else torch.as_tensor(np.asarray(weights), dtype=torch.float32)
        self.size_average = size_average

    def updateOutput(self, input, target):
        eps = 1e-12
        x, t = input.float(), target.float().reshape(input.shape)
        l = -(t * torch.log(x + eps) + (1 - t) * torch.log(1 - x + eps))
        if self.weights is not None:
            l = l * self.weights.to(x.device).expand_as(l)
        return l.mean() if self.size_average else l.sum()


class SmoothL1Criterion(Criterion):
    def __init__(self, size_average: bool = True):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target):
        return F.smooth_l1_loss(input.float(), target.float().reshape(input.shape),
                                reduction="mean" if self.size_average else "sum")


class SmoothL1CriterionWithWeights(Criterion):
    """Fast-RCNN box loss: input table target = (bbox_target, inside_w, outside_w); sigma-scaled smooth-L1, sum divided
    by ``num`` (S/nn/SmoothL1CriterionWithWeights.scala)."""

    def __init__(self, sigma: float, num: int = 0):
        super().__init__()
        self.sigma2, self.num = sigma * sigma, num

    def updateOutput(self, input, target):
        ts = to_table(target).to_list()
        bt = ts[0].float().reshape(input.shape)
        d = input.float() - bt
        if len(ts) > 1:
            d = d * ts[1].float().reshape(input.shape)
        ad = d.abs()
        l = torch.where(ad < 1.0 / self.sigma2, 0.5 * self.sigma2 * d * d, ad - 0.5 / self.sigma2)
        if len(ts) > 2:
            l = l * ts[2].float().reshape(input.shape)
        s = l.sum()
        return s / self.num if self.num > 0 else s


class SoftmaxWithCriterion(Criterion):
    """Caffe SoftmaxWithLoss over dim 1 of [N,C,(H,W)] with ignore label and normalize modes
    (S/nn/SoftmaxWithCriterion.scala)."""

    def __init__(self, ignore_label=None, normalize_mode: str = "VALID"):
        super().__init__()
        self.ignore_label, self.normalize_mode = ignore_label, normalize_mode.upper()

    def updateOutput(self, input, target):
        x = input.float()
        lp = F.log_softmax(x, dim=1)
        t = target.reshape(x.shape[0], *x.shape[2:]).long() - 1
        valid = torch.ones_like(t, dtype=torch.bool) if self.ignore_label is None else (t != self.ignore_label - 1)
        picked = lp.gather(1, torch.clamp(t, 0, x.shape[1] - 1).unsqueeze(1)).squeeze(1)
        loss = -(picked * valid.to(lp.dtype)).sum()
        outer, inner = x.shape[0], int(np.prod(x.shape[2:])) if x.dim() > 2 else 1
        if self.normalize_mode == "FULL":
            norm = outer * inner
        elif self.normalize_mode == "VALID":
            norm = int(valid.sum()) if self.ignore_label is not None else outer * inner
        elif self.normalize_mode == "BATCH_SIZE":
            norm = outer
        else:
            norm = 1
        return loss / max(norm, 1)


class MarginCriterion(Criterion):
    def __init__(self, margin: float = 1.0, size_average: bool = True, squared: bool = False):
        super().__init__()
        self.margin, self.size_average, self.squared = margin, size_average, squared

    def updateOutput(self, input, target):
        z = torch.clamp(self.margin - input.float() * target.float().reshape(input.shape), min=0)
        if self.squared:
            z = z * z
        return z.mean() if self.size_average else z.sum()


class MarginRankingCriterion(Criterion):
    def __init__(self, margin: float = 1.0, size_average: bool = True):
        super().__init__()
        self.margin, self.size_average = margin, size_average

    def updateOutput(self, input, target):
        a, b = to_table(input).to_list()
        y = target if isinstance(target, torch.Tensor) else to_table(target).to_list()[0]
        z = torch.clamp(-y.float().reshape(a.shape) * (a.float() - b.float()) + self.margin, min=0)
        return z.mean() if self.size_average else z.sum()


class MultiMarginCriterion(Criterion):
    def __init__(self, p: int = 1, weights=None, margin: float = 1.0, size_average: bool = True):
        super().__init__()
        self.p, self.margin, self.size_average = p, margin, size_average
        self.weights = None if weights is None else torch.as_tensor(np.asarray(weights), dtype=torch.float32)

    def updateOutput(self, input, target):
        x = input.float() if input.dim() > 1 else input.float().unsqueeze(0)
        w = None if self.weights is None else self.weights.to(x.device)
        l = F.multi_margin_loss(x, _class_index(target), p=self.p, margin=self.margin, weight=w, reduction="none")
        return l.mean() if self.size_average else l.sum()


class MultiLabelMarginCriterion(Criterion):
    def __init__(self, size_average: bool = True):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target):
        x = input.float() if input.dim() > 1 else input.float().unsqueeze(0)
        t = target.reshape(x.shape).long() - 1   # 0 terminates in the reference → -1 terminates in torch
        return F.multilabel_margin_loss(x, t, reduction="mean" if self.size_average else "sum")


class MultiLabelSoftMarginCriterion(Criterion):
    def __init__(self, weights=None, size_average: bool = True):
        super().__init__()
        self.size_average = size_average
        self.weights = None if weights is None else torch.as_tensor(np.asarray(weights), dtype=torch.float32)

    def updateOutput(self, input, target):
        x, t = input.float(), target.float().reshape(input.shape)
        l = -(t * F.logsigmoid(x) + (1 - t) * F.logsigmoid(-x))
        if self.weights is not None:
            l = l * self.weights.to(x.device)
        return l.sum() / x.numel() if self.size_average else l.sum()


class SoftMarginCriterion(Criterion):
    def __init__(self, size_average: bool = True):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target):
        l = torch.log1p(torch.exp(-input.float() * target.float().reshape(input.shape)))
        return l.mean() if self.size_average else l.sum()


class HingeEmbeddingCriterion(Criterion):
    def __init__(self, margin: float = 1.0, size_average: bool = True):
        super().__init__()
        self.margin, self.size_average = margin, size_average

    def updateOutput(self, input, target):
        x, t = input.float(), target.float().reshape(input.shape)
        l = torch.where(t > 0, x, torch.clamp(self.margin - x, min=0))
        return l.mean() if self.size_average else l.sum()


class L1HingeEmbeddingCriterion(Criterion):
    def __init__(self, margin: float = 1.0):
        super().__init__()
        self.margin = margin

    def updateOutput(self, input, target):
        a, b = to_table(input).to_list()
        d = (a.float() - b.float()).abs().sum()
        y = float(target.reshape(-1)[0]) if isinstance(target, torch.Tensor) else float(target)
        return d if y > 0 else torch.clamp(self.margin - d, min=0)


class CosineEmbeddingCriterion(Criterion):
    def __init__(self, margin: float = 0.0, size_average: bool = True):
        super().__init__()
        self.margin, self.size_average = margin, size_average

    def updateOutput(self, input, target):
        a, b = to_table(input).to_list()
        if a.dim() == 1:
            a, b = a.unsqueeze(0), b.unsqueeze(0)
        y = target if isinstance(target, torch.Tensor) else to_table(target).to_list()[0]
        cos = F.cosine_similarity(a.float(), b.float(), dim=1, eps=1e-12)
        l = torch.where(y.float().reshape(-1) > 0, 1 - cos, torch.clamp(cos - self.margin, min=0))
        return l.mean() if self.size_average else l.sum()


class CosineDistanceCriterion(Criterion):
    def __init__(self, size_average: bool = True):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target):
        x = input.float() if input.dim() > 1 else input.float().unsqueeze(0)
        t = target.float().reshape(x.shape)
        l = 1 - F.cosine_similarity(x, t, dim=1, eps=1e-12)
        return l.mean() if self.size_average else l.sum()


class CosineProximityCriterion(Criterion):
    def updateOutput(self, input, target):
        x = F.normalize(input.float(), dim=-1, eps=1e-12)
        t = F.normalize(target.float().reshape(input.shape), dim=-1, eps=1e-12)
        return -(x * t).sum(dim=-1).mean()


class DistKLDivCriterion(Criterion):
    def __init__(self, size_average: bool = True):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target):
        t = target.float().reshape(input.shape)
        l = torch.where(t > 0, t * (torch.log(torch.clamp(t, min=1e-38)) - input.float()), torch.zeros_like(t))
        return l.sum() / input.numel() if self.size_average else l.sum()


class KLDCriterion(Criterion):
    """KL(N(mean, exp(logvar)) || N(0,1)) for VAEs; input = table(mean, logvar) (S/nn/KLDCriterion.scala)."""

    def __init__(self, size_average: bool = True):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target=None):
        mean, logvar = to_table(input).to_list()
        l = 0.5 * (mean.float() ** 2 + logvar.float().exp() - 1 - logvar.float()).sum()
        return l / mean.shape[0] if self.size_average and mean.dim() > 1 else l


class KullbackLeiblerDivergenceCriterion(Criterion):
    """Keras kld: sum(y_true * log(y_true / y_pred)) with both clipped to [1e-7, 1], mean over batch."""

    def updateOutput(self, input, target):
        eps = 1e-7
        y = torch.clamp(target.float().reshape(input.shape), eps, 1.0)
        p = torch.clamp(input.float(), eps, 1.0)
        s = (y * torch.log(y / p)).sum()
        return s / (input.shape[0] if input.dim() > 1 else 1)


class GaussianCriterion(Criterion):
    """-log N(target; mean, exp(logvar)) summed; input = table(mean, logvar) (S/nn/GaussianCriterion.scala)."""

    def updateOutput(self, input, target):
        mean, logvar = to_table(input).to_list()
        t = target.float().reshape(mean.shape)
        return (0.5 * math.log(2 * math.pi) + 0.5 * logvar.float() +
                0.5 * (t - mean.float()) ** 2 / logvar.float().exp()).sum()


class PoissonCriterion(Criterion):
    def updateOutput(self, input, target):
        x, t = input.float(), target.float().reshape(input.shape)
        return (x - t * torch.log(x + 1e-7)).mean()


class DiceCoefficientCriterion(Criterion):
    def __init__(self, size_average: bool = True, epsilon: float = 1.0):
        super().__init__()
        self.size_average, self.epsilon = size_average, epsilon

    def updateOutput(self, input, target):
        x = input.float().reshape(input.shape[0], -1) if input.dim() > 1 else input.float().reshape(1, -1)
        t = target.float().reshape(x.shape)
        inter = (x * t).sum(dim=1)
        dice = 1 - (2 * inter + self.epsilon) / (x.sum(dim=1) + t.sum(dim=1) + self.epsilon)
        return dice.mean() if self.size_average else dice.sum()


class ClassSimplexCriterion(Criterion):
    """MSE against the vertices of a regular simplex embedding of the classes (S/nn/ClassSimplexCriterion.scala)."""

    def __init__(self, n_classes: int):
        super().__init__()
        self.n_classes = n_classes
        self.simplex = self._regsplex(n_classes - 1)

    @staticmethod
    def _regsplex(n: int) -> torch.Tensor:
        a = torch.zeros(n + 1, n)
        for k in range(n):
            a[k, k] = 1.0 if k == 0 else math.sqrt(max(1 - float((a[k, :k] ** 2).sum()), 0.0))
            c = (a[k, k] ** 2 - 1 - 1.0 / n) / a[k, k]
            a[k + 1:, k] = c
        return a

    def updateOutput(self, input, target):
        x = input.float() if input.dim() > 1 else input.float().unsqueeze(0)
        emb = torch.zeros(x.shape[0], self.n_classes, device=x.device)
        emb[:, :self.n_classes - 1] = self.simplex.to(x.device)[_class_index(target)]
        return ((x - emb) ** 2).mean()


class L1Cost(Criterion):
    def updateOutput(self, input, target=None):
        return input.float().abs().sum()


class MeanAbsolutePercentageCriterion(Criterion):
    def updateOutput(self, input, target):
        t = target.float().reshape(input.shape)
        d = (t - input.float()).abs() / torch.clamp(t.abs(), 1e-7, float("inf"))
        return 100.0 * d.mean()


class MeanSquaredLogarithmicCriterion(Criterion):
    def updateOutput(self, input, target):
        t = target.float().reshape(input.shape)
        a = torch.log(torch.clamp(input.float(), min=1e-7) + 1)
        b = torch.log(torch.clamp(t, min=1e-7) + 1)
        return ((a - b) ** 2).mean()


class CategoricalCrossEntropy(Criterion):
    """Keras categorical_crossentropy on probabilities with one-hot targets (S/nn/CategoricalCrossEntropy.scala)."""

    def updateOutput(self, input, target):
        x = input.float() if input.dim() > 1 else input.float().unsqueeze(0)
        t = target.float().reshape(x.shape)
        x = x / x.sum(dim=-1, keepdim=True)
        return -(t * torch.log(torch.clamp(x, 1e-7, 1 - 1e-7))).sum(dim=-1).mean()


class DotProductCriterion(Criterion):
    def __init__(self, size_average: bool = False):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target):
        s = (input.float() * target.float().reshape(input.shape)).sum()
        return s / input.shape[0] if (self.size_average and input.dim() > 1) else s


class PGCriterion(Criterion):
    """Policy-gradient loss -sum(reward * log(prob)) (S/nn/PGCriterion.scala)."""

    def __init__(self, size_average: bool = False):
        super().__init__()
        self.size_average = size_average

    def updateOutput(self, input, target):
        s = -(target.float().reshape(input.shape) * torch.log(input.float())).sum()
        return s / input.shape[0] if (self.size_average and input.dim() > 1) else s


class TransformerCriterion(Criterion):
    """Apply optional transformers to input and target before an inner criterion (S/nn/TransformerCriterion.scala)."""

    def __init__(self, criterion: Criterion, input_transformer=None, target_transformer=None):
        super().__init__()
        self.criterion, self.input_transformer, self.target_transformer = criterion, input_transformer, target_transformer

    def updateOutput(self, input, target):
        x = self.input_transformer._call(input) if self.input_transformer is not None else input
        t = self.target_transformer._call(target) if self.target_transformer is not None else target
        return self.criterion._call(x, map_activity(lambda v: v.detach(), t))


class MultiCriterion(Criterion):
    """Weighted sum of criterions on the same (input, target) (S/nn/MultiCriterion.scala)."""

    def __init__(self):
        super().__init__()
        self.criterions, self.weights = [], []

    def add(self, criterion: Criterion, weight: float = 1.0):
        self.criterions.append(criterion)
        self.weights.append(weight)
        return self

    def updateOutput(self, input, target):
        total = 0
        for c, w in zip(self.criterions, self.weights):
            total = total + w * c._call(input, target)
        return total


class ParallelCriterion(Criterion):
    """i-th criterion on i-th (input, target); ``repeat_target`` shares one target (S/nn/ParallelCriterion.scala)."""

    def __init__(self, repeat_target: bool = False):
        super().__init__()
        self.repeat_target = repeat_target
        self.criterions, self.weights = [], []

    def add(self, criterion: Criterion, weight: float = 1.0):
        self.criterions.append(criterion)
        self.weights.append(weight)
        return self

    def updateOutput(self, input, target):
        xs = to_table(input).to_list()
        ts = [target] * len(xs) if self.repeat_target else to_table(target).to_list()
        total = 0
        for c, w, x, t in zip(self.criterions, self.weights, xs, ts):
            total = total + w * c._call(x, t)
        return total


class TimeDistributedCriterion(Criterion):
    """Apply a criterion at every step of dim ``dimension`` (1-based, default 2) and sum (or average over steps)
    (S/nn/TimeDistributedCriterion.scala)."""

    def __init__(self, criterion: Criterion, size_average: bool = False, dimension: int = 2):
        super().__init__()
        self.criterion, self.size_average, self.dimension = criterion, size_average, dimension

    def updateOutput(self, input, target):
        d = self.dimension - 1
        n = input.shape[d]
        total = 0
        for i in range(n):
            total = total + self.criterion._call(input.select(d, i), target.select(d, i))
        return total / n if self.size_average else total


class TimeDistributedMaskCriterion(Criterion):
    """Per-step criterion averaged over the non-padding target entries (S/nn/TimeDistributedMaskCriterion.scala)."""

    def __init__(self, criterion: Criterion, padding_value: int = 0):
        super().__init__()
        self.criterion, self.padding_value = criterion, padding_value

    def updateOutput(self, input, target):
        x = input.reshape(-1, input.shape[-1])
        t = target.reshape(-1)
        mask = t != self.padding_value
        crit = self.criterion
        old = getattr(crit, "size_average", None)
        if old is not None:
            crit.size_average = False
        pv = getattr(crit, "paddingValue", None)
        if pv is not None:
            crit.paddingValue = self.padding_value
        try:
            total = crit._call(x, t)
        finally:
            if old is not None:
                crit.size_average = old
            if pv is not None:
                crit.paddingValue = pv
        return total / torch.clamp(mask.sum().to(total.dtype), min=1)


class NormMode:
    """How SoftmaxWithCriterion normalises the summed loss (SoftmaxWithCriterion.scala:152-155)."""
    FULL, VALID, BATCH_SIZE, NONE = "FULL", "VALID", "BATCH_SIZE", "NONE"
