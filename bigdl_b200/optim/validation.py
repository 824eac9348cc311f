"""Validation methods and their monoid results (S/optim/ValidationMethod.scala:1-1117).

Every ``ValidationResult`` supports ``+`` so partial results reduce across batches and, via
``all_reduce_result``, across ranks (the role of the Spark ``reduce`` in S/optim/AbstractOptimizer.scala:118-172).
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

from ..utils.table import Table, to_table


class ValidationResult:
    def result(self):  # pragma: no cover
        raise NotImplementedError

    def __add__(self, other):  # pragma: no cover
        raise NotImplementedError

    def as_vector(self) -> list[float]:  # pragma: no cover
        raise NotImplementedError

    @classmethod
    def from_vector(cls, v):  # pragma: no cover
        raise NotImplementedError


class AccuracyResult(ValidationResult):
    def __init__(self, correct: int, count: int):
        self.correct, self.count = int(correct), int(count)

    def result(self):
        return (self.correct / self.count if self.count else 0.0), self.count

    def __add__(self, other):
        return AccuracyResult(self.correct + other.correct, self.count + other.count)

    def as_vector(self):
        return [float(self.correct), float(self.count)]

    @classmethod
    def from_vector(cls, v):
        return cls(int(round(v[0])), int(round(v[1])))

    def __eq__(self, o):
        return isinstance(o, AccuracyResult) and (self.correct, self.count) == (o.correct, o.count)

    def __repr__(self):
        return f"Accuracy(correct: {self.correct}, count: {self.count}, accuracy: {self.result()[0]})"


class LossResult(ValidationResult):
    def __init__(self, loss: float, count: int):
        self.loss, self.count = float(loss), int(count)

    def result(self):
        return (self.loss / self.count if self.count else 0.0), self.count

    def __add__(self, other):
        return LossResult(self.loss + other.loss, self.count + other.count)

    def as_vector(self):
        return [self.loss, float(self.count)]

    @classmethod
    def from_vector(cls, v):
        return cls(v[0], int(round(v[1])))

    def __repr__(self):
        return f"(Loss: {self.loss}, count: {self.count}, Average Loss: {self.result()[0]})"


class ContiguousResult(ValidationResult):
    """Sum of a per-sample metric (HitRatio / NDCG / MAE)."""

    def __init__(self, total: float, count: int, name: str):
        self.total, self.count, self.name = float(total), int(count), name

    def result(self):
        return (self.total / self.count if self.count else 0.0), self.count

    def __add__(self, other):
        return ContiguousResult(self.total + other.total, self.count + other.count, self.name)

    def as_vector(self):
        return [self.total, float(self.count)]

    def __repr__(self):
        return f"{self.name}: {self.result()[0]} (count {self.count})"


class ValidationMethod:
    name = "ValidationMethod"

    def __call__(self, output, target) -> ValidationResult:  # pragma: no cover
        raise NotImplementedError

    def clone(self):
        import copy

        return copy.deepcopy(self)

    def __repr__(self):
        return self.name


def _logits_2d(output):
    if isinstance(output, Table):
        output = output.to_list()[0]
    return output if output.dim() > 1 else output.unsqueeze(0)


class Top1Accuracy(ValidationMethod):
    """Arg-max equals the (1-based) target; a single-column output is thresholded at 0.5 (ValidationMethod.scala:173)."""

    name = "Top1Accuracy"

    def __init__(self, zero_based_label: bool = False):
        self.zero_based_label = zero_based_label

    def __call__(self, output, target):
        o = _logits_2d(output)
        t = target.reshape(-1)
        if o.shape[1] == 1:
            pred = (o.reshape(-1) >= 0.5).long()
            correct = (pred == t.long()).sum()
            return AccuracyResult(int(correct), t.numel())
        if o.is_cuda and not self.zero_based_label:
            from ..nn.layers_core import native

            if native(o) and o.dtype in (torch.float32, torch.bfloat16):
                from ..ops.functions import topk_correct

                c = topk_correct(o, t)
                return AccuracyResult(int(c[0]), t.numel())
        pred = o.float().argmax(dim=1) + (0 if self.zero_based_label else 1)
        return AccuracyResult(int((pred == t.long()).sum()), t.numel())


class Top5Accuracy(ValidationMethod):
    name = "Top5Accuracy"

    def __init__(self, zero_based_label: bool = False):
        self.zero_based_label = zero_based_label

    def __call__(self, output, target):
        o = _logits_2d(output)
        t = target.reshape(-1)
        if o.is_cuda and not self.zero_based_label:
            from ..nn.layers_core import native

            if native(o) and o.dtype in (torch.float32, torch.bfloat16):
                from ..ops.functions import topk_correct

                c = topk_correct(o, t)
                return AccuracyResult(int(c[1]), t.numel())
        k = min(5, o.shape[1])
        top = o.float().topk(k, dim=1).indices + (0 if self.zero_based_label else 1)
        return AccuracyResult(int((top == t.long().unsqueeze(1)).any(dim=1).sum()), t.numel())


class Loss(ValidationMethod):
    name = "Loss"

    def __init__(self, criterion=None):
        if criterion is None:
            from ..nn.criterion import ClassNLLCriterion

            criterion = ClassNLLCriterion()
        self.criterion = criterion

    def __call__(self, output, target):
        with torch.no_grad():
            l = float(self.criterion._call(output, target))
        n = output.shape[0] if isinstance(output, torch.Tensor) and output.dim() > 1 else 1
        return LossResult(l * n, n)


class MAE(ValidationMethod):
    name = "MAE"

    def __call__(self, output, target):
        o = output.float()
        if o.dim() > 1 and o.shape[-1] > 1 and target.numel() == o.shape[0]:
            o = (o.argmax(dim=-1) + 1).float()
        d = (o.reshape(-1) - target.float().reshape(-1)).abs().mean()
        n = output.shape[0] if output.dim() > 1 else 1
        return LossResult(float(d) * n, n)


class TreeNNAccuracy(ValidationMethod):
    """Accuracy on the root node of a tree output [batch, nodes, classes] (ValidationMethod.scala:121)."""

    name = "TreeNNAccuracy"

    def __call__(self, output, target):
        o = output[:, 0, :] if output.dim() == 3 else output[0:1, :]
        t = target[:, 0] if target.dim() == 2 else target.reshape(-1)[0:1]
        if o.shape[1] == 1:
            pred = (o.reshape(-1) >= 0.5).long()
        else:
            pred = o.argmax(dim=1) + 1
        return AccuracyResult(int((pred == t.long()).sum()), t.numel())


class HitRatio(ValidationMethod):
    """Recommendation hit-ratio@k: each row group holds 1 positive + neg_num negatives (ValidationMethod.scala:882)."""

    def __init__(self, k: int = 10, neg_num: int = 100):
        self.k, self.neg_num = k, neg_num
        self.name = f"HitRatio@{k}"

    def __call__(self, output, target):
        o = output.reshape(-1).float()
        t = target.reshape(-1).float()
        group = self.neg_num + 1
        hits, n = 0.0, 0
        for s in range(0, o.numel(), group):
            oo, tt = o[s:s + group], t[s:s + group]
            pos = int(tt.argmax())
            rank = int((oo > oo[pos]).sum()) + 1
            hits += 1.0 if rank <= self.k else 0.0
            n += 1
        return ContiguousResult(hits, n, self.name)


class NDCG(ValidationMethod):
    def __init__(self, k: int = 10, neg_num: int = 100):
        self.k, self.neg_num = k, neg_num
        self.name = f"NDCG@{k}"

    def __call__(self, output, target):
        o = output.reshape(-1).float()
        t = target.reshape(-1).float()
        group = self.neg_num + 1
        tot, n = 0.0, 0
        for s in range(0, o.numel(), group):
            oo, tt = o[s:s + group], t[s:s + group]
            pos = int(tt.argmax())
            rank = int((oo > oo[pos]).sum()) + 1
            tot += math.log(2) / math.log(rank + 1) if rank <= self.k else 0.0
            n += 1
        return ContiguousResult(tot, n, self.name)


class MAPResult(ValidationResult):
    """Per-class (confidence, is_positive) lists + positive counts; merged by concatenation."""

    def __init__(self, n_class: int, k: int, records, pos_cnt):
        self.n_class, self.k = n_class, k
        self.records = records          # list per class of (score, label)
        self.pos_cnt = pos_cnt

    def _ap(self, c):
        recs = sorted(self.records[c], key=lambda r: -r[0])
        if self.k > 0:
            recs = recs[:self.k]
        tp, ap = 0, 0.0
        for i, (_, lab) in enumerate(recs, 1):
            if lab:
                tp += 1
                ap += tp / i
        return ap / self.pos_cnt[c] if self.pos_cnt[c] else 0.0

    def result(self):
        aps = [self._ap(c) for c in range(self.n_class)]
        valid = [a for a, p in zip(aps, self.pos_cnt) if p > 0]
        return (sum(valid) / len(valid) if valid else 0.0), sum(self.pos_cnt)

    def __add__(self, other):
        return MAPResult(self.n_class, self.k, [a + b for a, b in zip(self.records, other.records)],
                         [a + b for a, b in zip(self.pos_cnt, other.pos_cnt)])

    def as_vector(self):
        raise NotImplementedError("MAPResult is gathered with all_gather_object")

    def __repr__(self):
        return f"MeanAveragePrecision: {self.result()[0]}"


class MeanAveragePrecision(ValidationMethod):
    """Classification mAP@k over ``classes`` classes (ValidationMethod.scala:230)."""

    def __init__(self, k: int, classes: int):
        self.k, self.classes = k, classes
        self.name = f"MAP@{k}"

    def __call__(self, output, target):
        o = _logits_2d(output).float().cpu()
        t = target.reshape(-1).long().cpu() - 1
        recs = [[] for _ in range(self.classes)]
        pos = [0] * self.classes
        for i in range(o.shape[0]):
            for c in range(self.classes):
                recs[c].append((float(o[i, c]), bool(t[i] == c)))
            if 0 <= int(t[i]) < self.classes:
                pos[int(t[i])] += 1
        return MAPResult(self.classes, self.k, recs, pos)


class PrecisionRecallAUC(ValidationMethod):
    name = "PrecisionRecallAUC"

    class _Res(ValidationResult):
        def __init__(self, scores, labels):
            self.scores, self.labels = scores, labels

        def result(self):
            s = np.asarray(self.scores)
            l = np.asarray(self.labels)
            order = np.argsort(-s, kind="stable")
            l = l[order]
            tp = np.cumsum(l)
            fp = np.cumsum(1 - l)
            npos = max(l.sum(), 1)
            prec = np.concatenate([[1.0], tp / np.maximum(tp + fp, 1)])
            rec = np.concatenate([[0.0], tp / npos])
            return float(np.trapz(prec, rec)), len(s)

        def __add__(self, o):
            return PrecisionRecallAUC._Res(self.scores + o.scores, self.labels + o.labels)

        def __repr__(self):
            return f"Precision Recall AUC is {self.result()[0]}"

    def __call__(self, output, target):
        return self._Res(output.reshape(-1).float().cpu().tolist(), target.reshape(-1).float().cpu().tolist())


def all_reduce_result(res: ValidationResult) -> ValidationResult:
    """Sum a result over all ranks (K18 in SURVEY.md §2.M)."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return res
    try:
        vec = res.as_vector()
    except NotImplementedError:
        objs = [None] * dist.get_world_size()
        dist.all_gather_object(objs, res)
        out = objs[0]
        for o in objs[1:]:
            out = out + o
        return out
    t = torch.tensor(vec, dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.all_reduce(t)
    v = t.cpu().tolist()
    if isinstance(res, ContiguousResult):
        return ContiguousResult(v[0], int(round(v[1])), res.name)
    return type(res).from_vector(v)
