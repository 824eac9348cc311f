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


class PRAUCResult(ValidationResult):
    """(score, label) pairs; the area under the precision-recall curve is integrated by trapezoids from
    (recall 0, precision 1) over the pairs in descending score order (S/optim/PrecisionRecallAUC.scala:47-92)."""

    def __init__(self, results=None, labels=None):
        if labels is not None:                       # (scores, labels) form
            results = list(zip(results, labels))
        self.results = [(float(a), float(b)) for a, b in (results or [])]

    @property
    def scores(self):
        return [r[0] for r in self.results]

    @property
    def labels(self):
        return [r[1] for r in self.results]

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
        trapezoid = getattr(np, "trapezoid", None) or np.trapz
        return float(trapezoid(prec, rec)), len(s)

    def __add__(self, o):
        return PRAUCResult(self.results + o.results)

    def __repr__(self):
        r = self.result()
        return f"Precision Recall AUC is {r[0]} on {r[1]}"


class PrecisionRecallAUC(ValidationMethod):
    name = "PrecisionRecallAUC"
    _Res = PRAUCResult

    def __call__(self, output, target):
        if output.numel() == 0 or target.numel() == 0:
            raise ValueError("PrecisionRecallAUC: output and target must not be empty")
        if output.numel() != target.numel():
            raise ValueError(f"PrecisionRecallAUC: {output.numel()} scores but {target.numel()} labels")
        return PRAUCResult(output.reshape(-1).float().cpu().tolist(), target.reshape(-1).float().cpu().tolist())


class EvaluateMethods:
    """(correct, count) helpers (S/optim/EvaluateMethods.scala:21-95); targets are 1-based class ids."""

    @staticmethod
    def calcAccuracy(output, target):
        output, target = torch.as_tensor(output), torch.as_tensor(target)
        if output.dim() == 2:
            pred = output.argmax(1) + 1
            return int((pred == target.reshape(-1).to(pred.dtype)).sum()), output.shape[0]
        if output.dim() == 1:
            if target.numel() != 1:
                raise ValueError("a single output needs a single target")
            return int(int(output.argmax()) + 1 == int(target.reshape(-1)[0])), 1
        raise ValueError("output must have 1 or 2 dimensions")

    @staticmethod
    def calcTop5Accuracy(output, target):
        output, target = torch.as_tensor(output), torch.as_tensor(target)
        if output.dim() == 2:
            top = output.topk(min(5, output.shape[1]), 1).indices + 1
            hit = (top == target.reshape(-1, 1).to(top.dtype)).any(1)
            return int(hit.sum()), output.shape[0]
        if output.dim() == 1:
            if target.numel() != 1:
                raise ValueError("a single output needs a single target")
            top = output.topk(min(5, output.shape[0])).indices + 1
            return int(bool((top == int(target.reshape(-1)[0])).any())), 1
        raise ValueError("output must have 1 or 2 dimensions")


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


# ---------------------------------------------------------------------------------------------- detection mAP
class MAPType:
    """Which interpolation the average precision uses (ValidationMethod.scala:410-413)."""
    VOC2007, VOC2010, COCO = "voc2007", "voc2010", "coco"


MAPPascalVoc2007, MAPPascalVoc2010, MAPCOCO = MAPType.VOC2007, MAPType.VOC2010, MAPType.COCO


class MAPValidationResult(ValidationResult):
    """Per-class (confidence, hit) lists + ground-truth counts; the AP is only computed in ``result()``
    (ValidationMethod.scala:418-545). ``-1`` marks a COCO class without ground truth; such classes are left out of
    the mean."""

    def __init__(self, n_class, k, predict_for_class, gt_cnt, the_type=MAPPascalVoc2010, skip_class=-1):
        assert skip_class == -1 or 0 <= skip_class < n_class, f"Invalid skipClass {skip_class}"
        self.n_class, self.k, self.type, self.skip = n_class, k, the_type, skip_class
        self.predict, self.gt_cnt = predict_for_class, list(gt_cnt)

    def class_ap(self, c: int) -> float:
        recs = sorted(self.predict[c], key=lambda r: -r[0])
        if self.k > 0:
            recs = recs[:self.k]
        pos, tp, pnr = self.gt_cnt[c], 0, []
        for j, (_, hit) in enumerate(recs):
            if hit:
                tp += 1
                pnr.append((tp / pos, tp / (j + 1)))          # (recall, precision)

        def max_p(recall):
            return max((p for r, p in pnr if r >= recall), default=0.0)

        if self.type == MAPPascalVoc2007:
            return sum(max_p(0.1 * r) for r in range(11)) / 11
        if self.type == MAPPascalVoc2010:
            return sum(max_p(r / pos) for r in range(1, pos + 1)) / pos if pos else 0.0
        if pos == 0:
            return -1.0
        return sum(max_p(0.01 * r) for r in range(101)) / 101

    def result(self):
        aps = [self.class_ap(c) for c in range(self.n_class) if c != self.skip]
        if self.type == MAPCOCO:
            valid = [a for a in aps if a != -1.0]
            return (sum(valid) / len(valid) if valid else 0.0), 1
        return sum(aps) / (self.n_class - (0 if self.skip == -1 else 1)), 1

    def __add__(self, other):
        return MAPValidationResult(self.n_class, self.k, [a + b for a, b in zip(self.predict, other.predict)],
                                   [a + b for a, b in zip(self.gt_cnt, other.gt_cnt)], self.type, self.skip)

    def as_vector(self):
        raise NotImplementedError("detection results are gathered with all_gather_object")

    def __repr__(self):
        return f"MAPObjectDetection: {self.result()[0]}"


class MAPMultiIOUValidationResult(ValidationResult):
    """One MAPValidationResult per IoU threshold, averaged (COCO's mAP@[.5:.95]; ValidationMethod.scala:602-660)."""

    def __init__(self, impl):
        self.impl = list(impl)

    def result(self):
        return sum(r.result()[0] for r in self.impl) / len(self.impl), 1

    def __add__(self, other):
        assert len(self.impl) == len(other.impl)
        return MAPMultiIOUValidationResult([a + b for a, b in zip(self.impl, other.impl)])

    def as_vector(self):
        raise NotImplementedError("detection results are gathered with all_gather_object")

    def __repr__(self):
        return f"MAPObjectDetection(multi-IoU): {self.result()[0]}"


class GroundTruthRegion:
    """A ground-truth region and, per IoU threshold, whether a detection has claimed it
    (ValidationMethod.scala:546-573)."""

    __slots__ = ("label", "diff", "box", "rle", "occupied", "coco")

    def __init__(self, coco, n_iou, label, diff, box=None, rle=None):
        self.coco, self.label, self.diff, self.box, self.rle = coco, label, diff, box, rle
        self.occupied = [False] * n_iou

    def can_occupy(self, i):        # a COCO crowd region may absorb any number of detections
        return (self.coco and self.diff == 1) or not self.occupied[i]

    def occupy(self, i):
        self.occupied[i] = True

    def iou(self, x1, y1, x2, y2, det_rle=None):
        if self.rle is not None:
            from ..dataset.segmentation import MaskUtils

            return MaskUtils.rleIOU(det_rle, self.rle, self.diff != 0)
        gx1, gy1, gx2, gy2 = self.box
        inter = max(min(gx2, x2) - max(gx1, x1) + 1, 0) * max(min(gy2, y2) - max(gy1, y1) + 1, 0)
        det_area = (x2 - x1 + 1) * (y2 - y1 + 1)
        union = det_area if (self.coco and self.diff != 0) else det_area + (gx2 - gx1 + 1) * (gy2 - gy1 + 1) - inter
        return inter / union

    canOccupy, getIOURate = can_occupy, iou


_GroundTruth = GroundTruthRegion


class GroundTruthBBox(GroundTruthRegion):
    """Box region (ValidationMethod.scala:575-591)."""
    __slots__ = ()

    def __init__(self, isCOCO, numIOU, label, diff, xmin, ymin, xmax, ymax):
        super().__init__(isCOCO, numIOU, label, diff, (float(xmin), float(ymin), float(xmax), float(ymax)))

    xmin = property(lambda self: self.box[0])
    ymin = property(lambda self: self.box[1])
    xmax = property(lambda self: self.box[2])
    ymax = property(lambda self: self.box[3])


class GroundTruthRLE(GroundTruthRegion):
    """Mask region; always under COCO rules (ValidationMethod.scala:593-600)."""
    __slots__ = ()

    def __init__(self, numIOU, label, diff, rle):
        super().__init__(True, numIOU, label, diff, None, rle)


def _field(tab, key):
    if isinstance(tab, dict):
        return tab.get(key)
    try:
        return tab[key]
    except (KeyError, IndexError):
        return None


def _table_items(t):
    if isinstance(t, (list, tuple)):
        return list(t)
    if hasattr(t, "to_list"):
        return t.to_list()
    return [t[i] for i in range(1, len(t) + 1)]


class MeanAveragePrecisionObjectDetection(ValidationMethod):
    """Detection mAP, Pascal VOC 2007 / 2010 or COCO rules (ValidationMethod.scala:674-760, MAPUtil :286-408).

    ``output``: per image a table ``{bboxes [n,4], classes [n], scores [n] (, masks)}`` — or one 2-D tensor whose
    rows are ``[n, (label, score, x1, y1, x2, y2) * n]``; ``target``: per image ``{bboxes, classes ([n] or [2,n] with a
    "difficult" row), is_crowd (, masks)}``. Class ids start at 0. Detections are matched greedily in descending score
    order to the best still-free ground truth of their class with IoU >= threshold; a match to a difficult / crowd
    region counts neither as hit nor as miss."""

    def __init__(self, classes: int, topK: int = -1, iouThres=(0.5,), theType=MAPPascalVoc2010, skipClass: int = -1,
                 isSegmentation: bool = False):
        self.classes, self.k, self.iou = classes, topK, [float(v) for v in iouThres]
        self.type, self.skip, self.seg = theType, skipClass, isSegmentation
        self.name = "MAPObjectDetection"

    def _ground_truth(self, target):
        coco = self.type == MAPCOCO
        gt_cnt = [0] * self.classes
        images = []
        for roi in _table_items(target):
            regs = []
            bbox = _field(roi, "bboxes") if roi is not None and len(roi) > 0 else None
            if bbox is not None and len(bbox):
                bbox = torch.as_tensor(bbox).float().reshape(-1, 4)
                cls = torch.as_tensor(_field(roi, "classes")).float()
                crowd = _field(roi, "is_crowd")
                crowd = torch.zeros(bbox.shape[0]) if crowd is None else torch.as_tensor(crowd).float().reshape(-1)
                masks = _field(roi, "masks") if self.seg else None
                assert (cls.shape[-1] if cls.dim() == 2 else cls.numel()) == bbox.shape[0] == crowd.numel()
                for j in range(bbox.shape[0]):
                    label, d = (int(cls[0, j]), float(cls[1, j])) if cls.dim() == 2 else (int(cls.reshape(-1)[j]), 0.0)
                    assert 0 <= label < self.classes, f"Bad label id {label}"
                    diff = 1.0 if (float(crowd[j]) != 0 or d != 0) else 0.0
                    if masks is not None:
                        regs.append(GroundTruthRLE(len(self.iou), label, diff, masks[j]))
                    else:
                        regs.append(GroundTruthBBox(coco, len(self.iou), label, diff, *(float(v) for v in bbox[j])))
                    if diff == 0:
                        gt_cnt[label] += 1
            images.append(regs)
        return images, gt_cnt

    def _match(self, regs, label, score, x1, y1, x2, y2, mask, predict):
        assert 0 <= label < self.classes, f"Bad label id {label}"
        for i, thr in enumerate(self.iou):
            best = None
            for gt in regs:
                if gt.label != label or not gt.can_occupy(i):
                    continue
                r = gt.iou(x1, y1, x2, y2, mask)
                if r < thr:
                    continue
                if best is None or (gt.diff < best[0].diff if gt.diff != best[0].diff else r >= best[1]):
                    best = (gt, r)
            if best is not None:
                best[0].occupied[i] = True
            if best is None or best[0].diff == 0:
                predict[i][label].append((score, best is not None))

    def __call__(self, output, target):
        images, gt_cnt = self._ground_truth(target)
        predict = [[[] for _ in range(self.classes)] for _ in self.iou]
        if isinstance(output, torch.Tensor):
            assert not self.seg, "Cannot get segmentation data from tensor output for MAP"
            out = output.float().cpu()
            assert out.dim() == 2
            for img in range(out.shape[0]):
                n = int(out[img, 0])
                for b in range(n):
                    lab, sc, x1, y1, x2, y2 = (float(v) for v in out[img, 1 + 6 * b:7 + 6 * b])
                    self._match(images[img], int(lab), sc, x1, y1, x2, y2, None, predict)
        else:
            outs = _table_items(output)
            assert len(outs) == len(images), "The number of images in the output and in the target should be the same"
            for regs, det in zip(images, outs):
                if det is None or len(det) == 0:
                    continue
                bbox = torch.as_tensor(_field(det, "bboxes")).float().reshape(-1, 4)
                scores = torch.as_tensor(_field(det, "scores")).float().reshape(-1)
                labels = torch.as_tensor(_field(det, "classes")).float().reshape(-1)
                masks = _field(det, "masks") if self.seg else None
                assert bbox.shape[0] == labels.numel() == scores.numel()
                for j in sorted(range(bbox.shape[0]), key=lambda q: -float(scores[q])):
                    x1, y1, x2, y2 = (float(v) for v in bbox[j])
                    self._match(regs, int(labels[j]), float(scores[j]), x1, y1, x2, y2,
                                masks[j] if masks is not None else None, predict)
        res = [MAPValidationResult(self.classes, self.k, p, gt_cnt, self.type, self.skip) for p in predict]
        return res[0] if len(res) == 1 else MAPMultiIOUValidationResult(res)

    # ---- factories of the reference's companion object (ValidationMethod.scala:762-840)
    @staticmethod
    def coco(nClasses: int, topK: int = -1, skipClass: int = 0, iouThres=(0.5, 0.05, 10), isSegmentation=False):
        start, step, n = iouThres
        return MeanAveragePrecisionObjectDetection(nClasses, topK, [start + i * step for i in range(n)], MAPCOCO,
                                                   skipClass, isSegmentation)

    @staticmethod
    def pascalVOC(nClasses: int, useVoc2007: bool = False, topK: int = -1, skipClass: int = 0):
        return MeanAveragePrecisionObjectDetection(nClasses, topK, (0.5,),
                                                   MAPPascalVoc2007 if useVoc2007 else MAPPascalVoc2010, skipClass)


MeanAveragePrecision.cocoBBox = staticmethod(
    lambda nClasses, topK=-1, skipClass=0, iouThres=(0.5, 0.05, 10):
    MeanAveragePrecisionObjectDetection.coco(nClasses, topK, skipClass, iouThres, False))
MeanAveragePrecision.cocoSegmentation = staticmethod(
    lambda nClasses, topK=-1, skipClass=0, iouThres=(0.5, 0.05, 10):
    MeanAveragePrecisionObjectDetection.coco(nClasses, topK, skipClass, iouThres, True))
MeanAveragePrecision.pascalVOC = staticmethod(MeanAveragePrecisionObjectDetection.pascalVOC)
MeanAveragePrecision.classification = staticmethod(lambda nClasses, topK=-1: MeanAveragePrecision(topK, nClasses))


class MAPUtil:
    """The building blocks of the detection mAP as free functions (ValidationMethod.scala:286-408)."""

    @staticmethod
    def findTopK(k: int, arr, column: int):
        """The ``k`` largest values of one column of a row-major matrix → [(row, value)], best first."""
        col = [(i, float(r[column])) for i, r in enumerate(arr)]
        col.sort(key=lambda p: -p[1])
        return col[:min(k, len(col))]

    @staticmethod
    def gtTablesToGroundTruthRegions(gtTable, classes: int, numIOU: int, isCOCO: bool, isSegmentation: bool):
        m = MeanAveragePrecisionObjectDetection(classes, -1, [0.5] * numIOU, MAPCOCO if isCOCO else MAPPascalVoc2010,
                                                -1, isSegmentation)
        return m._ground_truth(gtTable)

    @staticmethod
    def parseDetection(gtBbox, label, score, x1, y1, x2, y2, mask, classes, iou, predictByClasses):
        m = MeanAveragePrecisionObjectDetection(classes, -1, list(iou))
        m._match(gtBbox, int(label), float(score), x1, y1, x2, y2, mask, predictByClasses)

    @staticmethod
    def parseSegmentationTensorResult(outTensor, func):
        """Calls ``func(imgId, label, score, x1, y1, x2, y2)`` for every detection of a
        ``[n, (label, score, x1, y1, x2, y2) * n]``-per-row output."""
        out = torch.as_tensor(outTensor).float().cpu()
        if out.dim() != 2:
            raise ValueError("the output tensor should have 2 dimensions")
        for img in range(out.shape[0]):
            for b in range(int(out[img, 0])):
                lab, sc, x1, y1, x2, y2 = (float(v) for v in out[img, 1 + 6 * b:7 + 6 * b])
                func(img, int(lab), sc, x1, y1, x2, y2)
