"""Box geometry used by the detection pipeline: ``BoundingBox`` (S/transform/vision/image/util/BoundingBox.scala:28-167),
the tensor helpers of ``BboxUtil`` (util/BboxUtil.scala:23-581) and the SSD ``BatchSampler``
(label/roi/BatchSampler.scala:39-134).

The reference walks JVM arrays element by element; here every helper is one vectorised torch expression over the
[N, 4] box tensor, so the same functions serve as host-side post-processing and (given CUDA tensors) run on the device.
Indices handed back to callers are 0-based tensors / dicts; the one place the reference's 1-based convention is part
of a file format (the ``[num, (label, score, x1, y1, x2, y2) * num]`` detection vector) is kept as is."""
from __future__ import annotations

import math

import numpy as np
import torch

from ...utils.random_generator import RNG


class BoundingBox:
    """(x1, y1, x2, y2), normalised to [0, 1] unless ``normalized`` is False."""

    __slots__ = ("x1", "y1", "x2", "y2", "normalized", "label", "difficult")

    def __init__(self, x1=0.0, y1=0.0, x2=1.0, y2=1.0, normalized: bool = True):
        if isinstance(x1, BoundingBox):
            x1, y1, x2, y2 = x1.x1, x1.y1, x1.x2, x1.y2
        self.x1, self.y1, self.x2, self.y2 = float(x1), float(y1), float(x2), float(y2)
        self.normalized, self.label, self.difficult = normalized, -1.0, -1.0

    def setLabel(self, l):
        self.label = float(l)
        return self

    def setDifficult(self, d):
        self.difficult = float(d)
        return self

    def centerX(self): return (self.x1 + self.x2) / 2
    def centerY(self): return (self.y1 + self.y2) / 2
    def width(self): return self.x2 - self.x1
    def height(self): return self.y2 - self.y1
    def area(self): return self.width() * self.height()

    def clipBox(self, clipped: "BoundingBox"):
        if self.normalized:
            clipped.x1, clipped.y1 = max(min(self.x1, 1.0), 0.0), max(min(self.y1, 1.0), 0.0)
            clipped.x2, clipped.y2 = max(min(self.x2, 1.0), 0.0), max(min(self.y2, 1.0), 0.0)

    def scaleBox(self, height, width, scaled: "BoundingBox"):
        scaled.x1, scaled.y1, scaled.x2, scaled.y2 = self.x1 * width, self.y1 * height, self.x2 * width, self.y2 * height

    def meetEmitCenterConstraint(self, bbox: "BoundingBox") -> bool:
        """Does the centre of ``bbox`` lie inside this box?"""
        return self.x1 <= bbox.centerX() <= self.x2 and self.y1 <= bbox.centerY() <= self.y2

    def isOverlap(self, bbox: "BoundingBox") -> bool:
        return not (bbox.x1 >= self.x2 or bbox.x2 <= self.x1 or bbox.y1 >= self.y2 or bbox.y2 <= self.y1)

    def jaccardOverlap(self, bbox: "BoundingBox") -> float:
        w = min(self.x2, bbox.x2) - max(self.x1, bbox.x1)
        h = min(self.y2, bbox.y2) - max(self.y1, bbox.y1)
        if w < 0 or h < 0:
            return 0.0
        inter = w * h
        return inter / (self.area() + bbox.area() - inter)

    def projectBbox(self, bbox: "BoundingBox", proj: "BoundingBox") -> bool:
        """``bbox`` expressed in the coordinate frame of this box, clipped; False when nothing is left."""
        if not self.isOverlap(bbox):
            return False
        sw, sh = self.width(), self.height()
        proj.x1, proj.y1 = (bbox.x1 - self.x1) / sw, (bbox.y1 - self.y1) / sh
        proj.x2, proj.y2 = (bbox.x2 - self.x1) / sw, (bbox.y2 - self.y1) / sh
        proj.clipBox(proj)
        return proj.area() > 0

    def locateBBox(self, box: "BoundingBox", loc: "BoundingBox"):
        """The inverse of ``projectBbox``: ``box`` (relative to this box) in this box's parent frame."""
        sw, sh = self.width(), self.height()
        x1, y1 = self.x1 + box.x1 * sw, self.y1 + box.y1 * sh
        x2, y2 = self.x1 + box.x2 * sw, self.y1 + box.y2 * sh
        loc.x1, loc.y1, loc.x2, loc.y2 = x1, y1, x2, y2

    def toTuple(self):
        return self.x1, self.y1, self.x2, self.y2

    def __iter__(self):
        return iter(self.toTuple())

    def __eq__(self, o):
        return isinstance(o, BoundingBox) and self.toTuple() == o.toTuple()

    def __hash__(self):
        return hash(self.toTuple())

    def __repr__(self):
        return f"BoundingBox ({self.x1}, {self.y1}, {self.x2}, {self.y2})"


class BatchSampler:
    """Draws up to ``maxSample`` boxes (≤ ``maxTrials`` tries) of random scale / aspect ratio whose Jaccard overlap
    with SOME ground-truth box lies inside [minOverlap, maxOverlap] (either bound may be absent)."""

    def __init__(self, maxSample: int = 1, maxTrials: int = 50, minScale: float = 1, maxScale: float = 1,
                 minAspectRatio: float = 1, maxAspectRatio: float = 1, minOverlap=None, maxOverlap=None):
        if not minScale <= maxScale:
            raise ValueError("minScale must <= maxScale")
        if not (0 < minScale <= 1 and 0 < maxScale <= 1):
            raise ValueError("minScale and maxScale must be in (0, 1]")
        if not 0 < minAspectRatio <= 1:
            raise ValueError("minAspectRatio must be in (0, 1]")
        if not maxAspectRatio >= 1:
            raise ValueError("maxAspectRatio must be >= 1")
        if minOverlap is not None and not 0 <= minOverlap <= 1:
            raise ValueError("minOverlap must be in [0, 1]")
        self.maxSample, self.maxTrials = maxSample, maxTrials
        self.minScale, self.maxScale = minScale, maxScale
        self.minAspectRatio, self.maxAspectRatio = minAspectRatio, maxAspectRatio
        self.minOverlap, self.maxOverlap = minOverlap, maxOverlap

    def satisfySampleConstraint(self, sampled: BoundingBox, target) -> bool:
        if self.minOverlap is None and self.maxOverlap is None:
            return True
        b = np.asarray(target.bboxes, dtype=np.float32).reshape(-1, 4)
        if b.shape[0] == 0:
            return False
        w = np.minimum(b[:, 2], sampled.x2) - np.maximum(b[:, 0], sampled.x1)
        h = np.minimum(b[:, 3], sampled.y2) - np.maximum(b[:, 1], sampled.y1)
        inter = np.where((w < 0) | (h < 0), 0.0, w * h)
        union = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1]) + sampled.area() - inter
        ov = inter / np.where(union == 0, 1.0, union)
        ok = np.ones_like(ov, dtype=bool)
        if self.minOverlap is not None:
            ok &= ov >= self.minOverlap
        if self.maxOverlap is not None:
            ok &= ov <= self.maxOverlap
        return bool(ok.any())

    def _sampleBox(self) -> BoundingBox:
        scale = RNG.uniform(self.minScale, self.maxScale)
        ratio = RNG.uniform(self.minAspectRatio, self.maxAspectRatio)
        ratio = min(max(ratio, scale * scale), 1.0 / scale / scale)
        w, h = scale * math.sqrt(ratio), scale / math.sqrt(ratio)
        x1, y1 = RNG.uniform(0, 1 - w), RNG.uniform(0, 1 - h)
        return BoundingBox(x1, y1, x1 + w, y1 + h)

    def sample(self, source: BoundingBox, target, sampled: list):
        found = 0
        for _ in range(self.maxTrials):
            if found >= self.maxSample:
                return
            box = self._sampleBox()
            source.locateBBox(box, box)
            if self.satisfySampleConstraint(box, target):
                found += 1
                sampled.append(box)

    @staticmethod
    def generateBatchSamples(label, batchSamplers, sampledBoxes: list):
        sampledBoxes.clear()
        unit = BoundingBox(0, 0, 1, 1)
        for s in batchSamplers:
            s.sample(unit, label, sampledBoxes)


def _t(x) -> torch.Tensor:
    return x if isinstance(x, torch.Tensor) else torch.as_tensor(np.asarray(x), dtype=torch.float32)


class BboxUtil:
    """Static tensor helpers; every function takes / returns ``torch.Tensor`` (numpy arrays are accepted)."""

    @staticmethod
    def decodeRois(output):
        """``[num, (label, score, x1, y1, x2, y2) × num, …]`` → [num, 6]; 2-D input is already decoded."""
        output = _t(output)
        if output.numel() < 6 or output.dim() == 2:
            return output
        num = int(output.reshape(-1)[0])
        if num < 0:
            raise ValueError("negative detection count")
        if num == 0:
            return output.new_zeros(0)
        return output.reshape(-1)[1:1 + num * 6].view(num, 6)

    @staticmethod
    def scaleBBox(bboxes, height: float, width: float):
        """In place: x *= width, y *= height."""
        if bboxes.numel() == 0:
            return bboxes
        bboxes[:, 0::2] *= width
        bboxes[:, 1::2] *= height
        return bboxes

    @staticmethod
    def bboxTransformInv(boxes, deltas, normalized: bool = False):
        """Apply (dx, dy, dw, dh) regression deltas ([N, 4a]) to ``boxes`` [N, 4]; pixel boxes count both ends
        (width = x2 − x1 + 1) unless ``normalized``."""
        boxes, deltas = _t(boxes), _t(deltas)
        if boxes.shape[0] == 0:
            return boxes
        if boxes.shape[1] != 4 or deltas.shape[1] % 4:
            raise ValueError(f"boxes {tuple(boxes.shape)} / deltas {tuple(deltas.shape)} are not N*4 / N*4a")
        off = 0.0 if normalized else 1.0
        w = (boxes[:, 2] - boxes[:, 0] + off)[:, None]
        h = (boxes[:, 3] - boxes[:, 1] + off)[:, None]
        cx, cy = boxes[:, 0:1] + w / 2, boxes[:, 1:2] + h / 2
        pcx, pcy = deltas[:, 0::4] * w + cx, deltas[:, 1::4] * h + cy
        pw, ph = torch.exp(deltas[:, 2::4]) * w / 2, torch.exp(deltas[:, 3::4]) * h / 2
        out = torch.empty_like(deltas)
        out[:, 0::4], out[:, 1::4], out[:, 2::4], out[:, 3::4] = pcx - pw, pcy - ph, pcx + pw, pcy + ph
        return out

    @staticmethod
    def clipBoxes(boxes, height: float | None = None, width: float | None = None, minH: float = 0, minW: float = 0,
                  scores=None):
        """With a size: clamp pixel boxes ([N, 4a], in place) to the image and, given ``scores``, zero the score of
        every box smaller than (minH, minW) — returns how many survive. Without a size: clamp normalised boxes to
        [0, 1] and return them (the reference's two overloads)."""
        if height is None:
            return boxes.clamp_(0.0, 1.0)
        if boxes.shape[1] % 4:
            raise ValueError("boxes should have the shape N*4a")
        boxes[:, 0::2].clamp_(0, width - 1)
        boxes[:, 1::2].clamp_(0, height - 1)
        if scores is None:
            return 0
        keep = ((boxes[:, 2::4] - boxes[:, 0::4] + 1) >= minW) & ((boxes[:, 3::4] - boxes[:, 1::4] + 1) >= minH)
        flat = scores.view(-1)[: keep.numel()]
        flat[~keep.reshape(-1)] = 0
        return int(keep.sum())

    @staticmethod
    def getLocPredictions(loc, numPredsPerClass: int, numClasses: int, shareLocation: bool, locPredsBuf=None):
        """[batch, P·C·4] → per image, per class [P, 4] (``shareLocation``: one class, stored LAST)."""
        loc = _t(loc)
        v = loc.reshape(loc.shape[0], numPredsPerClass, numClasses, 4)
        out = []
        for i in range(loc.shape[0]):
            if shareLocation:
                n_slots = len(locPredsBuf[i]) if locPredsBuf is not None else numClasses
                per = [v.new_zeros(numPredsPerClass, 4) for _ in range(n_slots)]
                per[-1] = v[i, :, numClasses - 1].clone()
            else:
                per = [v[i, :, c].clone() for c in range(numClasses)]
            out.append(per)
        return out

    @staticmethod
    def getConfidenceScores(conf, numPredsPerClass: int, numClasses: int, confBuf=None):
        """[batch, P·C] → per image, per class [P]."""
        conf = _t(conf)
        v = conf.reshape(conf.shape[0], numPredsPerClass, numClasses)
        return [[v[i, :, c].clone() for c in range(numClasses)] for i in range(conf.shape[0])]

    @staticmethod
    def getPriorBboxes(prior, nPriors: int):
        """The PriorBox output holds the boxes then their variances, both [nPriors, 4]."""
        flat = _t(prior).reshape(-1)
        return flat[: nPriors * 4].view(nPriors, 4), flat[nPriors * 4: nPriors * 8].view(nPriors, 4)

    @staticmethod
    def decodeBoxes(priorBoxes, priorVariances, isClipBoxes: bool, bboxes, varianceEncodedInTarget: bool, output=None):
        """SSD centre-size decoding of ``bboxes`` [N, 4] against the priors."""
        pb, pv, bb = _t(priorBoxes), _t(priorVariances), _t(bboxes)
        if pb.shape[0] != pv.shape[0] or pb.shape[0] != bb.shape[0]:
            raise ValueError("priors, variances and predictions must have the same number of rows")
        pw, ph = pb[:, 2] - pb[:, 0], pb[:, 3] - pb[:, 1]
        if pb.shape[0] and (bool((pw <= 0).any()) or bool((ph <= 0).any())):
            raise ValueError("prior boxes must have positive width and height")
        pcx, pcy = (pb[:, 0] + pb[:, 2]) / 2, (pb[:, 1] + pb[:, 3]) / 2
        d = bb if varianceEncodedInTarget else bb * pv
        cx, cy = d[:, 0] * pw + pcx, d[:, 1] * ph + pcy
        w, h = torch.exp(d[:, 2]) * pw, torch.exp(d[:, 3]) * ph
        res = torch.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)
        if isClipBoxes:
            res.clamp_(0.0, 1.0)
        if output is not None:
            output.resize_(res.shape).copy_(res)
            return output
        return res

    @staticmethod
    def decodeBboxesAll(allLocPreds, priorBoxes, priorVariances, nClasses: int, bgLabel: int, clipBoxes: bool,
                        varianceEncodedInTarget: bool, shareLocation: bool, output=None):
        out = output if output is not None else [[None] * nClasses for _ in allLocPreds]
        for i, per in enumerate(allLocPreds):
            for c in range(nClasses):
                if shareLocation or c != bgLabel:
                    out[i][c] = BboxUtil.decodeBoxes(priorBoxes, priorVariances, clipBoxes, per[c],
                                                     varianceEncodedInTarget)
        return out

    @staticmethod
    def getAreas(boxes, areas=None, startInd: int = 1, normalized: bool = False):
        """Areas of the boxes whose x1 sits in (1-based) column ``startInd``."""
        boxes = _t(boxes)
        if boxes.numel() == 0:
            return areas if areas is not None else boxes.new_zeros(0)
        s = startInd - 1
        off = 0.0 if normalized else 1.0
        a = (boxes[:, s + 2] - boxes[:, s] + off) * (boxes[:, s + 3] - boxes[:, s + 1] + off)
        if areas is not None:
            areas.resize_(a.shape).copy_(a)
            return areas
        return a

    @staticmethod
    def bboxVote(scoresNms, bboxNms, scoresAll, bboxAll, areasBuf=None):
        """Box voting: every kept box becomes the score-weighted mean of all candidates overlapping it by ≥ 0.5."""
        from .image import RoiLabel

        sn, bn, sa, ba = _t(scoresNms), _t(bboxNms), _t(scoresAll), _t(bboxAll)
        area_all = BboxUtil.getAreas(ba, areasBuf)
        area_nms = BboxUtil.getAreas(bn)
        iw = torch.min(bn[:, None, 2], ba[None, :, 2]) - torch.max(bn[:, None, 0], ba[None, :, 0]) + 1
        ih = torch.min(bn[:, None, 3], ba[None, :, 3]) - torch.max(bn[:, None, 1], ba[None, :, 1]) + 1
        inter = iw * ih
        ov = inter / (area_nms[:, None] + area_all[None, :] - inter)
        wgt = ((iw > 0) & (ih > 0) & (ov >= 0.5)).to(sa.dtype) * sa[None, :]
        voted = (wgt @ ba[:, :4]) / wgt.sum(1, keepdim=True)
        bn[:, :4] = voted
        return RoiLabel(sn.numpy() if sn.device.type == "cpu" else sn.cpu().numpy(),
                        bn.numpy() if bn.device.type == "cpu" else bn.cpu().numpy())

    @staticmethod
    def getGroundTruthIndices(result):
        """Rows of a [N, 7] ground-truth tensor are grouped by the image id in column 0 → {id: (start, length)}
        (0-based start)."""
        result = _t(result)
        if result.numel() == 0:
            return {}
        ids = result[:, 0].tolist()
        out, start = {}, 0
        for i in range(1, len(ids) + 1):
            if i == len(ids) or ids[i] != ids[start]:
                out[int(ids[start])] = (start, i - start)
                start = i
        return out

    @staticmethod
    def getGroundTruths(result):
        """{running index: rows of that image}; an image whose only row has label −1 carries no ground truth and is
        left out (its index is still consumed)."""
        result = _t(result)
        out = {}
        for ind, (_, (s, n)) in enumerate(sorted(BboxUtil.getGroundTruthIndices(result).items())):
            gt = result[s:s + n]
            if n > 1 or float(gt[0, 1]) != -1:
                out[ind] = gt
        return out

    @staticmethod
    def decodeWithWeight(encodeBox, bbox, weight, decodeBox=None):
        """Mask R-CNN box-head decoding: deltas ([N, 4a]) divided by ``weight`` = (wx, wy, ww, wh), dw / dh capped at
        62.5, pixel boxes with both ends counted."""
        eb, bb = _t(encodeBox), _t(bbox)
        if eb.shape[0] != bb.shape[0]:
            raise ValueError("encodeBox and bbox must have the same number of rows")
        w, h = (bb[:, 2] - bb[:, 0] + 1)[:, None], (bb[:, 3] - bb[:, 1] + 1)[:, None]
        cx, cy = bb[:, 0:1] + w / 2, bb[:, 1:2] + h / 2
        dx, dy = eb[:, 0::4] / weight[0], eb[:, 1::4] / weight[1]
        dw, dh = (eb[:, 2::4] / weight[2]).clamp(max=62.5), (eb[:, 3::4] / weight[3]).clamp(max=62.5)
        pcx, pcy = dx * w + cx, dy * h + cy
        pw, ph = torch.exp(dw) * w * 0.5, torch.exp(dh) * h * 0.5
        res = torch.empty_like(eb)
        res[:, 0::4], res[:, 1::4], res[:, 2::4], res[:, 3::4] = pcx - pw, pcy - ph, pcx + pw - 1, pcy + ph - 1
        if decodeBox is not None:
            decodeBox.resize_(res.shape).copy_(res)
            return decodeBox
        return res


class OpenCVMat:
    """The image container of the vision pipeline (opencv/OpenCVMat.scala:31-230). Here an image IS a numpy HWC array
    (what ``cv2`` itself uses); this class keeps the reference's constructor helpers and pickles as an encoded PNG the
    way the JVM class serialises."""

    def __init__(self, mat=None):
        self.mat = None if mat is None else np.array(mat.mat if isinstance(mat, OpenCVMat) else mat)
        self.isReleased = False

    # -- instance
    def shape(self):
        m = self.mat
        return (m.shape[0], m.shape[1], 1 if m.ndim == 2 else m.shape[2])

    def height(self): return self.mat.shape[0]
    def width(self): return self.mat.shape[1]
    def channels(self): return 1 if self.mat.ndim == 2 else self.mat.shape[2]

    def release(self):
        self.mat, self.isReleased = None, True

    def __array__(self, dtype=None, copy=None):
        return self.mat if dtype is None else self.mat.astype(dtype)

    def drawBoundingBox(self, bbox: BoundingBox, text: str, font=None, boxColor=(0, 255, 0),
                        textColor=(255, 255, 255)):
        import cv2

        self.mat = np.ascontiguousarray(self.mat)
        cv2.rectangle(self.mat, (int(bbox.x1), int(bbox.y1)), (int(bbox.x2), int(bbox.y2)), boxColor, 3)
        cv2.putText(self.mat, text, (int(bbox.x1), int(bbox.y1 - 2)),
                    cv2.FONT_HERSHEY_COMPLEX_SMALL if font is None else font, 1, textColor, 1)
        return self

    def __getstate__(self):
        m = self.mat
        try:
            data = OpenCVMat.imencode(self) if m is not None and m.size else b""
        except Exception:
            data = b""
        return {"dtype": None if m is None else str(m.dtype), "png": data}

    def __setstate__(self, st):
        self.isReleased = False
        if not st["png"]:
            self.mat = np.zeros((0, 0), dtype=st["dtype"] or "uint8")
        else:
            self.mat = OpenCVMat.fromImageBytes(st["png"]).mat.astype(st["dtype"])

    # -- constructors / converters
    @staticmethod
    def read(path: str) -> "OpenCVMat":
        with open(path, "rb") as f:
            return OpenCVMat.fromImageBytes(f.read())

    @staticmethod
    def fromImageBytes(fileContent: bytes) -> "OpenCVMat":
        import cv2

        m = cv2.imdecode(np.frombuffer(fileContent, dtype=np.uint8), cv2.IMREAD_UNCHANGED)
        if m is None:
            raise ValueError("cannot decode image bytes")
        return OpenCVMat(m)

    @staticmethod
    def imencode(mat, encoding: str = "png") -> bytes:
        import cv2

        m = np.asarray(mat.mat if isinstance(mat, OpenCVMat) else mat)
        if m.dtype != np.uint8:
            m = np.clip(m, 0, 255).astype(np.uint8)
        ok, buf = cv2.imencode("." + encoding, m)
        if not ok:
            raise ValueError(f"cannot encode image as {encoding}")
        return buf.tobytes()

    @staticmethod
    def fromFloats(floats, height: int, width: int, channel: int = 3) -> "OpenCVMat":
        if not 1 <= channel <= 4:
            raise ValueError(f"channel {channel} is out of range [1,4]")
        a = np.asarray(floats, dtype=np.float32).reshape(-1)
        if a.size < height * width * channel:
            raise ValueError(f"pixels array length {a.size} is less than height*width*channel")
        return OpenCVMat(a[: height * width * channel].reshape(height, width, channel))

    @staticmethod
    def fromPixelsBytes(pixels, height: int, width: int, channel: int = 3) -> "OpenCVMat":
        if not 1 <= channel <= 4:
            raise ValueError(f"channel {channel} is out of range [1,4]")
        a = np.frombuffer(bytes(pixels), dtype=np.uint8)
        if a.size < height * width * channel:
            raise ValueError(f"pixels array length {a.size} is less than height*width*channel")
        return OpenCVMat(a[: height * width * channel].reshape(height, width, channel))

    @staticmethod
    def toBytePixels(input, buffer=None):
        m = np.asarray(input.mat if isinstance(input, OpenCVMat) else input)
        m8 = m if m.dtype == np.uint8 else np.clip(np.rint(m), 0, 255).astype(np.uint8)
        h, w = m8.shape[:2]
        return m8.tobytes(), h, w, (1 if m8.ndim == 2 else m8.shape[2])

    @staticmethod
    def toFloatPixels(input, buffer=None):
        m = np.asarray(input.mat if isinstance(input, OpenCVMat) else input)
        h, w = m.shape[:2]
        return m.astype(np.float32).reshape(-1), h, w, (1 if m.ndim == 2 else m.shape[2])

    @staticmethod
    def fromTensor(tensor, format: str = "HWC") -> "OpenCVMat":
        if format not in ("HWC", "CHW"):
            raise ValueError("the format should be HWC or CHW")
        t = tensor.detach().cpu().float() if isinstance(tensor, torch.Tensor) else torch.as_tensor(np.asarray(tensor))
        if format == "CHW":
            t = t.permute(1, 2, 0)
        return OpenCVMat(t.contiguous().numpy())
