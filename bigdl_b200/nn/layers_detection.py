"""Object-detection layers: Nms, Anchor, PriorBox, Proposal, RegionProposal, DetectionOutputSSD, DetectionOutputFrcnn,
FPN, BoxHead, MaskHead (reference: the same-named files under S/nn/, S/transform/vision/image/util/BboxUtil.scala, and
S/models/maskrcnn/MaskRCNN.scala). Forward-only post-processing layers run under ``no_grad`` like the reference's
``updateGradInput`` stubs.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from ..utils.table import Table, to_table
from .layers_conv_extra import Pooler
from .layers_core import Linear, ReLU, SpatialConvolution
from .module import Container, Module, Sequential


# ---------------------------------------------------------------- box utilities
def box_iou(a: torch.Tensor, b: torch.Tensor, normalized: bool = False) -> torch.Tensor:
    """Pairwise IoU of [N,4] and [M,4] boxes (x1,y1,x2,y2); pixel boxes use the +1 convention of the reference."""
    off = 0.0 if normalized else 1.0
    area_a = (a[:, 2] - a[:, 0] + off) * (a[:, 3] - a[:, 1] + off)
    area_b = (b[:, 2] - b[:, 0] + off) * (b[:, 3] - b[:, 1] + off)
    lt = torch.max(a[:, None, :2], b[None, :, :2])
    rb = torch.min(a[:, None, 2:], b[None, :, 2:])
    wh = (rb - lt + off).clamp(min=0)
    inter = wh[..., 0] * wh[..., 1]
    return inter / (area_a[:, None] + area_b[None, :] - inter).clamp(min=1e-12)


def nms(boxes: torch.Tensor, scores: torch.Tensor, thresh: float, top_k: int = -1, normalized: bool = False):
    """Greedy NMS; returns kept indices sorted by descending score (S/nn/Nms.scala)."""
    if boxes.numel() == 0:
        return torch.zeros(0, dtype=torch.long, device=boxes.device)
    order = scores.argsort(descending=True)
    if top_k > 0:
        order = order[:top_k]
    b = boxes[order].float()
    iou = box_iou(b, b, normalized).cpu()
    n = b.shape[0]
    suppressed = torch.zeros(n, dtype=torch.bool)
    keep = []
    for i in range(n):
        if suppressed[i]:
            continue
        keep.append(i)
        suppressed |= iou[i] > thresh
    return order[torch.tensor(keep, dtype=torch.long, device=order.device)]


def bbox_transform_inv(boxes: torch.Tensor, deltas: torch.Tensor, weights=(1.0, 1.0, 1.0, 1.0),
                       clip: float = math.log(1000.0 / 16)) -> torch.Tensor:
    """Apply (dx, dy, dw, dh) deltas to boxes (BboxUtil.bboxTransformInv)."""
    w = boxes[:, 2] - boxes[:, 0] + 1.0
    h = boxes[:, 3] - boxes[:, 1] + 1.0
    cx, cy = boxes[:, 0] + 0.5 * w, boxes[:, 1] + 0.5 * h
    dx, dy = deltas[:, 0::4] / weights[0], deltas[:, 1::4] / weights[1]
    dw = (deltas[:, 2::4] / weights[2]).clamp(max=clip)
    dh = (deltas[:, 3::4] / weights[3]).clamp(max=clip)
    pcx, pcy = dx * w[:, None] + cx[:, None], dy * h[:, None] + cy[:, None]
    pw, ph = torch.exp(dw) * w[:, None], torch.exp(dh) * h[:, None]
    out = torch.zeros_like(deltas)
    out[:, 0::4], out[:, 1::4] = pcx - 0.5 * pw, pcy - 0.5 * ph
    out[:, 2::4], out[:, 3::4] = pcx + 0.5 * pw - 1, pcy + 0.5 * ph - 1
    return out


def clip_boxes(boxes: torch.Tensor, height: float, width: float) -> torch.Tensor:
    b = boxes.clone()
    b[:, 0::4] = b[:, 0::4].clamp(0, width - 1)
    b[:, 1::4] = b[:, 1::4].clamp(0, height - 1)
    b[:, 2::4] = b[:, 2::4].clamp(0, width - 1)
    b[:, 3::4] = b[:, 3::4].clamp(0, height - 1)
    return b


class Nms(Module):
    """input = table(boxes [N,4], scores [N]); output = kept indices (1-based, like the reference)."""

    def __init__(self, thresh: float = 0.3, top_k: int = -1):
        super().__init__()
        self.thresh, self.top_k = thresh, top_k

    def updateOutput(self, input):
        boxes, scores = to_table(input).to_list()
        with torch.no_grad():
            return nms(boxes, scores, self.thresh, self.top_k) + 1


class Anchor(Module):
    """Faster-RCNN anchor generator (S/nn/Anchor.scala): ``generateAnchors(width, height, feat_stride)`` →
    [H*W*A, 4]."""

    def __init__(self, ratios, scales, base_size: int = 16):
        super().__init__()
        self.ratios, self.scales, self.base_size = list(ratios), list(scales), base_size
        self.basic = self._basic_anchors()
        self.anchor_num = self.basic.shape[0]

    def _basic_anchors(self):
        base = torch.tensor([0.0, 0.0, self.base_size - 1.0, self.base_size - 1.0])
        w, h = base[2] - base[0] + 1, base[3] - base[1] + 1
        cx, cy = base[0] + 0.5 * (w - 1), base[1] + 0.5 * (h - 1)
        out = []
        for r in self.ratios:
            ws = torch.round(torch.sqrt(w * h / r))
            hs = torch.round(ws * r)
            for s in self.scales:
                wss, hss = ws * s, hs * s
                out.append(torch.stack([cx - 0.5 * (wss - 1), cy - 0.5 * (hss - 1), cx + 0.5 * (wss - 1),
                                        cy + 0.5 * (hss - 1)]))
        return torch.stack(out)

    def generateAnchors(self, width: int, height: int, feat_stride: float = 16.0, device="cpu"):
        sx = torch.arange(width, dtype=torch.float32) * feat_stride
        sy = torch.arange(height, dtype=torch.float32) * feat_stride
        yy, xx = torch.meshgrid(sy, sx, indexing="ij")
        shifts = torch.stack([xx, yy, xx, yy], dim=-1).reshape(-1, 1, 4)
        return (shifts + self.basic.reshape(1, -1, 4)).reshape(-1, 4).to(device)

    generate_anchors = generateAnchors

    def updateOutput(self, input):
        h, w = input.shape[-2], input.shape[-1]
        return self.generateAnchors(w, h, device=input.device)


class PriorBox(Module):
    """SSD prior boxes for one feature map: output [1, 2, H*W*P*4] (boxes, variances) (S/nn/PriorBox.scala)."""

    def __init__(self, min_sizes, max_sizes=None, aspect_ratios=None, is_flip=True, is_clip=False, variances=None,
                 offset=0.5, img_h=0, img_w=0, img_size=0, step_h=0.0, step_w=0.0, step=0.0):
        super().__init__()
        self.min_sizes, self.max_sizes = list(min_sizes), list(max_sizes or [])
        ars = [1.0]
        for ar in (aspect_ratios or []):
            if not any(abs(ar - a) < 1e-6 for a in ars):
                ars.append(ar)
                if is_flip:
                    ars.append(1.0 / ar)
        self.aspect_ratios, self.is_clip = ars, is_clip
        self.variances = list(variances) if variances else [0.1]
        self.offset = offset
        self.img_h, self.img_w = (img_size, img_size) if img_size else (img_h, img_w)
        self.step_h, self.step_w = (step, step) if step else (step_h, step_w)
        self.num_priors = len(ars) * len(self.min_sizes) + len(self.max_sizes)

    def updateOutput(self, input):
        feat = input if isinstance(input, torch.Tensor) else to_table(input).to_list()[0]
        H, W = feat.shape[-2], feat.shape[-1]
        img_h, img_w = self.img_h, self.img_w
        if (img_h == 0 or img_w == 0) and not isinstance(input, torch.Tensor):
            img = to_table(input).to_list()[1]
            img_h, img_w = img.shape[-2], img.shape[-1]
        step_h = self.step_h or img_h / H
        step_w = self.step_w or img_w / W
        boxes = []
        for i in range(H):
            for j in range(W):
                cx, cy = (j + self.offset) * step_w, (i + self.offset) * step_h
                for k, ms in enumerate(self.min_sizes):
                    bw = bh = ms
                    boxes.append([cx - bw / 2, cy - bh / 2, cx + bw / 2, cy + bh / 2])
                    if self.max_sizes:
                        bw = bh = math.sqrt(ms * self.max_sizes[k])
                        boxes.append([cx - bw / 2, cy - bh / 2, cx + bw / 2, cy + bh / 2])
                    for ar in self.aspect_ratios:
                        if abs(ar - 1.0) < 1e-6:
                            continue
                        bw, bh = ms * math.sqrt(ar), ms / math.sqrt(ar)
                        boxes.append([cx - bw / 2, cy - bh / 2, cx + bw / 2, cy + bh / 2])
        b = torch.tensor(boxes, dtype=torch.float32, device=feat.device)
        b[:, 0::2] /= img_w
        b[:, 1::2] /= img_h
        if self.is_clip:
            b = b.clamp(0, 1)
        var = torch.tensor(self.variances if len(self.variances) == 4 else self.variances * 4, dtype=torch.float32,
                           device=feat.device).repeat(b.shape[0], 1)
        return torch.stack([b.reshape(-1), var.reshape(-1)]).unsqueeze(0)


class Proposal(Module):
    """RPN proposals (S/nn/Proposal.scala): input = table(cls scores [1,2A,H,W], bbox deltas [1,4A,H,W],
    im_info [1,>=2] (h, w, ...)); output [n, 5] (batch index 0, x1, y1, x2, y2)."""

    def __init__(self, pre_nms_topn, post_nms_topn, ratios, scales, rpn_pre_nms_topn_train=12000,
                 rpn_post_nms_topn_train=2000, min_size=16, nms_thresh=0.7):
        super().__init__()
        self.pre_test, self.post_test = pre_nms_topn, post_nms_topn
        self.pre_train, self.post_train = rpn_pre_nms_topn_train, rpn_post_nms_topn_train
        self.anchor = Anchor(ratios, scales)
        self.min_size, self.nms_thresh = min_size, nms_thresh

    def updateOutput(self, input):
        scores, deltas, im_info = to_table(input).to_list()
        with torch.no_grad():
            A = self.anchor.anchor_num
            H, W = scores.shape[-2], scores.shape[-1]
            anchors = self.anchor.generateAnchors(W, H, device=scores.device)
            sc = scores[0, A:].permute(1, 2, 0).reshape(-1).float()
            dl = deltas[0].permute(1, 2, 0).reshape(-1, 4).float()
            props = clip_boxes(bbox_transform_inv(anchors, dl), float(im_info.reshape(-1)[0]), float(im_info.reshape(-1)[1]))
            scale = float(im_info.reshape(-1)[2]) if im_info.numel() > 2 else 1.0
            ws, hs = props[:, 2] - props[:, 0] + 1, props[:, 3] - props[:, 1] + 1
            keep = torch.nonzero((ws >= self.min_size * scale) & (hs >= self.min_size * scale)).reshape(-1)
            props, sc = props[keep], sc[keep]
            pre, post = (self.pre_train, self.post_train) if self.train else (self.pre_test, self.post_test)
            k = nms(props, sc, self.nms_thresh, pre)
            k = k[:post] if post > 0 else k
            out = torch.zeros(k.numel(), 5, device=scores.device)
            out[:, 1:] = props[k]
            return out


class DetectionOutputParam:
    """The SSD post-processing settings as one value (DetectionOutputSSD.scala:305-308); ``DetectionOutputSSD(param)``
    / ``DetectionOutputSSD.apply(param, postProcess)`` accept it."""

    def __init__(self, nClasses: int = 21, shareLocation: bool = True, bgLabel: int = 0, nmsThresh: float = 0.45,
                 nmsTopk: int = 400, keepTopK: int = 200, confThresh: float = 0.01,
                 varianceEncodedInTarget: bool = False):
        self.nClasses, self.shareLocation, self.bgLabel, self.nmsThresh = nClasses, shareLocation, bgLabel, nmsThresh
        self.nmsTopk, self.keepTopK, self.confThresh = nmsTopk, keepTopK, confThresh
        self.varianceEncodedInTarget = varianceEncodedInTarget


class DetectionOutputSSD(Module):
    """Decode SSD predictions: input = table(loc [B, P*4], conf [B, P*C], priors [1,2,P*4]) → [B, 1 + keep*6]
    (count, then rows (label, score, x1, y1, x2, y2)) (S/nn/DetectionOutputSSD.scala)."""

    def __init__(self, n_classes=21, share_location=True, bg_label=0, nms_thresh=0.45, nms_topk=400, keep_top_k=200,
                 conf_thresh=0.01, variance_encoded_in_target=False, conf_post_process=True):
        super().__init__()
        if isinstance(n_classes, DetectionOutputParam):
            prm, post = n_classes, (share_location if isinstance(share_location, bool) and share_location is not True
                                    else conf_post_process)
            n_classes, share_location, bg_label = prm.nClasses, prm.shareLocation, prm.bgLabel
            nms_thresh, nms_topk, keep_top_k, conf_thresh = prm.nmsThresh, prm.nmsTopk, prm.keepTopK, prm.confThresh
            variance_encoded_in_target, conf_post_process = prm.varianceEncodedInTarget, post
            # what the serializer stores: the unpacked scalars, not the parameter object
            self.__dict__["_ctor_args"] = dict(
                n_classes=n_classes, share_location=share_location, bg_label=bg_label, nms_thresh=nms_thresh,
                nms_topk=nms_topk, keep_top_k=keep_top_k, conf_thresh=conf_thresh,
                variance_encoded_in_target=variance_encoded_in_target, conf_post_process=conf_post_process)
        self.n_classes, self.share_location, self.bg_label = n_classes, share_location, bg_label
        self.nms_thresh, self.nms_topk, self.keep_top_k, self.conf_thresh = nms_thresh, nms_topk, keep_top_k, conf_thresh
        self.variance_encoded_in_target, self.conf_post_process = variance_encoded_in_target, conf_post_process

    @staticmethod
    def apply(param: DetectionOutputParam, postProcess: bool = True):
        return DetectionOutputSSD(param.nClasses, param.shareLocation, param.bgLabel, param.nmsThresh, param.nmsTopk,
                                  param.keepTopK, param.confThresh, param.varianceEncodedInTarget, postProcess)

    def _decode(self, loc, priors, var):
        pw, ph = priors[:, 2] - priors[:, 0], priors[:, 3] - priors[:, 1]
        pcx, pcy = (priors[:, 0] + priors[:, 2]) / 2, (priors[:, 1] + priors[:, 3]) / 2
        if self.variance_encoded_in_target:
            var = torch.ones_like(var)
        cx, cy = var[:, 0] * loc[:, 0] * pw + pcx, var[:, 1] * loc[:, 1] * ph + pcy
        w, h = torch.exp(var[:, 2] * loc[:, 2]) * pw, torch.exp(var[:, 3] * loc[:, 3]) * ph
        return torch.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], dim=1)

    def updateOutput(self, input):
        loc, conf, prior = to_table(input).to_list()
        with torch.no_grad():
            B = loc.shape[0]
            priors = prior[0, 0].reshape(-1, 4).float()
            var = prior[0, 1].reshape(-1, 4).float()
            P = priors.shape[0]
            conf = conf.reshape(B, P, self.n_classes).float()
            if self.conf_post_process:
                conf = F.softmax(conf, dim=2)
            results = []
            for b in range(B):
                boxes = self._decode(loc[b].reshape(P, 4).float(), priors, var)
                dets = []
                for c in range(self.n_classes):
                    if c == self.bg_label:
                        continue
                    sc = conf[b, :, c]
                    m = torch.nonzero(sc > self.conf_thresh).reshape(-1)
                    if m.numel() == 0:
                        continue
                    k = nms(boxes[m], sc[m], self.nms_thresh, self.nms_topk, normalized=True)
                    idx = m[k]
                    dets.append(torch.cat([torch.full((idx.numel(), 1), float(c), device=loc.device),
                                           sc[idx].unsqueeze(1), boxes[idx]], dim=1))
                d = torch.cat(dets) if dets else torch.zeros(0, 6, device=loc.device)
                if self.keep_top_k > 0 and d.shape[0] > self.keep_top_k:
                    d = d[d[:, 1].argsort(descending=True)[:self.keep_top_k]]
                results.append(d)
            n_max = max((r.shape[0] for r in results), default=0)
            out = torch.zeros(B, 1 + n_max * 6, device=loc.device)
            for b, r in enumerate(results):
                out[b, 0] = r.shape[0]
                out[b, 1:1 + r.numel()] = r.reshape(-1)
            return out


class DetectionOutputFrcnn(Module):
    """Faster-RCNN post-processing: input = table(im_info, rois [R,5], box deltas [R, 4C], scores [R, C]) →
    [1, 1 + n*6] rows (label, score, box) (S/nn/DetectionOutputFrcnn.scala)."""

    def __init__(self, nms_thresh=0.3, n_classes=21, bbox_vote=False, max_per_image=100, thresh=0.05):
        super().__init__()
        self.nms_thresh, self.n_classes, self.bbox_vote = nms_thresh, n_classes, bbox_vote
        self.max_per_image, self.thresh = max_per_image, thresh

    def updateOutput(self, input):
        im_info, rois, deltas, scores = to_table(input).to_list()
        with torch.no_grad():
            info = im_info.reshape(-1).float()
            boxes = rois[:, 1:5].float()
            if info.numel() > 2 and float(info[2]) != 0:
                boxes = boxes / info[2]
            pred = clip_boxes(bbox_transform_inv(boxes, deltas.float()), float(info[0]) / max(float(info[2]), 1e-9)
                              if info.numel() > 2 else float(info[0]), float(info[1]) / max(float(info[2]), 1e-9)
                              if info.numel() > 2 else float(info[1]))
            dets = []
            for c in range(1, self.n_classes):
                sc = scores[:, c].float()
                m = torch.nonzero(sc > self.thresh).reshape(-1)
                if m.numel() == 0:
                    continue
                bx = pred[m, 4 * c:4 * c + 4]
                k = nms(bx, sc[m], self.nms_thresh)
                dets.append(torch.cat([torch.full((k.numel(), 1), float(c), device=scores.device),
                                       sc[m][k].unsqueeze(1), bx[k]], dim=1))
            d = torch.cat(dets) if dets else torch.zeros(0, 6, device=scores.device)
            if self.max_per_image > 0 and d.shape[0] > self.max_per_image:
                d = d[d[:, 1].argsort(descending=True)[:self.max_per_image]]
            out = torch.zeros(1, 1 + d.numel(), device=scores.device)
            out[0, 0] = d.shape[0]
            out[0, 1:] = d.reshape(-1)
            return out


class FPN(Container):
    """Feature pyramid network: lateral 1x1 convs, top-down nearest upsampling + add, 3x3 output convs, optional extra
    levels (S/nn/FPN.scala). input = table(C2..C5 features, fine → coarse); output = table(P2..P5[, P6…])."""

    def __init__(self, in_channels, out_channels, top_blocks: int = 0, in_channels_of_p6p7: int = 0,
                 out_channels_of_p6p7: int = 0):
        super().__init__()
        self.inner = [SpatialConvolution(c, out_channels, 1, 1) if c > 0 else None for c in in_channels]
        self.layer = [SpatialConvolution(out_channels, out_channels, 3, 3, 1, 1, 1, 1) if c > 0 else None
                      for c in in_channels]
        self.top_blocks = top_blocks
        self.p6 = self.p7 = None
        if top_blocks == 2:
            self.p6 = SpatialConvolution(in_channels_of_p6p7, out_channels_of_p6p7, 3, 3, 2, 2, 1, 1)
            self.p7 = SpatialConvolution(out_channels_of_p6p7, out_channels_of_p6p7, 3, 3, 2, 2, 1, 1)
        self.modules = [m for m in self.inner + self.layer + [self.p6, self.p7] if m is not None]

    def updateOutput(self, input):
        feats = to_table(input).to_list()
        n = len(feats)
        last = self.inner[n - 1]._call(feats[n - 1])
        results = [self.layer[n - 1]._call(last)]
        for i in range(n - 2, -1, -1):
            if self.inner[i] is None:
                continue
            lateral = self.inner[i]._call(feats[i])
            top = F.interpolate(last, size=lateral.shape[-2:], mode="nearest")
            last = lateral + top
            results.insert(0, self.layer[i]._call(last))
        if self.top_blocks == 1:
            results.append(F.max_pool2d(results[-1], 1, 2))
        elif self.top_blocks == 2:
            p6 = self.p6._call(feats[-1])
            results.extend([p6, self.p7._call(F.relu(p6))])
        return Table(*results)


class RegionProposal(Container):
    """FPN RPN: shared 3x3 conv + 1x1 objectness/box heads on every level, per-level top-k + NMS, merged top-k
    (S/nn/RegionProposal.scala). input = table(features table, im_info [h, w]); output [n, 4] boxes."""

    def __init__(self, in_channels, anchor_sizes, aspect_ratios, anchor_stride, pre_nms_top_n_test=1000,
                 post_nms_top_n_test=1000, pre_nms_top_n_train=2000, post_nms_top_n_train=2000, nms_thread=0.7,
                 min_size=0):
        super().__init__()
        self.anchor_sizes, self.aspect_ratios, self.anchor_stride = list(anchor_sizes), list(aspect_ratios), list(anchor_stride)
        A = len(aspect_ratios)
        self.conv = SpatialConvolution(in_channels, in_channels, 3, 3, 1, 1, 1, 1)
        self.cls_logits = SpatialConvolution(in_channels, A, 1, 1)
        self.bbox_pred = SpatialConvolution(in_channels, A * 4, 1, 1)
        self.modules = [self.conv, self.cls_logits, self.bbox_pred]
        self.pre_test, self.post_test, self.pre_train, self.post_train = (pre_nms_top_n_test, post_nms_top_n_test,
                                                                          pre_nms_top_n_train, post_nms_top_n_train)
        self.nms_thresh, self.min_size = nms_thread, min_size
        self.anchors = [Anchor(self.aspect_ratios, [s / st], base_size=st) for s, st in zip(self.anchor_sizes, self.anchor_stride)]

    def updateOutput(self, input):
        feats, im_info = to_table(input).to_list()
        feats = to_table(feats).to_list()
        info = im_info.reshape(-1).float()
        pre, post = (self.pre_train, self.post_train) if self.train else (self.pre_test, self.post_test)
        all_boxes, all_scores = [], []
        for f, anc, stride in zip(feats, self.anchors, self.anchor_stride):
            t = F.relu(self.conv._call(f))
            logits, deltas = self.cls_logits._call(t), self.bbox_pred._call(t)
            with torch.no_grad():
                H, W = logits.shape[-2:]
                a = anc.generateAnchors(W, H, stride, device=f.device)
                sc = torch.sigmoid(logits[0].permute(1, 2, 0).reshape(-1).float())
                dl = deltas[0].reshape(-1, 4, H, W).permute(2, 3, 0, 1).reshape(-1, 4).float()
                k = min(pre, sc.numel())
                top, idx = sc.topk(k)
                boxes = clip_boxes(bbox_transform_inv(a[idx], dl[idx]), float(info[0]), float(info[1]))
                ws, hs = boxes[:, 2] - boxes[:, 0] + 1, boxes[:, 3] - boxes[:, 1] + 1
                ok = torch.nonzero((ws >= self.min_size) & (hs >= self.min_size)).reshape(-1)
                boxes, top = boxes[ok], top[ok]
                keep = nms(boxes, top, self.nms_thresh)[:post]
                all_boxes.append(boxes[keep])
                all_scores.append(top[keep])
        boxes, scores = torch.cat(all_boxes), torch.cat(all_scores)
        if scores.numel() > post:
            idx = scores.topk(post).indices
            boxes = boxes[idx]
        return boxes


class BoxHead(Container):
    """Mask-RCNN box head: Pooler → 2 FC → class scores + box deltas → per-class NMS (S/nn/BoxHead.scala).
    input = table(features table, proposals [R,4], im_info); output = table(labels, boxes, scores)."""

    def __init__(self, in_channels, resolution, scales, sampling_ratio, score_thresh, nms_thresh, max_per_image,
                 output_size, num_classes):
        super().__init__()
        self.pooler = Pooler(resolution, scales, sampling_ratio)
        self.fc6 = Linear(in_channels * resolution * resolution, output_size)
        self.fc7 = Linear(output_size, output_size)
        self.cls_score = Linear(output_size, num_classes)
        self.bbox_pred = Linear(output_size, num_classes * 4)
        self.modules = [self.fc6, self.fc7, self.cls_score, self.bbox_pred]
        self.score_thresh, self.nms_thresh, self.max_per_image, self.num_classes = (score_thresh, nms_thresh,
                                                                                  max_per_image, num_classes)

    def updateOutput(self, input):
        feats, proposals, im_info = to_table(input).to_list()
        x = self.pooler._call(Table(feats, proposals))
        x = F.relu(self.fc6._call(x.reshape(x.shape[0], -1)))
        x = F.relu(self.fc7._call(x))
        logits, deltas = self.cls_score._call(x), self.bbox_pred._call(x)
        with torch.no_grad():
            info = im_info.reshape(-1).float()
            prob = F.softmax(logits.float(), dim=1)
            boxes = clip_boxes(bbox_transform_inv(proposals.float(), deltas.float(), (10.0, 10.0, 5.0, 5.0)),
                               float(info[0]), float(info[1]))
            out_l, out_b, out_s = [], [], []
            for c in range(1, self.num_classes):
                sc = prob[:, c]
                m = torch.nonzero(sc > self.score_thresh).reshape(-1)
                if m.numel() == 0:
                    continue
                bx = boxes[m, 4 * c:4 * c + 4]
                k = nms(bx, sc[m], self.nms_thresh)
                out_l.append(torch.full((k.numel(),), float(c), device=sc.device))
                out_b.append(bx[k])
                out_s.append(sc[m][k])
            if out_l:
                l, b, s = torch.cat(out_l), torch.cat(out_b), torch.cat(out_s)
                if s.numel() > self.max_per_image:
                    idx = s.topk(self.max_per_image).indices
                    l, b, s = l[idx], b[idx], s[idx]
            else:
                l = torch.zeros(0, device=logits.device)
                b = torch.zeros(0, 4, device=logits.device)
                s = torch.zeros(0, device=logits.device)
        return Table(l, b, s)


class MaskHead(Container):
    """Mask-RCNN mask head: Pooler → conv stack → deconv x2 → per-class 1x1 logits → sigmoid, selecting each box's
    class channel (S/nn/MaskHead.scala). input = table(features table, boxes [R,4], labels [R])."""

    def __init__(self, in_channels, resolution, scales, sampling_ratio, layers, dilation, num_classes,
                 use_gn: bool = False):
        super().__init__()
        from .layers_conv_extra import SpatialFullConvolution

        self.pooler = Pooler(resolution, scales, sampling_ratio)
        convs, c = [], in_channels
        for l in layers:
            convs.append(SpatialConvolution(c, l, 3, 3, 1, 1, dilation, dilation, dilation_w=dilation, dilation_h=dilation))
            c = l
        self.convs = convs
        self.deconv = SpatialFullConvolution(c, c, 2, 2, 2, 2)
        self.logits = SpatialConvolution(c, num_classes, 1, 1)
        self.modules = convs + [self.deconv, self.logits]

    def updateOutput(self, input):
        feats, boxes, labels = to_table(input).to_list()
        x = self.pooler._call(Table(feats, boxes))
        for cv in self.convs:
            x = F.relu(cv._call(x))
        x = F.relu(self.deconv._call(x))
        m = torch.sigmoid(self.logits._call(x).float())
        idx = labels.long().reshape(-1)
        sel = m[torch.arange(m.shape[0], device=m.device), idx.clamp(0, m.shape[1] - 1)]
        return Table(x, sel.unsqueeze(1))
