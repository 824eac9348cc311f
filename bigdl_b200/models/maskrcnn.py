"""Mask R-CNN with a ResNet-50 + FPN backbone, inference only (S/models/maskrcnn/MaskRCNN.scala:30-260,
Utils.scala decodeMaskInImage). Input: Table(images [N,3,H,W], imageInfo [N,4] = (height, width, original height,
original width)); output per image: Table with masks (RLE), bboxes, classes, scores — the layout the COCO evaluator
(MeanAveragePrecisionObjectDetection) consumes."""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import torch

from ..dataset.segmentation import MaskUtils
from ..nn.layer import (BoxHead, CAddTable, ConcatTable, FPN, Identity, Input, MaskHead, Model, RegionProposal, ReLU,
                        Sequential, SpatialMaxPooling)
from ..nn.module import Container
from ..utils.table import Table, to_table
from .resnet import Convolution, Sbn


@dataclass
class MaskRCNNParams:
    anchorSizes: list = field(default_factory=lambda: [32.0, 64.0, 128.0, 256.0, 512.0])
    aspectRatios: list = field(default_factory=lambda: [0.5, 1.0, 2.0])
    anchorStride: list = field(default_factory=lambda: [4.0, 8.0, 16.0, 32.0, 64.0])
    preNmsTopNTest: int = 1000
    postNmsTopNTest: int = 1000
    preNmsTopNTrain: int = 2000
    postNmsTopNTrain: int = 2000
    rpnNmsThread: float = 0.7
    minSize: int = 0
    boxResolution: int = 7
    maskResolution: int = 14
    scales: list = field(default_factory=lambda: [0.25, 0.125, 0.0625, 0.03125])
    samplingRatio: int = 2
    boxScoreThresh: float = 0.05
    boxNmsThread: float = 0.5
    maxPerImage: int = 100
    outputSize: int = 1024
    layers: list = field(default_factory=lambda: [256, 256, 256, 256])
    dilation: int = 1
    useGn: bool = False


def _bottleneck(n_in, internal, n_out, stride, use_conv):
    s = (Sequential().add(Convolution(n_in, internal, 1, 1, stride, stride)).add(Sbn(internal)).add(ReLU(True))
         .add(Convolution(internal, internal, 3, 3, 1, 1, 1, 1)).add(Sbn(internal)).add(ReLU(True))
         .add(Convolution(internal, n_out, 1, 1)).add(Sbn(n_out)))
    sc = Sequential().add(Convolution(n_in, n_out, 1, 1, stride, stride)).add(Sbn(n_out)) if use_conv else Identity()
    return Sequential().add(ConcatTable().add(s).add(sc)).add(CAddTable(True)).add(ReLU(True))


def _stage(count, n_in, internal, n_out, stride=1):
    s = Sequential().add(_bottleneck(n_in, internal, n_out, stride, True))
    for _ in range(2, count + 1):
        s.add(_bottleneck(n_out, internal, n_out, 1, False))
    return s


def build_resnet50_fpn_backbone(in_channels: int = 256, out_channels: int = 256):
    stem = (Sequential().add(Convolution(3, 64, 7, 7, 2, 2, 3, 3, propagate_back=False)).add(Sbn(64)).add(ReLU(True))
            .add(SpatialMaxPooling(3, 3, 2, 2, 1, 1)))
    inp = Input()
    n0 = stem(inp)
    n1 = _stage(3, 64, 64, in_channels, 1)(n0)
    n2 = _stage(4, in_channels, 128, in_channels * 2, 2)(n1)
    n3 = _stage(6, in_channels * 2, 256, in_channels * 4, 2)(n2)
    n4 = _stage(3, in_channels * 4, 512, in_channels * 8, 2)(n3)
    resnet = Model([inp], [n1, n2, n3, n4])
    fpn = FPN([in_channels, in_channels * 2, in_channels * 4, in_channels * 8], out_channels, top_blocks=1)
    return Sequential().add(resnet).add(fpn)


def decode_mask_in_image(mask: torch.Tensor, box: torch.Tensor, height: int, width: int, thresh: float = 0.5):
    """Paste one [M,M] soft mask into its box of a [height,width] image (Utils.decodeMaskInImage): the box is padded by
    one pixel of mask, resized to the box size, thresholded and clipped to the image."""
    import cv2

    m = mask.detach().float().cpu().numpy()
    M = m.shape[-1]
    padded = np.zeros((M + 2, M + 2), dtype=np.float32)
    padded[1:-1, 1:-1] = m
    scale = (M + 2) / M
    x1, y1, x2, y2 = [float(v) for v in box.reshape(-1)[:4]]
    cx, cy, hw, hh = (x1 + x2) / 2, (y1 + y2) / 2, (x2 - x1) / 2 * scale, (y2 - y1) / 2 * scale
    bx1, by1, bx2, by2 = int(cx - hw), int(cy - hh), int(cx + hw), int(cy + hh)
    w, h = max(bx2 - bx1 + 1, 1), max(by2 - by1 + 1, 1)
    resized = cv2.resize(padded, (w, h), interpolation=cv2.INTER_LINEAR) > thresh
    out = np.zeros((height, width), dtype=np.uint8)
    xs, ys, xe, ye = max(bx1, 0), max(by1, 0), min(bx2 + 1, width), min(by2 + 1, height)
    if xe > xs and ye > ys:
        out[ys:ye, xs:xe] = resized[ys - by1:ye - by1, xs - bx1:xe - bx1]
    return out


class MaskRCNN(Container):
    def __init__(self, in_channels: int = 256, out_channels: int = 256, num_classes: int = 81,
                 config: MaskRCNNParams | None = None):
        super().__init__()
        c = config or MaskRCNNParams()
        self.config, self.num_classes = c, num_classes
        self.backbone = build_resnet50_fpn_backbone(in_channels, out_channels)
        self.rpn = RegionProposal(out_channels, c.anchorSizes, c.aspectRatios, c.anchorStride, c.preNmsTopNTest,
                                  c.postNmsTopNTest, c.preNmsTopNTrain, c.postNmsTopNTrain, c.rpnNmsThread, c.minSize)
        self.box_head = BoxHead(out_channels, c.boxResolution, c.scales, c.samplingRatio, c.boxScoreThresh,
                                c.boxNmsThread, c.maxPerImage, c.outputSize, num_classes)
        self.mask_head = MaskHead(out_channels, c.maskResolution, c.scales, c.samplingRatio, c.layers, c.dilation,
                                  num_classes, c.useGn)
        self.modules = [self.backbone, self.rpn, self.box_head, self.mask_head]

    def updateOutput(self, input):
        images, image_info = to_table(input).to_list()
        batch_info = torch.tensor([float(images.shape[2]), float(images.shape[3])], device=images.device)
        feats = self.backbone._call(images)
        outputs = Table()
        # the detection heads work image by image (variable number of boxes)
        feats_list = to_table(feats).to_list()
        for i in range(images.shape[0]):
            f_i = Table(*[f[i:i + 1] for f in feats_list])
            proposals = self.rpn._call(Table(f_i, batch_info))
            box_out = to_table(self.box_head._call(Table(f_i, proposals, batch_info)))
            labels, boxes_t, scores = box_out[1], box_out[2], box_out[3]
            masks = to_table(self.mask_head._call(Table(f_i, boxes_t, labels)))
            if self.isTraining():
                outputs[i + 1] = Table(boxes_t, labels, masks, scores)
                continue
            info = image_info[i].reshape(-1)
            h, w, oh, ow = [int(v) for v in info[:4]]
            bx = boxes_t.clone().float()
            if (h, w) != (oh, ow):
                bx = bx * torch.tensor([ow / w, oh / h, ow / w, oh / h], device=bx.device)
            soft = masks[2]
            rles = [MaskUtils.binaryToRLE(decode_mask_in_image(soft[j, 0], bx[j], oh, ow)) for j in range(bx.shape[0])]
            res = Table()
            res["masks"], res["bboxes"], res["classes"], res["scores"] = rles, bx, labels, scores
            outputs[i + 1] = res
        return outputs

    def updateGradInput(self, input, gradOutput):
        raise NotImplementedError("MaskRCNN model only support inference now")
