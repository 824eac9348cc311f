"""Segmentation masks and the COCO annotation model (S/dataset/segmentation/{MaskUtils,COCODataset}.scala, 944 LoC):
polygon masks, uncompressed and compressed ("string") RLE in COCO's column-major convention, RLE area / IoU / merge /
decode, polygon → RLE rasterisation, and a json loader producing per-image annotation lists."""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, field

import numpy as np


class SegmentationMasks:
    pass


class PolyMasks(SegmentationMasks):
    def __init__(self, poly: list[list[float]], height: int, width: int):
        self.poly, self.height, self.width = [np.asarray(p, dtype=np.float32) for p in poly], height, width

    def toRLE(self) -> "RLEMasks":
        return MaskUtils.mergeRLEs([MaskUtils.poly2RLE(p, self.height, self.width) for p in self.poly], False)


class RLEMasks(SegmentationMasks):
    """Run lengths over the column-major flattened mask, starting with a run of zeros."""

    def __init__(self, counts, height: int, width: int):
        self.counts, self.height, self.width = np.asarray(counts, dtype=np.int64), height, width

    def toRLE(self):
        return self

    def get(self, i: int) -> int:
        return int(self.counts[i])

    def size(self) -> int:
        return len(self.counts)


class MaskUtils:
    @staticmethod
    def binaryToRLE(mask: np.ndarray) -> RLEMasks:
        h, w = mask.shape
        flat = np.asarray(mask, dtype=np.uint8).T.reshape(-1)           # column-major
        change = np.flatnonzero(np.diff(flat)) + 1
        bounds = np.concatenate([[0], change, [flat.size]])
        counts = np.diff(bounds).tolist()
        if flat.size and flat[0] == 1:
            counts = [0] + counts
        return RLEMasks(counts, h, w)

    @staticmethod
    def RLE2Binary(rle: RLEMasks) -> np.ndarray:
        vals = np.zeros(len(rle.counts), dtype=np.uint8)
        vals[1::2] = 1
        flat = np.repeat(vals, rle.counts)
        return flat.reshape(rle.width, rle.height).T.copy()

    @staticmethod
    def rleArea(rle: RLEMasks) -> int:
        return int(rle.counts[1::2].sum())

    @staticmethod
    def rleIOU(a: RLEMasks, b: RLEMasks, iscrowd: bool = False) -> float:
        ma, mb = MaskUtils.RLE2Binary(a).astype(bool), MaskUtils.RLE2Binary(b).astype(bool)
        inter = float((ma & mb).sum())
        union = float(ma.sum()) if iscrowd else float((ma | mb).sum())
        return inter / union if union > 0 else 0.0

    @staticmethod
    def mergeRLEs(rles: list[RLEMasks], intersect: bool) -> RLEMasks:
        m = MaskUtils.RLE2Binary(rles[0]).astype(bool)
        for r in rles[1:]:
            o = MaskUtils.RLE2Binary(r).astype(bool)
            m = (m & o) if intersect else (m | o)
        return MaskUtils.binaryToRLE(m.astype(np.uint8))

    @staticmethod
    def poly2RLE(poly: np.ndarray, height: int, width: int) -> RLEMasks:
        import cv2

        pts = np.asarray(poly, dtype=np.float64).reshape(-1, 2)
        mask = np.zeros((height, width), dtype=np.uint8)
        cv2.fillPoly(mask, [np.round(pts).astype(np.int32)], 1)
        return MaskUtils.binaryToRLE(mask)

    # ---- COCO's compressed string form (LEB128-like with deltas)
    @staticmethod
    def RLE2String(rle: RLEMasks) -> str:
        out = []
        cnts = rle.counts
        for i, c in enumerate(cnts):
            x = int(c) - (int(cnts[i - 2]) if i > 2 else 0)
            more = True
            while more:
                ch = x & 0x1F
                x >>= 5
                more = (x != -1) if (ch & 0x10) else (x != 0)
                if more:
                    ch |= 0x20
                out.append(chr(ch + 48))
        return "".join(out)

    @staticmethod
    def string2RLE(s: str, height: int, width: int) -> RLEMasks:
        cnts, p = [], 0
        while p < len(s):
            x, k, more = 0, 0, True
            while more:
                c = ord(s[p]) - 48
                x |= (c & 0x1F) << (5 * k)
                more = bool(c & 0x20)
                p += 1
                k += 1
                if not more and (c & 0x10):
                    x |= -1 << (5 * k)
            if len(cnts) > 2:
                x += cnts[-2]
            cnts.append(x)
        return RLEMasks(cnts, height, width)


@dataclass
class COCOAnnotation:
    id: int
    image_id: int
    category_id: int
    bbox: tuple
    area: float
    iscrowd: bool
    segmentation: SegmentationMasks | None = None


@dataclass
class COCOImage:
    id: int
    height: int
    width: int
    file_name: str
    annotations: list = field(default_factory=list)


class COCODataset:
    """``COCODataset.load(jsonPath, imageRoot)`` (COCODataset.scala): images with their annotations, category id ↔
    contiguous 1-based index maps."""

    def __init__(self, images, categories):
        self.images, self.categories = images, categories
        ids = sorted(c["id"] for c in categories)
        self.categoryId2Idx = {cid: i + 1 for i, cid in enumerate(ids)}
        self.categoryIdx2Id = {i: cid for cid, i in self.categoryId2Idx.items()}

    @staticmethod
    def load(json_path: str, image_root: str = ".") -> "COCODataset":
        d = json.load(open(json_path))
        images = {im["id"]: COCOImage(im["id"], im["height"], im["width"], os.path.join(image_root, im["file_name"]))
                  for im in d.get("images", [])}
        for a in d.get("annotations", []):
            im = images[a["image_id"]]
            seg = a.get("segmentation")
            if isinstance(seg, list):
                masks = PolyMasks(seg, im.height, im.width)
            elif isinstance(seg, dict):
                cnt = seg["counts"]
                masks = MaskUtils.string2RLE(cnt, im.height, im.width) if isinstance(cnt, str) else \
                    RLEMasks(cnt, im.height, im.width)
            else:
                masks = None
            im.annotations.append(COCOAnnotation(a["id"], a["image_id"], a["category_id"], tuple(a.get("bbox", ())),
                                                 float(a.get("area", 0)), bool(a.get("iscrowd", 0)), masks))
        return COCODataset(list(images.values()), d.get("categories", []))
