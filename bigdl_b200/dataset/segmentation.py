"""Segmentation masks and the COCO annotation model (S/dataset/segmentation/{MaskUtils,COCODataset}.scala).

* ``PolyMasks`` / ``RLEMasks``: polygon lists and uncompressed run-length masks in COCO's column-major convention
  (runs alternate zeros / ones, starting with zeros).
* ``MaskUtils``: COCO's compressed "string" RLE codec, polygon rasterisation with the COCO API's sub-pixel rule
  (5x upsampled boundary walk, ``rleFrPoly``), union / intersection of RLEs, area, IoU with the crowd rule, bounding
  boxes — all computed on the run boundaries (interval sweeps in numpy), never through a dense mask.
* ``COCODataset``: the instances json (info, licences, categories, images, annotations) with id -> index maps, and the
  binary record format of ``COCOSeqFileGenerator`` (``COCOSerializeContext`` / ``COCODeserializer``: big-endian scalars,
  length-prefixed UTF-8 strings, a magic number closing every key).
"""
from __future__ import annotations

import json
import os
import struct
from dataclasses import dataclass, field

import numpy as np


class SegmentationMasks:
    """Base of the two mask encodings (MaskUtils.scala:23-35)."""

    def toRLE(self) -> "RLEMasks":
        raise NotImplementedError


class PolyMasks(SegmentationMasks):
    """A list of polygons ``[x0, y0, x1, y1, ...]`` describing ONE object (its parts may be disjoint)."""

    def __init__(self, poly, height: int, width: int):
        self.poly = [np.asarray(p, dtype=np.float32).reshape(-1) for p in poly]
        self.height, self.width = int(height), int(width)

    def toRLE(self) -> "RLEMasks":
        if self.height < 0 or self.width < 0:
            raise ValueError("the image size of this PolyMasks is not set")
        return MaskUtils.polyToSingleRLE(self, self.height, self.width)


class RLEMasks(SegmentationMasks):
    """Run lengths over the column-major flattened mask, starting with a run of zeros (COCO "uncompressed RLE")."""

    def __init__(self, counts, height: int, width: int):
        self.counts = np.asarray(counts, dtype=np.int64).reshape(-1)
        self.height, self.width = int(height), int(width)
        self._bbox = None
        self._area = None

    def toRLE(self) -> "RLEMasks":
        return self

    def get(self, i: int) -> int:
        return int(self.counts[i])

    def size(self) -> int:
        return int(self.counts.size)

    @property
    def bbox(self):
        """(x1, y1, x2, y2), inclusive pixel coordinates (MaskUtils.rleToOneBbox)."""
        if self._bbox is None:
            self._bbox = MaskUtils.rleToOneBbox(self)
        return self._bbox

    @property
    def area(self) -> int:
        if self._area is None:
            self._area = MaskUtils.rleArea(self)
        return self._area

    def __eq__(self, other):
        return (isinstance(other, RLEMasks) and self.height == other.height and self.width == other.width
                and np.array_equal(self.counts, other.counts))

    def __hash__(self):
        return hash((self.height, self.width, self.counts.tobytes()))

    def __repr__(self):
        return f"RLEMasks({self.height}x{self.width}, {self.counts.size} runs, area {self.area})"


def _intervals(rle: RLEMasks):
    """Half-open [start, end) index ranges of the runs of ones."""
    edges = np.cumsum(rle.counts)
    n = edges.size // 2
    return edges[0:2 * n:2], edges[1:2 * n:2]


def _from_intervals(starts, ends, height: int, width: int) -> RLEMasks:
    """Sorted, disjoint, non-touching [start, end) ranges -> counts (with the trailing run of zeros)."""
    total = height * width
    if len(starts) == 0:
        return RLEMasks([total], height, width)
    bounds = np.empty(2 * len(starts) + 1, dtype=np.int64)
    bounds[0:-1:2], bounds[1::2] = starts, ends
    bounds[-1] = total
    counts = np.diff(np.concatenate([[0], bounds]))
    if counts[-1] == 0:
        counts = counts[:-1]
    return RLEMasks(counts, height, width)


def _sweep(rles, need: int):
    """Ranges covered by at least ``need`` of the masks."""
    pos, delta = [], []
    for r in rles:
        s, e = _intervals(r)
        pos += [s, e]
        delta += [np.ones(s.size, dtype=np.int64), -np.ones(e.size, dtype=np.int64)]
    pos, delta = np.concatenate(pos), np.concatenate(delta)
    if pos.size == 0:
        return np.empty(0, np.int64), np.empty(0, np.int64)
    order = np.lexsort((-delta, pos))                 # at equal positions openings come first: touching runs merge
    pos, cover = pos[order], np.cumsum(delta[order])
    inside = cover >= need
    prev = np.concatenate([[False], inside[:-1]])
    starts, ends = pos[inside & ~prev], pos[~inside & prev]
    keep = ends > starts
    return starts[keep], ends[keep]


class MaskUtils:
    """Static helpers with the reference's names (MaskUtils.scala:125-580)."""

    @staticmethod
    def uint2long(i: int) -> int:
        return int(i) & 0xFFFFFFFF

    # ------------------------------------------------------------------ dense <-> RLE
    @staticmethod
    def binaryToRLE(mask) -> RLEMasks:
        mask = np.asarray(mask.detach().cpu().numpy() if hasattr(mask, "detach") else mask)
        h, w = mask.shape
        flat = (mask != 0).astype(np.uint8).T.reshape(-1)           # column-major
        change = np.flatnonzero(np.diff(flat)) + 1
        bounds = np.concatenate([[0], change, [flat.size]])
        counts = np.diff(bounds)
        if flat.size and flat[0] == 1:
            counts = np.concatenate([[0], counts])
        return RLEMasks(counts, h, w)

    @staticmethod
    def RLE2Binary(rle: RLEMasks) -> np.ndarray:
        vals = np.zeros(rle.counts.size, dtype=np.uint8)
        vals[1::2] = 1
        flat = np.repeat(vals, rle.counts)
        return flat.reshape(rle.width, rle.height).T.copy()

    # ------------------------------------------------------------------ COCO's compressed string form
    @staticmethod
    def RLE2String(rle: RLEMasks) -> str:
        """LEB128-like: 5 payload bits + continuation bit per char (offset 48); counts after the third are deltas to
        the count two places back (MaskUtils.scala:148-175, COCO rleToString)."""
        out = []
        cnts = rle.counts
        for i in range(cnts.size):
            x = int(cnts[i]) - (int(cnts[i - 2]) if i > 2 else 0)
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

    # ------------------------------------------------------------------ polygons
    @staticmethod
    def _one_poly(xy: np.ndarray, height: int, width: int) -> RLEMasks:
        scale = 5.0
        pts = np.asarray(xy, dtype=np.float64).reshape(-1, 2)
        x = np.floor(scale * pts[:, 0] + 0.5).astype(np.int64)
        y = np.floor(scale * pts[:, 1] + 0.5).astype(np.int64)
        x, y = np.append(x, x[0]), np.append(y, y[0])
        us, vs = [], []
        for j in range(len(pts)):                                     # dense integer walk along every edge
            xs, xe, ys, ye = x[j], x[j + 1], y[j], y[j + 1]
            dx, dy = abs(xe - xs), abs(ys - ye)
            flip = (dx >= dy and xs > xe) or (dx < dy and ys > ye)
            if flip:
                xs, xe, ys, ye = xe, xs, ye, ys
            if dx >= dy:
                t = np.arange(dx + 1, dtype=np.int64)
                t = dx - t if flip else t
                slope = (ye - ys) / dx if dx else 0.0
                us.append(t + xs)
                vs.append(np.floor(ys + slope * t + 0.5).astype(np.int64))
            else:
                t = np.arange(dy + 1, dtype=np.int64)
                t = dy - t if flip else t
                slope = (xe - xs) / dy
                vs.append(t + ys)
                us.append(np.floor(xs + slope * t + 0.5).astype(np.int64))
        u, v = np.concatenate(us), np.concatenate(vs)
        # points where the walk crosses a column boundary, back in pixel units
        j = np.flatnonzero(u[1:] != u[:-1]) + 1
        xd = np.where(u[j] < u[j - 1], u[j], u[j] - 1).astype(np.float64)
        xd = (xd + 0.5) / scale - 0.5
        ok = (np.floor(xd) == xd) & (xd >= 0) & (xd <= width - 1)
        yd = np.minimum(v[j], v[j - 1]).astype(np.float64)
        yd = np.ceil(np.clip((yd + 0.5) / scale - 0.5, 0, height))
        a = (xd[ok].astype(np.int64) * height + yd[ok].astype(np.int64))
        a = np.sort(np.append(a, height * width))
        runs = np.diff(np.concatenate([[0], a]))
        # a zero-length run joins its two neighbours
        out = [int(runs[0])]
        k = 1
        while k < runs.size:
            if runs[k] > 0:
                out.append(int(runs[k]))
                k += 1
            else:
                k += 1
                if k < runs.size:
                    out[-1] += int(runs[k])
                    k += 1
        return RLEMasks(out, height, width)

    @staticmethod
    def poly2RLE(poly, height: int, width: int):
        """One RLE per polygon of ``poly`` (a ``PolyMasks``, a list of polygons, or one flat polygon array)."""
        if isinstance(poly, PolyMasks):
            return [MaskUtils._one_poly(p, height, width) for p in poly.poly]
        arr = poly if isinstance(poly, (list, tuple)) and len(poly) and np.ndim(poly[0]) > 0 else [poly]
        res = [MaskUtils._one_poly(np.asarray(p), height, width) for p in arr]
        return res if arr is poly else res[0]

    @staticmethod
    def polyToSingleRLE(poly: PolyMasks, height: int, width: int) -> RLEMasks:
        return MaskUtils.mergeRLEs(MaskUtils.poly2RLE(poly, height, width), False)

    # ------------------------------------------------------------------ set operations on runs
    @staticmethod
    def mergeRLEs(rles, intersect: bool) -> RLEMasks:
        rles = list(rles)
        if len(rles) == 1:
            return rles[0]
        h, w = rles[0].height, rles[0].width
        for r in rles[1:]:
            if (r.height, r.width) != (h, w):
                raise ValueError("The height and width of the merged RLEs must be the same")
        s, e = _sweep(rles, len(rles) if intersect else 1)
        return _from_intervals(s, e, h, w)

    @staticmethod
    def rleArea(rle: RLEMasks) -> int:
        return int(rle.counts[1::2].sum())

    @staticmethod
    def rleIOU(detection: RLEMasks, groundTruth: RLEMasks, isCrowd: bool = False) -> float:
        """|D ∩ G| / |D ∪ G|; for a crowd ground truth the denominator is |D| (COCO's rule). The bounding boxes are
        compared first, like the reference, so disjoint masks cost two box computations."""
        if (detection.width, detection.height) != (groundTruth.width, groundTruth.height):
            raise ValueError("The sizes of RLEs must be the same to compute IOU")
        if MaskUtils.bboxIOU(groundTruth.bbox, detection.bbox, isCrowd) <= 0:
            return 0.0
        s, e = _sweep([detection, groundTruth], 2)
        inter = int((e - s).sum())
        if inter == 0:
            return 0.0
        union = detection.area if isCrowd else detection.area + groundTruth.area - inter
        return float(np.float32(inter) / np.float32(union))

    @staticmethod
    def bboxIOU(groundTruth, detection, isCrowd: bool = False, *rest) -> float:
        """bboxIOU((gx1,gy1,gx2,gy2), (dx1,dy1,dx2,dy2), isCrowd) or the 9-scalar form; inclusive pixel boxes."""
        if rest:
            vals = (groundTruth, detection, isCrowd) + rest
            gx1, gy1, gx2, gy2, dx1, dy1, dx2, dy2, isCrowd = vals
        else:
            (gx1, gy1, gx2, gy2), (dx1, dy1, dx2, dy2) = groundTruth, detection
        area = (gx2 - gx1 + 1) * (gy2 - gy1 + 1)
        iw = max(min(gx2, dx2) - max(gx1, dx1) + 1, 0)
        ih = max(min(gy2, dy2) - max(gy1, dy1) + 1, 0)
        inter = iw * ih
        det = (dx2 - dx1 + 1) * (dy2 - dy1 + 1)
        union = det if isCrowd else det + area - inter
        return float(inter / union) if union else 0.0

    @staticmethod
    def rleToOneBbox(rle: RLEMasks):
        s, e = _intervals(rle)
        if s.size == 0:
            return (0.0, 0.0, 0.0, 0.0)
        h = rle.height
        last = e - 1
        xs, xe = s // h, last // h
        ys, ye = s % h, last % h
        spans = xe > xs                                   # a run that continues into the next column covers full height
        y1 = 0 if spans.any() else int(ys.min())
        y2 = h - 1 if spans.any() else int(ye.max())
        return (float(xs.min()), float(y1), float(xe.max()), float(y2))


# ====================================================================== COCO record (de)serialisation
class COCOSerializeContext:
    """Byte sink of ``COCOSeqFileGenerator`` (COCODataset.scala:29-78): big-endian scalars, strings as int32 length +
    UTF-8 bytes, booleans as one byte."""

    def __init__(self):
        self._buf = bytearray()

    def dump(self, v):
        if isinstance(v, (bool, np.bool_)):
            self._buf += b"\x01" if v else b"\x00"
        elif isinstance(v, (int, np.integer)):
            self._buf += struct.pack(">i", int(v))
        elif isinstance(v, (float, np.floating)):
            self._buf += struct.pack(">f", float(v))
        elif isinstance(v, str):
            raw = v.encode("utf-8")
            self._buf += struct.pack(">i", len(raw)) + raw
        elif isinstance(v, (bytes, bytearray)):
            self._buf += bytes(v)
        else:
            raise TypeError(f"cannot serialise {type(v).__name__}")
        return self

    def dumpLong(self, v: int):
        self._buf += struct.pack(">q", int(v))
        return self

    def dumpDouble(self, v: float):
        self._buf += struct.pack(">d", float(v))
        return self

    def clear(self):
        self._buf.clear()

    def toByteArray(self) -> bytes:
        return bytes(self._buf)


@dataclass
class SimpleAnnotation:
    categoryId: int
    area: float
    bbox1: float
    bbox2: float
    bbox3: float
    bbox4: float
    isCrowd: bool
    masks: SegmentationMasks


class COCODeserializer:
    """Reader of the keys written through ``COCOSerializeContext`` (COCODataset.scala:80-135)."""

    def __init__(self, buffer: bytes):
        self._b, self._p = memoryview(bytes(buffer)), 0

    def _take(self, fmt: str):
        n = struct.calcsize(fmt)
        (v,) = struct.unpack_from(fmt, self._b, self._p)
        self._p += n
        return v

    def getInt(self) -> int: return self._take(">i")
    def getFloat(self) -> float: return self._take(">f")
    def getLong(self) -> int: return self._take(">q")
    def getDouble(self) -> float: return self._take(">d")
    def getBoolean(self) -> bool: return self._take(">b") != 0

    def getString(self) -> str:
        n = self.getInt()
        s = bytes(self._b[self._p:self._p + n]).decode("utf-8")
        self._p += n
        return s

    def getAnnotations(self):
        """(height, width, [SimpleAnnotation]) of one image."""
        height, width, n = self.getInt(), self.getInt(), self.getInt()
        return height, width, [self._annotation(height, width) for _ in range(n)]

    def _annotation(self, height: int, width: int) -> SimpleAnnotation:
        cat, area = self.getInt(), self.getFloat()
        b = [self.getFloat() for _ in range(4)]
        crowd = self.getBoolean()
        if crowd:
            masks = RLEMasks([self.getInt() for _ in range(self.getInt())], height, width)
        else:
            masks = PolyMasks([[self.getFloat() for _ in range(self.getInt())] for _ in range(self.getInt())],
                              height, width)
        return SimpleAnnotation(cat, area, b[0], b[1], b[2], b[3], crowd, masks)


class COCOSegmentation:
    """A mask that can write itself into a COCO record stream (COCODataset.scala:271-273)."""

    def dumpTo(self, context: "COCOSerializeContext"):  # pragma: no cover - interface
        raise NotImplementedError


class COCOPoly(PolyMasks, COCOSegmentation):
    def dumpTo(self, context: COCOSerializeContext):
        context.dump(len(self.poly))
        for p in self.poly:
            context.dump(int(p.size))
            for v in p:
                context.dump(float(v))


class COCORLE(RLEMasks, COCOSegmentation):
    def dumpTo(self, context: COCOSerializeContext):
        context.dump(int(self.counts.size))
        for c in self.counts:
            context.dump(int(c))


@dataclass
class COCOCategory:
    id: int
    name: str
    superCategory: str | None = None


@dataclass
class COCOLicence:
    id: int
    name: str
    url: str


@dataclass
class COCODatasetInfo:
    year: int = 0
    version: str = ""
    description: str = ""
    contributor: str = ""
    url: str = ""
    dateCreated: str = ""


@dataclass
class COCOAnotationOD:
    """One object annotation (the reference's spelling). bbox = (xmin, ymin, xmax, ymax), inclusive."""
    id: int
    imageId: int
    categoryId: int
    segmentation: SegmentationMasks | None
    area: float
    bbox: tuple
    isCrowd: bool
    image: "COCOImage | None" = None

    # names used by round-1 callers
    @property
    def image_id(self): return self.imageId
    @property
    def category_id(self): return self.categoryId
    @property
    def iscrowd(self): return self.isCrowd

    def dumpTo(self, context: COCOSerializeContext, dataset: "COCODataset"):
        context.dump(int(dataset.categoryId2Idx[self.categoryId]))
        context.dump(float(self.area))
        for v in self.bbox:
            context.dump(float(v))
        context.dump(bool(self.isCrowd))
        self.segmentation.dumpTo(context)


COCOAnnotation = COCOAnotationOD


@dataclass
class COCOImage:
    id: int
    height: int
    width: int
    file_name: str                      # full path (image root joined)
    license: int = 0
    flickrUrl: str | None = None
    cocoUrl: str | None = None
    dateCaptured: str | None = None
    fileName: str | None = None         # as written in the json
    annotations: list = field(default_factory=list)

    @property
    def path(self) -> str:
        return self.file_name

    @property
    def data(self) -> bytes:
        with open(self.file_name, "rb") as f:
            return f.read()

    def dumpTo(self, context: COCOSerializeContext, dataset: "COCODataset"):
        context.dump(int(self.height))
        context.dump(int(self.width))
        context.dump(len(self.annotations))
        for a in self.annotations:
            a.dumpTo(context, dataset)


class COCODataset:
    """``COCODataset.load(jsonPath, imageRoot)`` (COCODataset.scala:138-186, 298-352): images with their annotations
    attached, category id <-> contiguous 1-based index maps."""

    MAGIC_NUM = 0x1F3D4E5A

    def __init__(self, images, categories, info=None, licenses=None, annotations=None):
        self.info = info or COCODatasetInfo()
        self.images = list(images)
        self.licenses = list(licenses or [])
        self.categories = [c if isinstance(c, COCOCategory) else
                           COCOCategory(c["id"], c.get("name", ""), c.get("supercategory")) for c in categories]
        self.annotations = list(annotations) if annotations is not None else [a for im in self.images
                                                                              for a in im.annotations]
        self._by_id = {im.id: im for im in self.images}
        self.categoryId2Idx = {c.id: i + 1 for i, c in enumerate(self.categories)}
        self.categoryIdx2Id = {i: cid for cid, i in self.categoryId2Idx.items()}

    def getImageById(self, image_id: int) -> COCOImage:
        return self._by_id[image_id]

    def getCategoryByIdx(self, idx: int) -> COCOCategory:
        return self.categories[idx - 1]

    def imageKey(self, im: COCOImage) -> bytes:
        """The SequenceFile key of ``im``: file name, sizes, annotations, magic (COCOSeqFileGenerator.scala:97-102)."""
        ctx = COCOSerializeContext()
        ctx.dump(im.fileName if im.fileName is not None else os.path.basename(im.file_name))
        im.dumpTo(ctx, self)
        ctx.dump(COCODataset.MAGIC_NUM)
        return ctx.toByteArray()

    @staticmethod
    def parseKey(key: bytes):
        """Inverse of ``imageKey``: (file name, height, width, [SimpleAnnotation]); checks the magic number."""
        d = COCODeserializer(key)
        name = d.getString()
        h, w, anns = d.getAnnotations()
        if d.getInt() != COCODataset.MAGIC_NUM:
            raise ValueError("Corrupted metadata")
        return name, h, w, anns

    @staticmethod
    def load(json_path: str, image_root: str = ".") -> "COCODataset":
        with open(json_path) as f:
            d = json.load(f)
        images = {}
        for im in d.get("images", []):
            images[im["id"]] = COCOImage(im["id"], im["height"], im["width"], os.path.join(image_root, im["file_name"]),
                                         im.get("license", 0), im.get("flickr_url"), im.get("coco_url"),
                                         im.get("date_captured"), im["file_name"])
        annotations = []
        for a in d.get("annotations", []):
            im = images[a["image_id"]]
            crowd = int(a.get("iscrowd", 0)) == 1
            seg = a.get("segmentation")
            if isinstance(seg, dict):
                cnt, size = seg["counts"], seg.get("size", [im.height, im.width])
                if len(size) != 2:
                    raise ValueError("The size in the COCO annotation data should have 2 elements")
                rle = MaskUtils.string2RLE(cnt, size[0], size[1]) if isinstance(cnt, str) else None
                masks = COCORLE(rle.counts if rle is not None else cnt, size[0], size[1])
            elif isinstance(seg, list):
                masks = COCOPoly(seg, im.height, im.width)
            else:
                masks = None
            raw = a.get("bbox", ())
            if raw and len(raw) != 4:
                raise ValueError("The bbox in the COCO annotation data should have 4 elements")
            bbox = (raw[0], raw[1], raw[0] + raw[2] - 1, raw[1] + raw[3] - 1) if raw else ()
            ann = COCOAnotationOD(a["id"], a["image_id"], a["category_id"], masks, float(a.get("area", 0)), bbox, crowd,
                                  im)
            im.annotations.append(ann)
            annotations.append(ann)
        info = d.get("info") or {}
        lic = [COCOLicence(l.get("id", 0), l.get("name", ""), l.get("url", "")) for l in d.get("licenses", [])]
        cats = d.get("categories", [])         # index = position in the json + 1, like the reference (0 = background)
        return COCODataset(list(images.values()), cats,
                           COCODatasetInfo(info.get("year", 0), str(info.get("version", "")),
                                           info.get("description", ""), info.get("contributor", ""),
                                           info.get("url", ""), info.get("date_created", "")), lic, annotations)
