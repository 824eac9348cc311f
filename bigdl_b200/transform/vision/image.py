"""Vision 2.0 pipeline: ``ImageFeature`` (a keyed record travelling through the pipeline), ``ImageFrame`` (local or
rank-sharded collection) and ``FeatureTransformer``s chained with ``->`` / ``Pipeline``.

Reference: S/transform/vision/image/{ImageFeature,ImageFrame,FeatureTransformer,Convertor,MTImageFeatureToBatch}.scala,
S/transform/vision/image/augmentation/*.scala (19 augmentations), S/transform/vision/image/label/roi/*.scala, and the
Python mirror P/transform/vision/image.py whose class names and constructor arguments are kept. The reference runs every
op on an OpenCV ``Mat`` (uint8/float BGR, HWC); here the "mat" is a numpy HWC BGR array processed with cv2. A transformer
that fails on one image marks the feature invalid instead of aborting the frame, as FeatureTransformer.transform does
(FeatureTransformer.scala:78-100)."""
from __future__ import annotations

import logging
import math
import os
from typing import Iterable, Sequence

import numpy as np
import torch

from ...dataset.base import DataSet, MiniBatch, Sample
from ...utils.random_generator import RNG

log = logging.getLogger("bigdl_b200.transform")


class ImageFeature(dict):
    bytes, mat, floats, label, uri, originalSize = "bytes", "mat", "floats", "label", "uri", "originalSize"
    imageTensor, sample, predict, boundingBox, size, cropBbox, expandBbox = ("imageTensor", "sample", "predict",
                                                                              "boundingBox", "size", "cropBbox", "expandBbox")

    def __init__(self, image=None, label=None, path=None):
        super().__init__()
        self.isValid = True
        if isinstance(image, (bytes, bytearray)):
            self["bytes"] = bytes(image)
        elif image is not None:
            img = np.asarray(image)
            self["mat"] = img
            self["originalSize"] = img.shape
        if label is not None:
            self["label"] = np.asarray(label, dtype=np.float32)
        if path is not None:
            self["uri"] = path

    def get_image(self, float_key="floats", to_chw=True):
        img = self.get(float_key) if float_key in self else self.get("mat")
        img = np.asarray(img, dtype=np.float32)
        return img.transpose(2, 0, 1) if (to_chw and img.ndim == 3) else img

    def get_label(self):
        return self.get("label")

    def opencvMat(self) -> np.ndarray:
        return self["mat"]

    def getHeight(self): return self["mat"].shape[0]
    def getWidth(self): return self["mat"].shape[1]

    def getSize(self):
        """(height, width, channel) of the current mat."""
        m = self["mat"]
        return (m.shape[0], m.shape[1], 1 if m.ndim == 2 else m.shape[2])

    def getOriginalSize(self):
        return tuple(self["originalSize"]) if len(self["originalSize"]) == 3 else tuple(self["originalSize"]) + (1,)

    def getChannel(self): return self.getSize()[2]
    def getOriginalHeight(self): return self["originalSize"][0]
    def getOriginalWidth(self): return self["originalSize"][1]

    def clone(self):
        c = ImageFeature()
        for k, v in self.items():
            c[k] = v.copy() if isinstance(v, np.ndarray) else v
        c.isValid = self.isValid
        return c


class FeatureTransformer:
    ignoreException = True

    def transformMat(self, feature: ImageFeature):
        pass

    def transform(self, feature: ImageFeature) -> ImageFeature:
        if not feature.isValid:
            return feature
        try:
            self.transformMat(feature)
            m = feature.get("mat")
            if isinstance(m, np.ndarray) and m.ndim == 2:
                feature["mat"] = m[..., None]          # cv2 drops the channel axis of single-channel images
        except Exception as e:  # noqa: BLE001 - mirror of the reference's per-image isolation
            feature.isValid = False
            if not self.ignoreException:
                raise
            log.warning("failed in transformer %s for %s: %s", type(self).__name__, feature.get("uri"), e)
        return feature

    def __call__(self, data):
        if isinstance(data, ImageFrame):
            return data.transform(self)
        if isinstance(data, ImageFeature):
            return self.transform(data)
        return (self.transform(f) for f in data)

    def __rshift__(self, other: "FeatureTransformer") -> "Pipeline":
        return Pipeline([self, other])

    def setSkipException(self, skip: bool):
        self.ignoreException = skip
        return self


class Pipeline(FeatureTransformer):
    def __init__(self, transformers: Sequence[FeatureTransformer]):
        self.transformers = []
        for t in transformers:
            self.transformers.extend(t.transformers if isinstance(t, Pipeline) else [t])

    def transform(self, feature):
        for t in self.transformers:
            feature = t.transform(feature)
        return feature


class ChainedFeatureTransformer(Pipeline):
    """``first -> last`` (FeatureTransformer.scala:126-138)."""

    def __init__(self, first: FeatureTransformer, last: FeatureTransformer):
        super().__init__([first, last])
        self.first, self.last = first, last

    def enableIgnoreException(self):
        for t in self.transformers:
            if hasattr(t, "enableIgnoreException"):
                t.enableIgnoreException()
        return self


# ------------------------------------------------------------------------------------------------ frames
class ImageFrame:
    def __init__(self, features: Iterable[ImageFeature]):
        self.features = list(features)

    @staticmethod
    def read(path: str, sc=None, min_partitions: int = 1, distributed: bool = False) -> "ImageFrame":
        """Read one image file or every file of a folder (ImageFrame.read, ImageFrame.scala:80-135); ``sc`` is accepted
        for API parity — a non-None value selects the rank-sharded frame."""
        import glob

        if os.path.isfile(path):
            paths = [path]
        elif os.path.isdir(path):
            paths = sorted(os.path.join(path, f) for f in os.listdir(path) if os.path.isfile(os.path.join(path, f)))
        else:       # a glob pattern, or a path that matches nothing → an empty frame (ImageFrameSpec "read local")
            paths = sorted(p for p in glob.glob(path) if os.path.isfile(p))
        feats = []
        for p in paths:
            with open(p, "rb") as f:
                feats.append(ImageFeature(f.read(), path=p))
        cls = DistributedImageFrame if (distributed or sc is not None) else LocalImageFrame
        frame = cls(feats)
        return frame.transform(BytesToMat())

    @staticmethod
    def write_parquet(path: str, output: str, sc=None, partition_num: int = 1) -> None:
        """Pack the image files under ``path`` (a file, a folder or a glob) into parquet with the reference's two
        columns — ``uri`` (string) and ``bytes`` (binary) — split over ``partition_num`` files (ImageFrame.scala:170-178)."""
        import glob

        import pyarrow as pa
        import pyarrow.parquet as pq

        if os.path.isdir(path):
            paths = sorted(os.path.join(path, f) for f in os.listdir(path) if os.path.isfile(os.path.join(path, f)))
        else:
            paths = sorted(p for p in glob.glob(path) if os.path.isfile(p))
        os.makedirs(output, exist_ok=True)
        parts = max(1, min(int(partition_num), len(paths) or 1))
        for k in range(parts):
            chunk = paths[k::parts]
            data = []
            for p in chunk:
                with open(p, "rb") as f:
                    data.append(f.read())
            table = pa.table({"uri": pa.array(chunk, pa.string()), "bytes": pa.array(data, pa.binary())})
            pq.write_table(table, os.path.join(output, f"part-{k:05d}.parquet"))

    writeParquet = write_parquet

    @staticmethod
    def read_parquet(path: str, sc=None) -> "DistributedImageFrame":
        """Parquet written by ``write_parquet`` (or by the reference) -> rank-sharded frame of decoded images
        (ImageFrame.scala:153-161)."""
        import pyarrow.parquet as pq

        files = [path] if os.path.isfile(path) else sorted(
            os.path.join(path, f) for f in os.listdir(path) if f.endswith(".parquet"))
        feats = []
        for fn in files:
            t = pq.read_table(fn, columns=["uri", "bytes"]).to_pydict()
            feats += [ImageFeature(b, path=u) for u, b in zip(t["uri"], t["bytes"])]
        return DistributedImageFrame(feats).transform(BytesToMat())

    readParquet = read_parquet

    @staticmethod
    def array(images, labels=None) -> "LocalImageFrame":
        labels = labels if labels is not None else [None] * len(images)
        return LocalImageFrame([ImageFeature(i, l) for i, l in zip(images, labels)])

    def is_local(self): return isinstance(self, LocalImageFrame)
    def is_distributed(self): return isinstance(self, DistributedImageFrame)
    isLocal, isDistributed = is_local, is_distributed

    def transform(self, transformer: FeatureTransformer) -> "ImageFrame":
        self.features = [transformer.transform(f) for f in self.features]
        return self

    __rshift__ = transform

    def get_image(self, float_key="floats", to_chw=True):
        return [f.get_image(float_key, to_chw) for f in self.features]

    def get_label(self):
        return [f.get("label") for f in self.features]

    def get_predict(self, key="predict"):
        return [(f.get("uri"), f.get(key)) for f in self.features]

    def get_sample(self, key="sample"):
        return [f.get(key) for f in self.features]

    def get_uri(self, key="uri"):
        return [f.get(key) for f in self.features]

    def set_label(self, labels):
        for f, l in zip(self.features, labels):
            f["label"] = np.asarray(l, dtype=np.float32)
        return self

    def random_split(self, weights):
        idx = RNG.randperm(len(self.features))
        tot, out, pos = float(sum(weights)), [], 0
        for i, w in enumerate(weights):
            n = len(idx) - pos if i == len(weights) - 1 else int(round(len(idx) * w / tot))
            out.append(type(self)([self.features[j] for j in idx[pos:pos + n]]))
            pos += n
        return out

    def toDataSet(self, key="sample"):
        samples = [f[key] for f in self.features if f.isValid and key in f]
        return DataSet.array(samples, isinstance(self, DistributedImageFrame))

    def __len__(self):
        return len(self.features)

    def __iter__(self):
        return iter(self.features)


class LocalImageFrame(ImageFrame):
    def __init__(self, image_list=None, label_list=None):
        if image_list and not isinstance(image_list[0], ImageFeature):
            labels = label_list if label_list is not None else [None] * len(image_list)
            image_list = [ImageFeature(i, l) for i, l in zip(image_list, labels)]
        super().__init__(image_list or [])


class DistributedImageFrame(ImageFrame):
    """Each rank keeps ``features[rank::world]`` — the RDD partitioning of the reference becomes rank sharding."""

    def __init__(self, image_list=None, label_list=None):
        import torch.distributed as dist

        if image_list and not isinstance(image_list[0], ImageFeature):
            labels = label_list if label_list is not None else [None] * len(image_list)
            image_list = [ImageFeature(i, l) for i, l in zip(image_list, labels)]
        feats = list(image_list or [])
        if dist.is_initialized() and dist.get_world_size() > 1:
            feats = feats[dist.get_rank()::dist.get_world_size()]
        super().__init__(feats)


# ------------------------------------------------------------------------------------------------ decode / convert
class BytesToMat(FeatureTransformer):
    """Decode the encoded image bytes. ``image_codec`` is OpenCV's imread flag; the default (−1, "unchanged") keeps a
    grey image single-channel — the mat is then [H, W, 1] — exactly as the reference's
    ``BytesToMat(imageCodec = CV_LOAD_IMAGE_UNCHANGED)``; pass ``cv2.IMREAD_COLOR`` (1) to force three channels."""

    def __init__(self, byte_key="bytes", image_codec: int = -1):
        self.byte_key, self.image_codec = byte_key, image_codec

    def transformMat(self, f):
        import cv2

        img = cv2.imdecode(np.frombuffer(f[self.byte_key], dtype=np.uint8), self.image_codec)
        if img is None:
            raise ValueError("cannot decode image bytes")
        if img.ndim == 2:
            img = img[..., None]
        f["mat"] = img
        f["originalSize"] = img.shape


class PixelBytesToMat(FeatureTransformer):
    """Raw pixel bytes whose shape is in ``originalSize`` (PixelBytesToMat.scala)."""

    def __init__(self, byte_key="bytes"):
        self.byte_key = byte_key

    def transformMat(self, f):
        if self.byte_key not in f and f.get("mat") is not None:
            return              # SeqFileFolder.files_to_image_frame already decoded the record into the mat
        h, w, c = f["originalSize"]
        f["mat"] = np.frombuffer(f[self.byte_key], dtype=np.uint8).reshape(h, w, c).copy()


class MatToFloats(FeatureTransformer):
    def __init__(self, valid_height=300, valid_width=300, valid_channel=3, out_key="floats", share_buffer=True):
        self.out_key = out_key

    def transformMat(self, f):
        f[self.out_key] = np.asarray(f["mat"], dtype=np.float32)
        f["size"] = f["mat"].shape


class MatToTensor(FeatureTransformer):
    """mat → float tensor under ``tensor_key``; ``grey_to_rgb`` repeats a single-channel image three times
    (Convertor.scala ``MatToTensor(toRGB, tensorKey, shareBuffer, greyToRGB)``)."""

    def __init__(self, to_rgb=False, tensor_key="imageTensor", share_buffer=True, format="NCHW", grey_to_rgb=False,
                 greyToRGB=None):
        self.to_rgb, self.tensor_key, self.format = to_rgb, tensor_key, format
        self.grey_to_rgb = bool(grey_to_rgb if greyToRGB is None else greyToRGB)

    def transformMat(self, f):
        m = np.asarray(f["mat"], dtype=np.float32)
        if m.ndim == 2:
            m = m[..., None]
        if self.grey_to_rgb and m.shape[2] == 1:
            m = np.repeat(m, 3, axis=2)
        if self.to_rgb:
            m = m[..., ::-1]
        if self.format == "NCHW":
            m = m.transpose(2, 0, 1)
        f[self.tensor_key] = torch.from_numpy(np.ascontiguousarray(m))


class ImageFrameToSample(FeatureTransformer):
    def __init__(self, input_keys=("imageTensor",), target_keys=None, sample_key="sample"):
        self.input_keys, self.target_keys, self.sample_key = list(input_keys), list(target_keys or []), sample_key

    def transformMat(self, f):
        ins = [torch.as_tensor(f[k]) for k in self.input_keys]
        tg = [torch.as_tensor(f[k]) for k in self.target_keys if k in f]
        f[self.sample_key] = Sample(ins if len(ins) > 1 else ins[0], (tg if len(tg) > 1 else tg[0]) if tg else None)


# ------------------------------------------------------------------------------------------------ geometry
class Resize(FeatureTransformer):
    """Resize to (resize_h, resize_w); ``resize_mode`` is the cv2 interpolation flag, −1 picks one at random
    (Resize.scala:38-86)."""

    def __init__(self, resize_h, resize_w, resize_mode=1, use_scale_factor=True):
        self.h, self.w, self.mode = resize_h, resize_w, resize_mode

    def transformMat(self, f):
        import cv2

        modes = [cv2.INTER_LINEAR, cv2.INTER_CUBIC, cv2.INTER_AREA, cv2.INTER_NEAREST, cv2.INTER_LANCZOS4]
        mode = modes[int(RNG.uniform(0, len(modes)))] if self.mode == -1 else self.mode
        f["mat"] = cv2.resize(f["mat"], (self.w, self.h), interpolation=mode)


class AspectScale(FeatureTransformer):
    """Scale the shorter side to ``min_size`` keeping aspect, capped so the longer side ≤ ``max_size``; sizes rounded
    down to a multiple of ``scale_multiple_of`` (ScaleResize/AspectScale in Resize.scala:100-170)."""

    def __init__(self, min_size, scale_multiple_of=1, max_size=1000, resize_mode=1, use_scale_factor=True,
                 min_scale=-1.0):
        self.min_size, self.mult, self.max_size, self.mode, self.min_scale = (min_size, scale_multiple_of, max_size,
                                                                              resize_mode, min_scale)

    def _scale(self, h, w, size):
        short, long_ = min(h, w), max(h, w)
        s = size / short
        if self.min_scale > 0 and s < self.min_scale:
            s = self.min_scale
        if round(s * long_) > self.max_size:
            s = self.max_size / long_
        nh, nw = int(h * s + 0.5), int(w * s + 0.5)
        if self.mult > 1:
            nh, nw = nh // self.mult * self.mult, nw // self.mult * self.mult
        return max(nh, 1), max(nw, 1)

    @staticmethod
    def getHeightWidthAfterRatioScale(img, scaleTo: float, maxSize: int, scaleMultipleOf: int = 1, minScale=None):
        """(height, width) after the aspect-preserving scale (Resize.scala:117-141). ``img``: a mat / OpenCVMat."""
        m = np.asarray(getattr(img, "mat", img))
        t = AspectScale(scaleTo, scaleMultipleOf, maxSize, min_scale=-1.0 if minScale is None else float(minScale))
        return t._scale(m.shape[0], m.shape[1], scaleTo)

    def transformMat(self, f):
        import cv2

        h, w = f["mat"].shape[:2]
        nh, nw = self._scale(h, w, self.min_size)
        f["mat"] = cv2.resize(f["mat"], (nw, nh), interpolation=self.mode)


class RandomAspectScale(AspectScale):
    def __init__(self, scales, scale_multiple_of=1, max_size=1000):
        super().__init__(scales[0], scale_multiple_of, max_size)
        self.scales = list(scales)

    def transformMat(self, f):
        self.min_size = self.scales[int(RNG.uniform(0, len(self.scales)))]
        super().transformMat(f)


class RandomResize(FeatureTransformer):
    """Shorter side → U{min_size..max_size} (RandomResize.scala)."""

    def __init__(self, min_size, max_size):
        self.min_size, self.max_size = min_size, max_size

    def transformMat(self, f):
        import cv2

        h, w = f["mat"].shape[:2]
        size = int(RNG.uniform(self.min_size, self.max_size + 1))
        s = size / min(h, w)
        f["mat"] = cv2.resize(f["mat"], (max(1, int(round(w * s))), max(1, int(round(h * s)))))


class _Crop(FeatureTransformer):
    def __init__(self, normalized=True, is_clip=True):
        self.normalized, self.is_clip = normalized, is_clip

    def generateRoi(self, f):  # → (x1, y1, x2, y2)
        raise NotImplementedError

    def transformMat(self, f):
        h, w = f["mat"].shape[:2]
        x1, y1, x2, y2 = self.generateRoi(f)
        from .bbox import BoundingBox

        f["boundingBox"] = BoundingBox(x1, y1, x2, y2, self.normalized)      # what RoiProject maps the labels into
        if self.normalized:
            x1, x2, y1, y2 = x1 * w, x2 * w, y1 * h, y2 * h
        if self.is_clip:
            x1, y1, x2, y2 = (max(min(x1, w), 0), max(min(y1, h), 0), max(min(x2, w), 0), max(min(y2, h), 0))
        # origin and extent are truncated separately, like the reference's Rect(x1.toInt, y1.toInt, (x2-x1).toInt, …)
        ox, oy, cw, ch = int(x1), int(y1), int(x2 - x1), int(y2 - y1)
        x1, y1, x2, y2 = ox, oy, ox + cw, oy + ch
        f["mat"] = np.ascontiguousarray(f["mat"][y1:y2, x1:x2])
        f["cropBbox"] = (x1, y1, x2, y2)


Crop = _Crop          # the reference's public name of the abstract base (augmentation/Crop.scala:32)


class CenterCrop(_Crop):
    def __init__(self, crop_width, crop_height, is_clip=True):
        super().__init__(False, is_clip)
        self.cw, self.ch = crop_width, crop_height

    def generateRoi(self, f):
        h, w = f["mat"].shape[:2]
        x1, y1 = (w - self.cw) / 2.0, (h - self.ch) / 2.0          # kept fractional: the crop truncates, the label
        return x1, y1, x1 + self.cw, y1 + self.ch                  # projection uses the exact window (Crop.scala)


class RandomCrop(_Crop):
    def __init__(self, crop_width, crop_height, is_clip=True):
        super().__init__(False, is_clip)
        self.cw, self.ch = crop_width, crop_height

    def generateRoi(self, f):
        h, w = f["mat"].shape[:2]
        x1 = int(RNG.uniform(0, max(w - self.cw, 0) + 1))
        y1 = int(RNG.uniform(0, max(h - self.ch, 0) + 1))
        return x1, y1, x1 + self.cw, y1 + self.ch


class FixedCrop(_Crop):
    def __init__(self, x1, y1, x2, y2, normalized=True, is_clip=True):
        super().__init__(normalized, is_clip)
        self.roi = (x1, y1, x2, y2)

    def generateRoi(self, f):
        return self.roi


class DetectionCrop(_Crop):
    """Crop to the first box stored under ``roi_key`` (rows of [.., x1, y1, x2, y2] with the box in the last 4)."""

    def __init__(self, roi_key, normalized=True):
        super().__init__(normalized, True)
        self.roi_key = roi_key

    def generateRoi(self, f):
        r = np.asarray(f[self.roi_key], dtype=np.float32).reshape(-1)
        return tuple(r[-4:])


class RandomCropper(FeatureTransformer):
    """Random/center crop to crop_w×crop_h plus optional mirror (RandomCropper.scala:40-110)."""

    def __init__(self, crop_w, crop_h, mirror, cropper_method="Random", channels=3):
        self.cw, self.ch, self.mirror, self.method = crop_w, crop_h, mirror, str(cropper_method).lower()

    def transformMat(self, f):
        h, w = f["mat"].shape[:2]
        if self.method.startswith("random"):
            y1, x1 = int(RNG.uniform(0, h - self.ch + 1)), int(RNG.uniform(0, w - self.cw + 1))
        else:
            y1, x1 = (h - self.ch) // 2, (w - self.cw) // 2
        m = f["mat"][y1:y1 + self.ch, x1:x1 + self.cw]
        if self.mirror and RNG.uniform(0, 1) > 0.5:
            m = m[:, ::-1]
        f["mat"] = np.ascontiguousarray(m)


class RandomAlterAspect(FeatureTransformer):
    """Inception-style crop: random area fraction in [min_area_ratio, max_area_ratio] and aspect in
    [min_aspect_ratio_change, 1/…], then resize to crop_length² (RandomAlterAspect.scala:45-110)."""

    def __init__(self, min_area_ratio=0.08, max_area_ratio=1, min_aspect_ratio_change=0.75, interp_mode="CUBIC",
                 crop_length=224):
        self.min_area, self.max_area, self.min_aspect, self.crop = (min_area_ratio, max_area_ratio,
                                                                    min_aspect_ratio_change, crop_length)
        self.interp = interp_mode

    def transformMat(self, f):
        import cv2

        interp = {"CUBIC": cv2.INTER_CUBIC, "LINEAR": cv2.INTER_LINEAR, "AREA": cv2.INTER_AREA,
                  "NEAREST": cv2.INTER_NEAREST}.get(str(self.interp).upper(), cv2.INTER_CUBIC)
        h, w = f["mat"].shape[:2]
        area = h * w
        for _ in range(10):
            target = RNG.uniform(self.min_area, self.max_area) * area
            ar = RNG.uniform(self.min_aspect, 1.0 / self.min_aspect)
            nw, nh = int(round(math.sqrt(target * ar))), int(round(math.sqrt(target / ar)))
            if RNG.uniform(0, 1) < 0.5:
                nw, nh = nh, nw
            if 0 < nw <= w and 0 < nh <= h:
                y1, x1 = int(RNG.uniform(0, h - nh + 1)), int(RNG.uniform(0, w - nw + 1))
                f["mat"] = cv2.resize(f["mat"][y1:y1 + nh, x1:x1 + nw], (self.crop, self.crop), interpolation=interp)
                return
        s = self.crop / min(h, w)                      # fallback: scale + centre crop
        m = cv2.resize(f["mat"], (max(self.crop, int(round(w * s))), max(self.crop, int(round(h * s)))),
                       interpolation=interp)
        y1, x1 = (m.shape[0] - self.crop) // 2, (m.shape[1] - self.crop) // 2
        f["mat"] = np.ascontiguousarray(m[y1:y1 + self.crop, x1:x1 + self.crop])


class Expand(FeatureTransformer):
    """Place the image at a random position of a mean-filled canvas ``ratio`` ∈ [min,max] times larger (SSD)."""

    def __init__(self, means_r=123, means_g=117, means_b=104, min_expand_ratio=1.0, max_expand_ratio=4.0):
        self.means, self.lo, self.hi = (means_b, means_g, means_r), min_expand_ratio, max_expand_ratio

    def transformMat(self, f):
        h, w = f["mat"].shape[:2]
        ratio = RNG.uniform(self.lo, self.hi)
        nh, nw = int(h * ratio), int(w * ratio)
        top, left = int(RNG.uniform(0, nh - h + 1)), int(RNG.uniform(0, nw - w + 1))
        canvas = np.empty((nh, nw, 3), dtype=f["mat"].dtype)
        canvas[:] = np.asarray(self.means, dtype=f["mat"].dtype)
        canvas[top:top + h, left:left + w] = f["mat"]
        f["mat"] = canvas
        f["expandBbox"] = (-left / w, -top / h, (nw - left) / w, (nh - top) / h)
        from .bbox import BoundingBox

        f["boundingBox"] = BoundingBox(*f["expandBbox"], True)


class FixExpand(FeatureTransformer):
    def __init__(self, expand_height, expand_width):
        self.eh, self.ew = expand_height, expand_width

    def transformMat(self, f):
        h, w = f["mat"].shape[:2]
        top, left = (self.eh - h) // 2, (self.ew - w) // 2
        canvas = np.zeros((self.eh, self.ew, f["mat"].shape[2]), dtype=f["mat"].dtype)
        canvas[top:top + h, left:left + w] = f["mat"]
        f["mat"] = canvas
        f["expandBbox"] = (-left / w, -top / h, (self.ew - left) / w, (self.eh - top) / h)
        from .bbox import BoundingBox

        f["boundingBox"] = BoundingBox(*f["expandBbox"], True)


class HFlip(FeatureTransformer):
    def transformMat(self, f):
        f["mat"] = np.ascontiguousarray(f["mat"][:, ::-1])


class Filler(FeatureTransformer):
    """Fill the (normalised) rectangle with ``value`` (Filler.scala)."""

    def __init__(self, start_x, start_y, end_x, end_y, value=255):
        assert 0 <= start_x <= end_x <= 1 and 0 <= start_y <= end_y <= 1
        self.r, self.value = (start_x, start_y, end_x, end_y), value

    def transformMat(self, f):
        h, w = f["mat"].shape[:2]
        x1, y1, x2, y2 = self.r
        f["mat"][int(math.ceil(y1 * h)):int(math.ceil(y2 * h)), int(math.ceil(x1 * w)):int(math.ceil(x2 * w))] = self.value


# ------------------------------------------------------------------------------------------------ colour
def _as_float(f):
    m = f["mat"]
    return m.astype(np.float32) if m.dtype != np.float32 else m


def _restore(f, m, was_uint8):
    f["mat"] = np.clip(m, 0, 255).astype(np.uint8) if was_uint8 else m


class Brightness(FeatureTransformer):
    def __init__(self, delta_low, delta_high):
        self.lo, self.hi = delta_low, delta_high

    def transformMat(self, f):
        u8 = f["mat"].dtype == np.uint8
        _restore(f, _as_float(f) + RNG.uniform(self.lo, self.hi), u8)


class Contrast(FeatureTransformer):
    def __init__(self, delta_low, delta_high):
        self.lo, self.hi = delta_low, delta_high

    def transformMat(self, f):
        u8 = f["mat"].dtype == np.uint8
        _restore(f, _as_float(f) * RNG.uniform(self.lo, self.hi), u8)


class Saturation(FeatureTransformer):
    def __init__(self, delta_low, delta_high):
        self.lo, self.hi = delta_low, delta_high

    def transformMat(self, f):
        import cv2

        u8 = f["mat"].dtype == np.uint8
        hsv = cv2.cvtColor(np.clip(_as_float(f), 0, 255).astype(np.uint8), cv2.COLOR_BGR2HSV).astype(np.float32)
        hsv[..., 1] = np.clip(hsv[..., 1] * RNG.uniform(self.lo, self.hi), 0, 255)
        out = cv2.cvtColor(hsv.astype(np.uint8), cv2.COLOR_HSV2BGR)
        f["mat"] = out if u8 else out.astype(np.float32)


class Hue(FeatureTransformer):
    def __init__(self, delta_low, delta_high):
        self.lo, self.hi = delta_low, delta_high

    def transformMat(self, f):
        import cv2

        u8 = f["mat"].dtype == np.uint8
        hsv = cv2.cvtColor(np.clip(_as_float(f), 0, 255).astype(np.uint8), cv2.COLOR_BGR2HSV).astype(np.float32)
        hsv[..., 0] = (hsv[..., 0] + RNG.uniform(self.lo, self.hi) / 2.0) % 180.0      # cv2 hue is degrees/2
        out = cv2.cvtColor(hsv.astype(np.uint8), cv2.COLOR_HSV2BGR)
        f["mat"] = out if u8 else out.astype(np.float32)


class ChannelOrder(FeatureTransformer):
    """Random permutation of the colour channels."""

    def transformMat(self, f):
        f["mat"] = np.ascontiguousarray(f["mat"][..., RNG.randperm(f["mat"].shape[2])])


class RandomTransformer(FeatureTransformer):
    def __init__(self, transformer, prob):
        self.inner, self.prob = transformer, prob

    def transform(self, feature):
        if RNG.uniform(0, 1) < self.prob:
            return self.inner.transform(feature)
        return feature


class ColorJitter(FeatureTransformer):
    """SSD-style photometric distortion (ColorJitter.scala:45-120): brightness, then contrast/saturation/hue either
    before or after, each applied with its probability; ``shuffle`` randomises the order of all four."""

    def __init__(self, brightness_prob=0.5, brightness_delta=32.0, contrast_prob=0.5, contrast_lower=0.5,
                 contrast_upper=1.5, hue_prob=0.5, hue_delta=18.0, saturation_prob=0.5, saturation_lower=0.5,
                 saturation_upper=1.5, random_order_prob=0.0, shuffle=False):
        self.b = RandomTransformer(Brightness(-brightness_delta, brightness_delta), brightness_prob)
        self.c = RandomTransformer(Contrast(contrast_lower, contrast_upper), contrast_prob)
        self.s = RandomTransformer(Saturation(saturation_lower, saturation_upper), saturation_prob)
        self.h = RandomTransformer(Hue(-hue_delta, hue_delta), hue_prob)
        self.order = RandomTransformer(ChannelOrder(), random_order_prob)
        self.shuffle = shuffle

    def transform(self, feature):
        if not feature.isValid:
            return feature
        if self.shuffle:
            seq = [self.b, self.c, self.s, self.h]
            seq = [seq[i] for i in RNG.randperm(4)]
        elif RNG.uniform(0, 1) > 0.5:
            seq = [self.b, self.c, self.s, self.h]
        else:
            seq = [self.b, self.s, self.h, self.c]
        for t in seq + [self.order]:
            feature = t.transform(feature)
        return feature


class ChannelNormalize(FeatureTransformer):
    """(x − mean) / std per channel; constants in R,G,B order while the mat is BGR (ChannelNormalize.scala)."""

    def __init__(self, mean_r, mean_g, mean_b, std_r=1.0, std_g=1.0, std_b=1.0):
        self.mean = np.array([mean_b, mean_g, mean_r], dtype=np.float32)
        self.std = np.array([std_b, std_g, std_r], dtype=np.float32)

    def transformMat(self, f):
        f["mat"] = (_as_float(f) - self.mean) / self.std


class ChannelScaledNormalizer(FeatureTransformer):
    """(x − mean) * scale (ChannelScaledNormalizer.scala)."""

    def __init__(self, mean_r, mean_g, mean_b, scale):
        self.mean, self.scale = np.array([mean_b, mean_g, mean_r], dtype=np.float32), scale

    def transformMat(self, f):
        f["mat"] = (_as_float(f) - self.mean) * self.scale


class PixelNormalize(FeatureTransformer):
    def __init__(self, means):
        self.means = np.asarray(means, dtype=np.float32)

    def transformMat(self, f):
        m = _as_float(f)
        f["mat"] = m - self.means.reshape(m.shape)


PixelNormalizer = PixelNormalize


# ------------------------------------------------------------------------------------------------ ROI labels
class RoiLabel:
    """classes [N] (or [2,N] with difficulty) + bboxes [N,4] (label/roi/RoiLabel.scala)."""

    CLASSES, BBOXES, MASKS, ISCROWD, ORIGSIZE, SCORES = "classes", "bboxes", "masks", "is_crowd", "size", "scores"

    def __init__(self, classes, bboxes, masks=None):
        self.classes = np.asarray(classes, dtype=np.float32)
        self.bboxes = np.asarray(bboxes, dtype=np.float32).reshape(-1, 4)
        self.masks = masks
        n = self.classes.shape[-1] if self.classes.ndim else 0
        if self.classes.size and n != self.bboxes.shape[0]:
            raise ValueError(f"the number of classes {n} should be equal to the number of boxes {self.bboxes.shape[0]}")
        if masks is not None and self.classes.size and len(masks) != n:
            raise ValueError(f"the number of classes {n} should be equal to the number of masks {len(masks)}")

    def size(self):
        return self.bboxes.shape[0]

    def toTable(self):
        """{classes, bboxes[, masks as RLE]} (RoiLabel.toTable, RoiLabel.scala:57-65)."""
        from ...utils.table import T

        t = T()
        if self.masks is not None:
            t[RoiLabel.MASKS] = [m.toRLE() if hasattr(m, "toRLE") else m for m in self.masks]
        t[RoiLabel.CLASSES] = torch.from_numpy(self.classes)
        t[RoiLabel.BBOXES] = torch.from_numpy(self.bboxes)
        return t

    @staticmethod
    def fromTensor(tensor) -> "RoiLabel":
        """[N, 6] rows of (label, difficult, x1, y1, x2, y2) → RoiLabel with [2, N] classes."""
        a = np.asarray(tensor, dtype=np.float32).reshape(-1, 6)
        return RoiLabel(np.ascontiguousarray(a[:, :2].T), a[:, 2:6])

    @staticmethod
    def getScores(tab): return tab[RoiLabel.SCORES]
    @staticmethod
    def getClasses(tab): return tab[RoiLabel.CLASSES]
    @staticmethod
    def getBBoxes(tab): return tab[RoiLabel.BBOXES]
    @staticmethod
    def getMasks(tab): return tab[RoiLabel.MASKS]
    @staticmethod
    def getIsCrowd(tab): return tab[RoiLabel.ISCROWD]
    @staticmethod
    def getOrigSize(tab): return tab[RoiLabel.ORIGSIZE]


class RoiNormalize(FeatureTransformer):
    def transformMat(self, f):
        h, w = f["mat"].shape[:2]
        f["label"].bboxes = f["label"].bboxes / np.array([w, h, w, h], dtype=np.float32)


class RoiHFlip(FeatureTransformer):
    def __init__(self, normalized=True):
        self.normalized = normalized

    def transformMat(self, f):
        w = 1.0 if self.normalized else f["mat"].shape[1]
        b = f["label"].bboxes
        x1 = w - b[:, 2].copy()
        b[:, 2] = w - b[:, 0]
        b[:, 0] = x1


class RoiResize(FeatureTransformer):
    def __init__(self, normalized=True):
        self.normalized = normalized

    def transformMat(self, f):
        if self.normalized:
            return
        oh, ow = f["originalSize"][:2]
        h, w = f["mat"].shape[:2]
        f["label"].bboxes = f["label"].bboxes * np.array([w / ow, h / oh, w / ow, h / oh], dtype=np.float32)


class ScaleResize(FeatureTransformer):
    """Resize so that the shorter side becomes ``min_size`` unless that would push the longer side past ``max_size``
    (then the longer side becomes ``max_size``); with ``resize_roi`` the RoiLabel boxes follow
    (augmentation/ScaleResize.scala:41-90 — the Mask R-CNN evaluation resize)."""

    def __init__(self, min_size: int, max_size: int = -1, resize_roi: bool = False):
        self.min_size, self.max_size, self.resize_roi = min_size, max_size, resize_roi

    def _size(self, h: int, w: int):
        size = self.min_size
        if self.max_size > 0:
            lo, hi = (h, w) if w > h else (w, h)
            if hi / lo * size > self.max_size:
                size = int(round(self.max_size * lo / hi))
        if (w <= h and w == size) or (h <= w and h == size):
            return h, w
        return (size * h // w, size) if w < h else (size, size * w // h)

    def transformMat(self, f):
        import cv2

        h, w = f["mat"].shape[:2]
        nh, nw = self._size(h, w)
        f["mat"] = cv2.resize(f["mat"], (nw, nh), interpolation=cv2.INTER_LINEAR)
        lab = f.get("label") if hasattr(f, "get") else None
        if self.resize_roi and isinstance(lab, RoiLabel):
            oh, ow = (f.get("originalSize") or (h, w))[:2]
            lab.bboxes = lab.bboxes * np.array([nw / ow, nh / oh, nw / ow, nh / oh], dtype=np.float32)


class RoiProject(FeatureTransformer):
    """Project normalised boxes into the crop/expand window stored by the preceding transformer and drop boxes that
    fall outside (RoiProject.scala; ``need_meet_center_constraint`` keeps only boxes whose centre is inside)."""

    def __init__(self, need_meet_center_constraint=True):
        self.center = need_meet_center_constraint

    def transformMat(self, f):
        bb = f.get("boundingBox")
        if bb is not None:
            x1, y1, x2, y2 = bb.toTuple()
            if not bb.normalized:            # a pixel window is expressed relative to the CURRENT image (RoiTransformer.scala)
                h, w = f["mat"].shape[:2]
                x1, y1, x2, y2 = x1 / w, y1 / h, x2 / w, y2 / h
        else:
            win = f.get("expandBbox") or f.get("cropBboxNormalized")
            if win is None:
                return
            x1, y1, x2, y2 = win
        ww, wh = x2 - x1, y2 - y1
        lab = f["label"]
        b = lab.bboxes
        keep = np.ones(len(b), dtype=bool)
        if self.center:
            cx, cy = (b[:, 0] + b[:, 2]) / 2, (b[:, 1] + b[:, 3]) / 2
            keep = (cx >= x1) & (cx <= x2) & (cy >= y1) & (cy <= y2)
        keep &= ~((b[:, 0] >= x2) | (b[:, 2] <= x1) | (b[:, 1] >= y2) | (b[:, 3] <= y1))        # must overlap the window
        nb = np.stack([(b[:, 0] - x1) / ww, (b[:, 1] - y1) / wh, (b[:, 2] - x1) / ww, (b[:, 3] - y1) / wh], 1)
        nb = np.clip(nb, 0, 1).astype(np.float32)
        keep &= (nb[:, 2] - nb[:, 0]) * (nb[:, 3] - nb[:, 1]) > 0
        lab.bboxes = nb[keep]
        lab.classes = lab.classes[..., keep]
        if lab.masks is not None and len(lab.masks) == len(keep):
            lab.masks = [m for m, k in zip(lab.masks, keep) if k]


class RandomSampler(_Crop):
    """SSD data-augmentation crop (label/roi/RandomSampler.scala:31-69): each of seven ``BatchSampler``s — the whole
    image, five with a minimum Jaccard overlap of 0.1 / 0.3 / 0.5 / 0.7 / 0.9 against some ground-truth box, and one
    unconstrained — proposes at most one box; one proposal is drawn uniformly and the image is cropped to it, then the
    labels are projected into the crop (the reference's ``RandomSampler() -> RoiProject()``)."""

    def __init__(self, max_trials: int = 50):
        super().__init__(True, True)
        from .bbox import BatchSampler

        kw = dict(minScale=0.3, minAspectRatio=0.5, maxAspectRatio=2.0, maxTrials=max_trials)
        self.batchSamplers = [BatchSampler(maxTrials=1)] + [BatchSampler(minOverlap=o, **kw)
                                                            for o in (0.1, 0.3, 0.5, 0.7, 0.9)]
        self.batchSamplers.append(BatchSampler(maxOverlap=1.0, **kw))

    def generateRoi(self, f):
        from .bbox import BatchSampler, BoundingBox

        boxes: list = []
        BatchSampler.generateBatchSamples(f["label"], self.batchSamplers, boxes)
        if boxes:
            return boxes[int(RNG.uniform(0, 1) * len(boxes))]
        return BoundingBox(0, 0, 1, 1)

    def transformMat(self, f):
        lab = f.get("label")
        if lab is None or not isinstance(lab, RoiLabel):
            return
        box = self.generateRoi(f)
        FixedCrop(box.x1, box.y1, box.x2, box.y2, normalized=True).transformMat(f)
        f["cropBboxNormalized"] = box.toTuple()
        f["boundingBox"] = box
        RoiProject(True).transformMat(f)


# ------------------------------------------------------------------------------------------------ batching
class MTImageFeatureToBatch:
    """ImageFeatures → MiniBatches of [N,3,H,W] (MTImageFeatureToBatch.scala:40-346): the optional ``transformer`` is
    applied per image by a thread pool, tensors are written into a reusable (optionally pinned) batch buffer."""

    def __init__(self, width, height, batch_size, transformer: FeatureTransformer | None = None, to_rgb=False,
                 extract_roi=False, pin: bool = False, threads: int | None = None):
        self.w, self.h, self.batch_size, self.transformer, self.to_rgb = width, height, batch_size, transformer, to_rgb
        self.extract_roi, self.pin = extract_roi, pin
        from ...utils.thread_pool import ThreadPool

        self.pool = ThreadPool(threads or min(8, os.cpu_count() or 1))

    def _one(self, f):
        if self.transformer is not None:
            f = self.transformer.transform(f.clone())     # keep the stored feature pristine for the next epoch
        return f

    def __call__(self, frame_or_iter):
        feats = frame_or_iter.features if isinstance(frame_or_iter, ImageFrame) else frame_or_iter
        buf = []
        for f in feats:
            buf.append(f)
            if len(buf) == self.batch_size:
                yield self._emit(buf)
                buf = []
        if buf:
            yield self._emit(buf)

    def _emit(self, feats):
        feats = [f for f in self.pool.invokeAndWait([lambda f=f: self._one(f) for f in feats]) if f.isValid]
        n = len(feats)
        out = torch.empty((n, 3, self.h, self.w), dtype=torch.float32)
        if self.pin and torch.cuda.is_available():
            out = out.pin_memory()
        labels = []
        for i, f in enumerate(feats):
            m = np.asarray(f["mat"], dtype=np.float32)
            if m.shape[:2] != (self.h, self.w):
                raise ValueError(f"image size {m.shape[:2]} does not match batch size {(self.h, self.w)}")
            if self.to_rgb:
                m = m[..., ::-1]
            out[i].copy_(torch.from_numpy(np.ascontiguousarray(m.transpose(2, 0, 1))))
            lab = f.get("label")
            labels.append(lab if not isinstance(lab, RoiLabel) else lab)
        if self.extract_roi:
            return self._roi_batch(out, feats)
        if all(l is not None for l in labels):
            return MiniBatch(out, torch.as_tensor(np.stack([np.asarray(l, dtype=np.float32).reshape(-1) for l in labels])).squeeze(-1))
        return MiniBatch(out, None)

    @staticmethod
    def _roi_batch(out, feats):
        labels, crowd, sizes = [], [], []
        for f in feats:
            lab = f.get("label")
            ic = f.get(RoiLabel.ISCROWD)
            ic = torch.zeros(lab.size()) if ic is None else torch.as_tensor(np.asarray(ic, dtype=np.float32))
            if ic.numel() != lab.size():
                raise ValueError("the number of detections in the ImageFeature's is_crowd should be equal to the "
                                 "number of detections in the RoiLabel")
            labels.append(lab)
            crowd.append(ic)
            sizes.append(tuple(f.get("originalSize") or np.asarray(f["mat"]).shape))
        return RoiMiniBatch(out, labels, crowd, sizes)


class ClassificationMTImageFeatureToBatch(MTImageFeatureToBatch):
    """The classification batcher the reference's factory returns for ``extractRoi = false``
    (MTImageFeatureToBatch.scala:141-168)."""

    def __init__(self, width, height, totalBatchSize, transformer=None, toRGB=True, **kw):
        super().__init__(width, height, totalBatchSize, transformer, toRGB, False, **kw)


class RoiMTImageFeatureToBatch(MTImageFeatureToBatch):
    """Fixed-size detection batcher → ``RoiMiniBatch`` (MTImageFeatureToBatch.scala:221-255)."""

    def __init__(self, width, height, totalBatchSize, transformer=None, toRGB=True, **kw):
        super().__init__(width, height, totalBatchSize, transformer, toRGB, True, **kw)


class RoiImageFeatureToBatchWithResize(MTImageFeatureToBatch):
    """Detection batcher for images of different sizes (MTImageFeatureToBatch.scala:268-346): the batch tensor takes the
    largest height and width of its images, rounded up to a multiple of ``sizeDivisible`` when that is positive; each
    image sits in the top-left corner and the rest is zero."""

    def __init__(self, sizeDivisible: int = -1, totalBatchSize: int = 1, transformer=None, toRGB: bool = False, **kw):
        super().__init__(0, 0, totalBatchSize, transformer, toRGB, True, **kw)
        self.size_divisible = sizeDivisible

    def _emit(self, feats):
        feats = [f for f in self.pool.invokeAndWait([lambda f=f: self._one(f) for f in feats]) if f.isValid]
        mats = [np.asarray(f["mat"], dtype=np.float32) for f in feats]
        h, w = max(m.shape[0] for m in mats), max(m.shape[1] for m in mats)
        d = self.size_divisible
        if d > 0:
            h, w = int(math.ceil(h / d) * d), int(math.ceil(w / d) * d)
        out = torch.zeros((len(mats), 3, h, w), dtype=torch.float32)
        if self.pin and torch.cuda.is_available():
            out = out.pin_memory()
        for i, m in enumerate(mats):
            if self.to_rgb:
                m = m[..., ::-1]
            out[i, :, :m.shape[0], :m.shape[1]].copy_(torch.from_numpy(np.ascontiguousarray(m.transpose(2, 0, 1))))
        return self._roi_batch(out, feats)


def MTImageFeatureToBatchWithResize(sizeDivisible: int = -1, batchSize: int = 1, transformer=None, toRGB: bool = False,
                                    **kw):
    """Factory of the reference's companion object (MTImageFeatureToBatch.scala:62-66)."""
    return RoiImageFeatureToBatchWithResize(sizeDivisible, batchSize, transformer, toRGB, **kw)


class RoiMiniBatch(MiniBatch):
    """A batch of images with one ``RoiLabel``, one is-crowd vector and one original (height, width, channel) per image
    (MTImageFeatureToBatch.scala:175-213). ``getTarget()`` is a table of per-image tables
    {classes, bboxes[, masks], is_crowd, size}."""

    def __init__(self, input, target, isCrowd, originalSizes):
        self.input, self.target, self.isCrowd, self.originalSizes = input, list(target), list(isCrowd), list(originalSizes)
        self.inputs, self.targets = [input], []

    def size(self):
        return self.input.shape[0]

    def getInput(self):
        return self.input

    def getTarget(self):
        from ...utils.table import T

        out = T()
        for i, (lab, crowd, size) in enumerate(zip(self.target, self.isCrowd, self.originalSizes)):
            t = lab.toTable()
            t[RoiLabel.ISCROWD] = crowd
            t[RoiLabel.ORIGSIZE] = size
            out[i + 1] = t
        return out

    def slice(self, offset: int, length: int):
        """1-based ``offset`` like every MiniBatch.slice."""
        o = offset - 1
        return RoiMiniBatch(self.input[o:o + length], self.target[o:o + length], self.isCrowd[o:o + length],
                            self.originalSizes[o:o + length])

    get_input, get_target = getInput, getTarget

    def set(self, samples):
        raise NotImplementedError("do not use Sample here")

    def pin(self):
        if torch.cuda.is_available():
            return RoiMiniBatch(self.input.pin_memory(), self.target, self.isCrowd, self.originalSizes)
        return self

    def to(self, device, non_blocking: bool = False):
        return RoiMiniBatch(self.input.to(device, non_blocking=non_blocking), self.target, self.isCrowd,
                            self.originalSizes)


class ImageFeatureToMiniBatch:
    """Batches the ``Sample`` stored in each ImageFeature under ``sampleKey`` (Convertor.scala:252-278)."""

    def __init__(self, batchSize: int, featurePaddingParam=None, labelPaddingParam=None, partitionNum=None,
                 sampleKey: str = "sample"):
        from ...dataset.base import SampleToMiniBatch

        self.toBatch = SampleToMiniBatch(batchSize, featurePaddingParam, labelPaddingParam, partitionNum)
        self.sampleKey = sampleKey

    def __call__(self, prev):
        feats = prev.features if isinstance(prev, ImageFrame) else prev
        return self.toBatch(f[self.sampleKey] for f in feats)


class SeqFileFolder:
    """``SeqFileFolder.files_to_image_frame(url, sc, class_num)`` (P/transform/vision/image.py:726-750) and the COCO
    form ``files_to_roi_image_frame`` (S/dataset/DataSet.scala:606-650)."""

    @staticmethod
    def files_to_roi_image_frame(url: str, sc=None, partition_num: int | None = None):
        """SequenceFiles written by ``cli.coco_seqfile_generator`` → ImageFrame whose features carry the decoded BGR
        pixels as raw ``bytes`` + ``originalSize`` (feed ``PixelBytesToMat`` or read ``mat``), ``uri``, a ``RoiLabel``
        (class indices, [x1, y1, x2, y2] boxes, masks) and the ``is_crowd`` vector. Grey images are expanded to three
        channels like the reference does."""
        import cv2

        from ...cli.coco_seqfile_generator import read_coco_seqfile
        from ...dataset.seqfile import seq_file_paths

        feats = []
        for fn in seq_file_paths(url):
            for name, h, w, anns, data in read_coco_seqfile(fn):
                img = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
                if img is None or img.shape[:2] != (h, w):
                    raise ValueError(f"{name}: image does not match the recorded size {(h, w)}")
                label = RoiLabel(np.array([a.categoryId for a in anns], dtype=np.float32),
                                 np.array([[a.bbox1, a.bbox2, a.bbox3, a.bbox4] for a in anns], dtype=np.float32).reshape(-1, 4),
                                 [a.masks for a in anns])
                f = ImageFeature()
                f["bytes"] = np.ascontiguousarray(img).tobytes()
                f["mat"] = img
                f["originalSize"] = (h, w, 3)
                f["label"] = label
                f["uri"] = name
                f[RoiLabel.ISCROWD] = np.array([1.0 if a.isCrowd else 0.0 for a in anns], dtype=np.float32)
                feats.append(f)
        return (DistributedImageFrame if sc is not None else LocalImageFrame)(feats)

    filesToRoiImageFrame = files_to_roi_image_frame

    @staticmethod
    def files_to_image_frame(url: str, sc=None, class_num: int = 1000, partition_num: int = -1):
        from ...dataset.seqfile import decode_image_value, read_label, read_name, read_sequence_file, seq_file_paths

        feats = []
        for fn in seq_file_paths(url):
            for k, v in read_sequence_file(fn):
                lab = float(read_label(k))
                if lab > class_num:
                    continue
                f = ImageFeature(decode_image_value(v).copy(), [lab])
                try:
                    f["uri"] = read_name(k)
                except ValueError:
                    pass
                feats.append(f)
        return (DistributedImageFrame if sc is not None else LocalImageFrame)(feats)
