"""Legacy image pipeline (S/dataset/image/*, 24 files): byte records → BGR / grey float images → crop / flip / colour
jitter / lighting / normalise → Sample or MiniBatch. Images are float32 HWC in BGR order with values 0..255 unless a
``normalize`` divisor is given, exactly as ``LabeledBGRImage`` (S/dataset/image/Types.scala:199-330).

The multi-threaded batcher ``MTLabeledBGRImgToBatch`` (S/dataset/image/MTLabeledBGRImgToBatch.scala:60-130) runs the
crop+flip+normalise+CHW conversion in the C++ host library (csrc/host/batcher.cpp) straight into a pinned buffer."""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass

import numpy as np
import torch

from ..utils.random_generator import RNG
from .base import MiniBatch, Sample, Transformer

CropRandom, CropCenter = "random", "center"


class CropperMethod:
    """The two crop policies of BGRImgCropper (S/dataset/image/BGRImgCropper.scala:26-36)."""
    CropRandom, CropCenter = CropRandom, CropCenter


@dataclass
class ByteRecord:
    data: bytes
    label: float


@dataclass
class LocalLabeledImagePath:
    label: float
    path: str


class LabeledBGRImage:
    NO_SCALE = -1

    def __init__(self, data: np.ndarray | None = None, label: float = 0.0):
        self.data = data          # float32 [H, W, 3], BGR
        self._label = label

    def copy(self, src, normalize: float = 255.0):
        """Fill from a byte record (8-byte BE width/height header + BGR bytes) or from another image
        (BGRImage.copy, Types.scala:137-210)."""
        if isinstance(src, (bytes, bytearray, memoryview)):
            w, h = struct.unpack_from(">ii", src, 0)
            raw = np.frombuffer(src, dtype=np.uint8, offset=8, count=w * h * 3).reshape(h, w, 3)
            self.data = raw.astype(np.float32) / normalize
        else:
            self.data = src.data.copy()
            if hasattr(src, "_label"):
                self._label = src._label
        return self

    def save(self, path: str, scale: float = 255.0):
        """Write the image file; values are multiplied by ``scale`` first (BGRImage.save, Types.scala:174-190)."""
        import cv2

        cv2.imwrite(path, np.clip(self.data * scale, 0, 255).astype(np.uint8))
        return self

    @staticmethod
    def readImage(path: str, scale_to_or_w: int = -1, resize_h: int | None = None) -> bytes:
        """Decode → byte record with the width/height header, shorter side scaled to ``scale_to`` or resized to
        w×h (object BGRImage.readImage, Types.scala:371-384)."""
        if resize_h is None:
            img = read_image_bgr(path, scale_to_or_w if scale_to_or_w > 0 else None)
        else:
            import cv2

            img = cv2.resize(read_image_bgr(path), (scale_to_or_w, resize_h), interpolation=cv2.INTER_LINEAR)
        return struct.pack(">ii", img.shape[1], img.shape[0]) + np.ascontiguousarray(img).tobytes()

    def width(self) -> int:
        return self.data.shape[1]

    def height(self) -> int:
        return self.data.shape[0]

    def label(self) -> float:
        return self._label

    def setLabel(self, v: float):
        self._label = v
        return self

    def content(self) -> np.ndarray:
        return self.data

    def hflip(self):
        self.data = np.ascontiguousarray(self.data[:, ::-1])
        return self

    def convertToByte(self, scale: float = 255.0) -> np.ndarray:
        return np.clip(self.data * (255.0 / scale), 0, 255).astype(np.uint8)

    def copyTo(self, dst: torch.Tensor, to_rgb: bool = False):
        """Write CHW float into ``dst`` ([3,H,W]); ``to_rgb`` swaps channel order (BGRImage.copyTo)."""
        d = self.data[..., ::-1] if to_rgb else self.data
        dst.copy_(torch.from_numpy(np.ascontiguousarray(d.transpose(2, 0, 1))))
        return dst

    def clone(self):
        return LabeledBGRImage(self.data.copy(), self._label)


class LabeledGreyImage:
    def __init__(self, data: np.ndarray | None = None, label: float = 0.0):
        self.data, self._label = data, label      # float32 [H, W]

    def copy(self, src, normalize: float = 1.0, offset: int = 0, width: int | None = None, height: int | None = None):
        """From raw bytes (``width``×``height`` taken from the current shape when not given) or another image
        (GreyImage.copy, Types.scala:57-90)."""
        if isinstance(src, (bytes, bytearray, memoryview)):
            h = height if height is not None else self.data.shape[0]
            w = width if width is not None else self.data.shape[1]
            raw = np.frombuffer(src, dtype=np.uint8, offset=offset, count=w * h).reshape(h, w)
            self.data = raw.astype(np.float32) / normalize
        else:
            self.data = src.data.copy()
            if hasattr(src, "_label"):
                self._label = src._label
        return self

    def setLabel(self, v: float):
        self._label = v
        return self

    def clone(self):
        return LabeledGreyImage(self.data.copy(), self._label)

    def width(self): return self.data.shape[1]
    def height(self): return self.data.shape[0]
    def label(self): return self._label
    def content(self): return self.data


# ------------------------------------------------------------------------------------------------ readers
def read_image_bgr(path: str, scale_to: int | None = None) -> np.ndarray:
    """Decode to uint8 BGR; ``scale_to`` resizes the SHORTER side to that length (BGRImage.readImage, Types.scala)."""
    import cv2

    img = cv2.imread(path, cv2.IMREAD_COLOR)
    if img is None:
        raise IOError(f"cannot decode image {path}")
    if scale_to is not None and scale_to > 0:
        h, w = img.shape[:2]
        if w < h:
            nw, nh = scale_to, max(1, int(round(scale_to * h / w)))
        else:
            nh, nw = scale_to, max(1, int(round(scale_to * w / h)))
        img = cv2.resize(img, (nw, nh), interpolation=cv2.INTER_LINEAR)
    return img


class LocalImageFiles:
    """Folder-per-class listing; labels are 1-based in sorted folder order (LocalImageFiles.scala:80-140)."""

    @staticmethod
    def readLabels(path: str) -> dict[str, int]:
        dirs = sorted(d for d in os.listdir(path) if os.path.isdir(os.path.join(path, d)))
        return {d: i + 1 for i, d in enumerate(dirs)}

    @staticmethod
    def readPaths(path: str, has_label: bool = True) -> list[LocalLabeledImagePath]:
        out = []
        if has_label:
            for d, lab in LocalImageFiles.readLabels(path).items():
                for fn in sorted(os.listdir(os.path.join(path, d))):
                    out.append(LocalLabeledImagePath(float(lab), os.path.join(path, d, fn)))
        else:
            for fn in sorted(os.listdir(path)):
                out.append(LocalLabeledImagePath(-1.0, os.path.join(path, fn)))
        return out


class LocalImgReader(Transformer):
    """path → LabeledBGRImage (LocalImgReader.scala): scale shorter side to ``scale_to`` (or resize to w×h),
    divide by ``normalize``."""

    def __init__(self, scale_to: int = 256, resize_w: int | None = None, resize_h: int | None = None,
                 normalize: float = 255.0):
        self.scale_to, self.resize_w, self.resize_h, self.normalize = scale_to, resize_w, resize_h, normalize

    def _read(self, p: LocalLabeledImagePath) -> LabeledBGRImage:
        import cv2

        if self.resize_w and self.resize_h:
            img = cv2.resize(read_image_bgr(p.path), (self.resize_w, self.resize_h))
        else:
            img = read_image_bgr(p.path, self.scale_to)
        return LabeledBGRImage(img.astype(np.float32) / self.normalize, p.label)

    def __call__(self, it):
        for p in it:
            yield self._read(p)


class LocalImgReaderWithName(LocalImgReader):
    def __call__(self, it):
        for p in it:
            yield self._read(p), os.path.basename(p.path)


class LocalScaleImgReader(LocalImgReader):
    """Shorter side → ``scaleTo`` (LocalImgReader.scala:47-60)."""

    def __init__(self, scaleTo: int = -1, normalize: float = 255.0):
        super().__init__(scale_to=scaleTo, normalize=normalize)


class LocalResizeImgReader(LocalImgReader):
    """Resize to exactly ``resizeW``×``resizeH`` (LocalImgReader.scala:70-83)."""

    def __init__(self, resizeW: int, resizeH: int, normalize: float = 255.0):
        super().__init__(resize_w=resizeW, resize_h=resizeH, normalize=normalize)


class LocalScaleImgWithName(LocalImgReaderWithName):
    def __init__(self, scaleTo: int = -1, normalize: float = 255.0):
        super().__init__(scale_to=scaleTo, normalize=normalize)


class LocalResizeImgWithName(LocalImgReaderWithName):
    def __init__(self, resizeW: int, resizeH: int, normalize: float = 255.0):
        super().__init__(resize_w=resizeW, resize_h=resizeH, normalize=normalize)


# unlabeled base names of the reference's class hierarchy (Types.scala:51,126): same behaviour, label defaults to 0
BGRImage = LabeledBGRImage
GreyImage = LabeledGreyImage


@dataclass
class LocalSeqFilePath:
    """A sequence-file path on the local disk (S/dataset/Types.scala:53)."""
    path: str

    def __fspath__(self):
        return str(self.path)

    def __str__(self):
        return str(self.path)


@dataclass
class LocalImagePath:
    """An image file without a label (S/dataset/Types.scala:46)."""
    path: str


class BytesToBGRImg(Transformer):
    """ByteRecord (8-byte BE width/height header + BGR bytes) → LabeledBGRImage (BytesToBGRImg.scala:33-60)."""

    def __init__(self, normalize: float = 255.0, resize_w: int = -1, resize_h: int = -1):
        self.normalize, self.resize_w, self.resize_h = normalize, resize_w, resize_h

    def __call__(self, it):
        for rec in it:
            w, h = struct.unpack_from(">ii", rec.data, 0)
            img = np.frombuffer(rec.data, dtype=np.uint8, offset=8, count=w * h * 3).reshape(h, w, 3)
            if self.resize_w > 0 and self.resize_h > 0:
                import cv2

                img = cv2.resize(img, (self.resize_w, self.resize_h))
            yield LabeledBGRImage(img.astype(np.float32) / self.normalize, rec.label)


class BytesToGreyImg(Transformer):
    def __init__(self, row: int, col: int):
        self.row, self.col = row, col

    def __call__(self, it):
        for rec in it:
            img = np.frombuffer(rec.data, dtype=np.uint8, count=self.row * self.col).reshape(self.row, self.col)
            yield LabeledGreyImage(img.astype(np.float32) / 255.0, rec.label)


# ------------------------------------------------------------------------------------------------ augmentations
def _crop_offsets(h, w, ch, cw, method):
    if method == CropRandom:
        return int(RNG.uniform(0, h - ch + 1)) if h > ch else 0, int(RNG.uniform(0, w - cw + 1)) if w > cw else 0
    return (h - ch) // 2, (w - cw) // 2


class BGRImgCropper(Transformer):
    def __init__(self, crop_width: int, crop_height: int, crop_method: str = CropRandom):
        self.cw, self.ch, self.method = crop_width, crop_height, crop_method

    def __call__(self, it):
        for img in it:
            oy, ox = _crop_offsets(img.height(), img.width(), self.ch, self.cw, self.method)
            img.data = np.ascontiguousarray(img.data[oy:oy + self.ch, ox:ox + self.cw])
            yield img


class BGRImgRdmCropper(Transformer):
    """Zero-pad by ``padding`` then take a random crop (BGRImgRdmCropper.scala — CIFAR-style augmentation)."""

    def __init__(self, crop_width: int, crop_height: int, padding: int):
        self.cw, self.ch, self.pad = crop_width, crop_height, padding

    def __call__(self, it):
        for img in it:
            p = self.pad
            d = np.pad(img.data, ((p, p), (p, p), (0, 0))) if p else img.data
            oy, ox = _crop_offsets(d.shape[0], d.shape[1], self.ch, self.cw, CropRandom)
            img.data = np.ascontiguousarray(d[oy:oy + self.ch, ox:ox + self.cw])
            yield img


class GreyImgCropper(BGRImgCropper):
    pass


class HFlip(Transformer):
    def __init__(self, threshold: float = 0.0):
        self.threshold = threshold

    def __call__(self, it):
        for img in it:
            if RNG.uniform(0, 1) >= self.threshold:
                img.hflip()
            yield img


class BGRImgNormalizer(Transformer):
    """Per-channel (x − mean) / std with means given in R,G,B order (BGRImgNormalizer.scala:30-110). ``from_dataset``
    computes the statistics over a dataset like the companion ``apply(dataSet)``."""

    def __init__(self, mean_r, mean_g, mean_b, std_r=1.0, std_g=1.0, std_b=1.0):
        self.mean = np.array([mean_b, mean_g, mean_r], dtype=np.float32)
        self.std = np.array([std_b, std_g, std_r], dtype=np.float32)

    @staticmethod
    def from_dataset(images, samples: int = -1):
        s, sq, n = np.zeros(3), np.zeros(3), 0
        for i, img in enumerate(images):
            if 0 <= samples <= i:
                break
            d = img.data.reshape(-1, 3).astype(np.float64)
            s += d.sum(0)
            sq += (d * d).sum(0)
            n += d.shape[0]
        mean = s / n
        std = np.sqrt(np.maximum(sq / n - mean * mean, 0))
        return BGRImgNormalizer(mean[2], mean[1], mean[0], std[2], std[1], std[0])

    def getMean(self):
        return float(self.mean[2]), float(self.mean[1]), float(self.mean[0])

    def getStd(self):
        return float(self.std[2]), float(self.std[1]), float(self.std[0])

    def __call__(self, it):
        for img in it:
            img.data = (img.data - self.mean) / self.std
            yield img


class GreyImgNormalizer(Transformer):
    def __init__(self, mean: float, std: float):
        self.mean, self.std = mean, std

    def __call__(self, it):
        for img in it:
            img.data = (img.data - self.mean) / self.std
            yield img


class BGRImgPixelNormalizer(Transformer):
    """Subtract a per-pixel mean image (BGRImgPixelNormalizer.scala)."""

    def __init__(self, means: np.ndarray):
        self.means = np.asarray(means, dtype=np.float32)

    def __call__(self, it):
        for img in it:
            img.data = img.data - self.means.reshape(img.data.shape)
            yield img


class ColorJitter(Transformer):
    """Random order of brightness / contrast / saturation jitter, each blending with strength U(−0.4, 0.4)
    (ColorJitter.scala:31-120, Torch fb.resnet recipe)."""

    def __init__(self, brightness: float = 0.4, contrast: float = 0.4, saturation: float = 0.4):
        self.b, self.c, self.s = brightness, contrast, saturation

    @staticmethod
    def _grey(d):
        return (0.114 * d[..., 0] + 0.587 * d[..., 1] + 0.299 * d[..., 2])[..., None]

    def _blend(self, d, other, var):
        alpha = 1.0 + RNG.uniform(-var, var)
        return d * alpha + other * (1 - alpha)

    def __call__(self, it):
        for img in it:
            order = RNG.shuffle([0, 1, 2])
            d = img.data
            for o in order:
                if o == 0:
                    d = self._blend(d, np.zeros_like(d), self.b)
                elif o == 1:
                    d = self._blend(d, np.full_like(d, self._grey(d).mean()), self.c)
                else:
                    d = self._blend(d, np.repeat(self._grey(d), 3, axis=2), self.s)
            img.data = d.astype(np.float32)
            yield img


class Lighting(Transformer):
    """AlexNet-style PCA lighting noise (Lighting.scala:30-80): add Σ_k α_k λ_k v_k with α ~ N(0, 0.1), RGB eigen
    basis of ImageNet."""

    alphastd = 0.1
    eigval = np.array([0.2175, 0.0188, 0.0045], dtype=np.float32)
    eigvec = np.array([[-0.5675, 0.7192, 0.4009], [-0.5808, -0.0045, -0.8140], [-0.5836, -0.6948, 0.4203]],
                      dtype=np.float32)

    def __call__(self, it):
        for img in it:
            alpha = np.array([RNG.normal(0, self.alphastd) for _ in range(3)], dtype=np.float32)
            rgb = (self.eigvec * (alpha * self.eigval)[None, :]).sum(1)
            img.data = img.data + rgb[::-1][None, None, :]        # data is BGR
            yield img


# ------------------------------------------------------------------------------------------------ to Sample / batch
class BGRImgToSample(Transformer):
    def __init__(self, to_rgb: bool = True):
        self.to_rgb = to_rgb

    def __call__(self, it):
        for img in it:
            d = img.data[..., ::-1] if self.to_rgb else img.data
            yield Sample(torch.from_numpy(np.ascontiguousarray(d.transpose(2, 0, 1))), torch.tensor([img.label()]))


class GreyImgToSample(Transformer):
    def __call__(self, it):
        for img in it:
            yield Sample(torch.from_numpy(np.ascontiguousarray(img.data)), torch.tensor([img.label()]))


class BGRImgToBatch(Transformer):
    def __init__(self, batch_size: int, to_rgb: bool = True):
        self.batch_size, self.to_rgb = batch_size, to_rgb

    def __call__(self, it):
        feats, labels = [], []
        for img in it:
            d = img.data[..., ::-1] if self.to_rgb else img.data
            feats.append(torch.from_numpy(np.ascontiguousarray(d.transpose(2, 0, 1))))
            labels.append(img.label())
            if len(feats) == self.batch_size:
                yield MiniBatch(torch.stack(feats), torch.tensor(labels))
                feats, labels = [], []
        if feats:
            yield MiniBatch(torch.stack(feats), torch.tensor(labels))


class GreyImgToBatch(Transformer):
    def __init__(self, batch_size: int):
        self.batch_size = batch_size

    def __call__(self, it):
        feats, labels = [], []
        for img in it:
            feats.append(torch.from_numpy(np.ascontiguousarray(img.data)))
            labels.append(img.label())
            if len(feats) == self.batch_size:
                yield MiniBatch(torch.stack(feats), torch.tensor(labels))
                feats, labels = [], []
        if feats:
            yield MiniBatch(torch.stack(feats), torch.tensor(labels))


class BGRImgToImageVector(Transformer):
    """Flatten to a dense feature vector for the ML-pipeline API (BGRImgToImageVector.scala)."""

    def __call__(self, it):
        for img in it:
            yield np.ascontiguousarray(img.data.transpose(2, 0, 1)).reshape(-1), img.label()


class MTLabeledBGRImgToBatch(Transformer):
    """ByteRecords → MiniBatches with decode + crop + flip + normalise done by the C++ pool in ONE pass per pixel into a
    pinned staging buffer (so the H2D copy can be asynchronous). ``mean``/``std`` are in R,G,B order, values 0..255
    unless ``scale`` says otherwise; output is NCHW float32 (or bf16 bits with ``bf16=True``), RGB when ``to_rgb``."""

    def __init__(self, width: int, height: int, batch_size: int, mean=(0, 0, 0), std=(1, 1, 1), scale: float = 1.0,
                 random_crop: bool = True, random_flip: bool = True, to_rgb: bool = False, pin: bool = False,
                 bf16: bool = False):
        self.w, self.h, self.batch_size = width, height, batch_size
        self.mean, self.std, self.scale = list(mean), list(std), scale
        self.random_crop, self.random_flip, self.to_rgb, self.pin, self.bf16 = random_crop, random_flip, to_rgb, pin, bf16

    def _emit(self, imgs, labels):
        from ..utils import native

        n = len(imgs)
        dt = torch.bfloat16 if self.bf16 else torch.float32
        out = torch.empty((n, 3, self.h, self.w), dtype=dt)
        if self.pin and torch.cuda.is_available():
            out = out.pin_memory()
        seeds = np.array([RNG.random() | (RNG.random() << 32) for _ in range(n)], dtype=np.uint64)
        if self.to_rgb:
            imgs = [np.ascontiguousarray(im[..., ::-1]) for im in imgs]
            mean, std = self.mean, self.std
        else:
            mean, std = self.mean[::-1], self.std[::-1]           # data is BGR, constants are given R,G,B
        native.batch_images(imgs, self.h, self.w, mean, std, scale=self.scale, random_crop=self.random_crop,
                            random_flip=self.random_flip, seeds=seeds, out=out, bf16=self.bf16)
        return MiniBatch(out, torch.tensor(labels, dtype=torch.float32))

    def __call__(self, it):
        imgs, labels = [], []
        for rec in it:
            if isinstance(rec, ByteRecord):
                w, h = struct.unpack_from(">ii", rec.data, 0)
                img = np.frombuffer(rec.data, dtype=np.uint8, offset=8, count=w * h * 3).reshape(h, w, 3)
                lab = rec.label
            else:                                   # LabeledBGRImage with 0..255 content
                img, lab = rec.convertToByte(255.0), rec.label()
            imgs.append(img)
            labels.append(lab)
            if len(imgs) == self.batch_size:
                yield self._emit(imgs, labels)
                imgs, labels = [], []
        if imgs:
            yield self._emit(imgs, labels)


class BGRImgToLocalSeqFile(Transformer):
    """(LabeledBGRImage, name) stream → ``base_<i>.seq`` files of ``block_size`` records; yields the file names."""

    def __init__(self, block_size: int, base_file_name: str, has_name: bool = False):
        self.block_size, self.base, self.has_name = block_size, base_file_name, has_name

    def __call__(self, it):
        from .seqfile import write_image_seq_files

        items = ((img.convertToByte(), img.label(), name) for img, name in it)
        yield from write_image_seq_files(items, self.base, self.block_size, self.has_name)


class LocalSeqFileToBytes(Transformer):
    def __call__(self, it):
        from .seqfile import read_label, read_sequence_file

        for path in it:
            for k, v in read_sequence_file(str(path)):
                yield ByteRecord(v, float(read_label(k)))


def image_folder_dataset(path: str, scale_to: int = 256, distributed: bool = False):
    from .base import DataSet

    paths = LocalImageFiles.readPaths(path)
    reader = LocalImgReader(scale_to, normalize=1.0)
    return DataSet.array(list(reader(iter(paths))), distributed)


# ------------------------------------------------------------------------------------------------ detection demo
class Visualizer:
    """Draw the detections of one class onto an image file (S/dataset/image/Utils.scala:31-100): boxes scoring above
    ``thresh`` get a red rectangle and a "<class> <score>" caption; the result is written to
    ``outPath/<class>_<image file name>``. Nothing is written when no box passes the threshold."""

    @staticmethod
    def _vis(imgPath, className, scores, bboxes, savePath, thresh=0.3):
        import cv2

        scores = np.asarray(scores, dtype=np.float32).reshape(-1)
        bboxes = np.asarray(bboxes, dtype=np.float32).reshape(-1, 4)
        img = None
        for score, (x1, y1, x2, y2) in zip(scores, bboxes):
            if score <= thresh:
                continue
            if img is None:
                img = cv2.imread(imgPath, cv2.IMREAD_COLOR)
                if img is None:
                    raise IOError(f"cannot decode image {imgPath}")
            x1, y1, x2, y2 = int(x1), int(y1), int(x2), int(y2)
            title = f"{className} {score:.3f}"
            cv2.rectangle(img, (x1, y1), (x2, y2), (0, 0, 255), 3)
            (tw, th), base = cv2.getTextSize(title, cv2.FONT_HERSHEY_SIMPLEX, 0.5, 1)
            cv2.rectangle(img, (x1, y1 - 2 - th - base), (x1 + tw, y1 - 2 + base), (127, 0, 0), -1)
            cv2.putText(img, title, (x1, y1 - 2), cv2.FONT_HERSHEY_SIMPLEX, 0.5, (255, 255, 255), 1)
        if img is not None:
            cv2.imwrite(savePath, img)
        return img is not None

    @staticmethod
    def visDetection(imagePath: str, clsname: str, scores, bboxes, thresh: float = 0.3, outPath: str = "data/demo"):
        os.makedirs(outPath, exist_ok=True)
        path = os.path.join(outPath, f"{clsname}_{os.path.basename(imagePath)}")
        return path if Visualizer._vis(imagePath, clsname, scores, bboxes, path, thresh) else None
