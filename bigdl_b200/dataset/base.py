"""Sample / MiniBatch / DataSet / Transformer (reference: S/dataset/Sample.scala, MiniBatch.scala, DataSet.scala,
Transformer.scala).

* ``Sample(features, labels)`` — one record; features/labels are tensors (or lists of tensors).
* ``MiniBatch`` — batched features/labels with 1-based ``slice(offset, length)``.
* ``DataSet.array(...)`` → ``LocalArrayDataSet``; ``DataSet.rdd(...)`` → ``DistributedDataSet`` which shards the
  records over ranks (one process per GPU replaces one Spark partition per node) and, like
  ``CachedDistriDataSet`` (DataSet.scala:247-322), serves an *infinite* shuffled training iterator that starts at a
  random offset.
* ``Transformer`` objects chain with ``->`` in Scala; here with ``>>`` (or ``.chain``).
"""
from __future__ import annotations

import math
from typing import Callable, Iterable, Iterator, Sequence

import numpy as np
import torch

from ..utils.random_generator import RNG
from ..utils.table import Table


def _t(x) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x
    if type(x).__name__ == "SparseTensor":          # a sparse feature travels as is (SparseMiniBatch concatenates them)
        return x
    a = np.asarray(x)
    if a.dtype == np.float64:
        a = a.astype(np.float32)
    return torch.from_numpy(np.ascontiguousarray(a))


class Sample:
    """One training record (S/dataset/Sample.scala ArraySample/TensorSample; Python P/util/common.py:294)."""

    def __init__(self, features, labels=None):
        self.features = [_t(f) for f in features] if isinstance(features, (list, tuple)) else [_t(features)]
        if labels is None:
            self.labels = []
        else:
            self.labels = [_t(l) for l in labels] if isinstance(labels, (list, tuple)) else [_t(labels)]

    @classmethod
    def from_ndarray(cls, features, labels=None):
        return cls(features, labels)

    def feature(self, index: int = 0) -> torch.Tensor:
        return self.features[index]

    def label(self, index: int = 0) -> torch.Tensor:
        return self.labels[index]

    def numFeature(self) -> int:
        return len(self.features)

    def numLabel(self) -> int:
        return len(self.labels)

    def __repr__(self):
        shp = lambda t: tuple(t.shape) if hasattr(t, "shape") else tuple(t.size())  # noqa: E731
        return f"Sample: features: {[shp(f) for f in self.features]}, labels: {[shp(l) for l in self.labels]}"


class PaddingStrategy:
    """How long the padded first dimension of feature / label ``i`` becomes (S/dataset/MiniBatch.scala:527-575).
    ``paddingSize(sizes)`` maps the per-input sizes ``[batch, longest, ...]`` to the padded sizes."""

    def paddingSize(self, sizes):
        raise NotImplementedError

    def length_for(self, idx: int, longest: int) -> int:
        """Padded length of input ``idx`` whose longest sample has ``longest`` steps."""
        out = self.paddingSize([[0, longest] if i == idx else [0, 0] for i in range(idx + 1)])
        return int(out[idx][1])


class DefaultPadding(PaddingStrategy):
    """Pad to the longest sample of the batch."""

    def paddingSize(self, sizes):
        return sizes


class PaddingLongest(PaddingStrategy):
    """Longest sample of the batch plus a constant per input."""

    def __init__(self, paddingLength):
        self.paddingLength = list(paddingLength)

    def paddingSize(self, sizes):
        for i, s in enumerate(sizes):
            if i < len(self.paddingLength):
                s[1] += self.paddingLength[i]
        return sizes


class FixedLength(PaddingStrategy):
    """A fixed length per input (negative = keep the longest); shorter than the data is an error."""

    def __init__(self, fixedLength):
        self.fixedLength = list(fixedLength) if isinstance(fixedLength, (list, tuple)) else [fixedLength]

    def paddingSize(self, sizes):
        for i, s in enumerate(sizes):
            if i >= len(self.fixedLength):
                continue
            fixed = self.fixedLength[i]
            if not (fixed >= s[1] or fixed < 0):
                raise ValueError(f"{i}th FixedLength={fixed} is smaller than its FeatureLength={s[1]}")
            if fixed >= s[1]:
                s[1] = fixed
        return sizes


class PaddingParam:
    """Variable-length padding policy (S/dataset/MiniBatch.scala:523): pad each feature to the longest in the batch
    (a ``PaddingStrategy`` — ``FixedLength``, ``PaddingLongest`` — changes the target length) with ``padding_tensor``
    values (default 0). ``fixed_length`` (an int or one per input) is the shorthand for ``FixedLength``."""

    def __init__(self, padding_tensor=None, padding_strategy=None, fixed_length=None):
        self.padding_tensor = padding_tensor
        self.strategy = padding_strategy if isinstance(padding_strategy, PaddingStrategy) else None
        self.fixed_length = fixed_length if fixed_length is not None else (
            None if self.strategy is not None else padding_strategy)


def _stack(tensors: Sequence[torch.Tensor], pad: PaddingParam | None, idx: int) -> torch.Tensor:
    shapes = {tuple(t.shape) for t in tensors}
    if len(shapes) == 1 and (pad is None or (pad.fixed_length is None and getattr(pad, "strategy", None) is None)):
        return torch.stack(list(tensors))
    if pad is None:
        raise ValueError("samples have different sizes; pass a PaddingParam to pad them")
    nd = tensors[0].dim()
    target = [max(t.shape[d] for t in tensors) for d in range(nd)]
    if pad.fixed_length is not None:
        fl = pad.fixed_length[idx] if isinstance(pad.fixed_length, (list, tuple)) else pad.fixed_length
        if fl is not None and fl > 0:
            target[0] = max(target[0], fl)
    if getattr(pad, "strategy", None) is not None and nd > 0:
        target[0] = max(target[0], pad.strategy.length_for(idx, target[0]))
    value, row = 0.0, None
    if pad.padding_tensor is not None:
        pt = pad.padding_tensor[idx] if isinstance(pad.padding_tensor, (list, tuple)) else pad.padding_tensor
        pt = _t(pt)
        if pt.numel() > 1 and nd > 1 and list(pt.shape) == target[1:]:
            row = pt          # one padding ROW per time step (e.g. the one-hot END token of the RNN example)
        else:
            value = float(pt.reshape(-1)[0])
    out = torch.full([len(tensors)] + target, value, dtype=tensors[0].dtype)
    if row is not None:
        out[:] = row.to(out.dtype)
    for i, t in enumerate(tensors):
        out[(i,) + tuple(slice(0, s) for s in t.shape)] = t
    return out


class MiniBatch:
    """Batched activity pair (S/dataset/MiniBatch.scala:34). ``getInput``/``getTarget`` return a Tensor, or a Table
    when there are several feature/label tensors."""

    def __init__(self, inputs, targets=None):
        self.inputs = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
        self.targets = [] if targets is None else (list(targets) if isinstance(targets, (list, tuple)) else [targets])

    def size(self) -> int:
        return self.inputs[0].shape[0]

    def getInput(self):
        return self.inputs[0] if len(self.inputs) == 1 else Table(*self.inputs)

    def getTarget(self):
        if not self.targets:
            return None
        return self.targets[0] if len(self.targets) == 1 else Table(*self.targets)

    get_input, get_target = getInput, getTarget

    def slice(self, offset: int, length: int) -> "MiniBatch":
        """1-based offset, like the reference."""
        s = slice(offset - 1, offset - 1 + length)
        return MiniBatch([t[s] for t in self.inputs], [t[s] for t in self.targets])

    def to(self, device, non_blocking: bool = True) -> "MiniBatch":
        return MiniBatch([t.to(device, non_blocking=non_blocking) for t in self.inputs],
                         [t.to(device, non_blocking=non_blocking) for t in self.targets])

    def pin(self) -> "MiniBatch":
        if torch.cuda.is_available():
            return MiniBatch([t.pin_memory() for t in self.inputs], [t.pin_memory() for t in self.targets])
        return self

    @staticmethod
    def from_samples(samples: Sequence[Sample], feature_padding: PaddingParam | None = None,
                     label_padding: PaddingParam | None = None) -> "MiniBatch":
        nf, nl = samples[0].numFeature(), samples[0].numLabel()
        ins = [_stack([s.features[i] for s in samples], feature_padding, i) for i in range(nf)]
        tgs = [_stack([s.labels[i] for s in samples], label_padding, i) for i in range(nl)]
        return MiniBatch(ins, tgs)


# ====================================================================== transformers
class Transformer:
    """Iterator → iterator stage (S/dataset/Transformer.scala). ``a >> b`` chains (Scala ``a -> b``)."""

    def __call__(self, prev: Iterator) -> Iterator:  # pragma: no cover
        raise NotImplementedError

    def chain(self, other: "Transformer") -> "Transformer":
        return ChainedTransformer(self, other)

    __rshift__ = chain

    def cloneTransformer(self):
        import copy

        return copy.deepcopy(self)


class ChainedTransformer(Transformer):
    def __init__(self, first: Transformer, last: Transformer):
        self.first, self.last = first, last

    def __call__(self, prev):
        return self.last(self.first(prev))


class Identity(Transformer):
    def __call__(self, prev):
        return prev


class FuncTransformer(Transformer):
    def __init__(self, fn: Callable):
        self.fn = fn

    def __call__(self, prev):
        return (self.fn(x) for x in prev)


class SampleToMiniBatch(Transformer):
    """Group ``batch_size`` samples into a MiniBatch (S/dataset/Transformer.scala:309). ``batch_size`` is the batch of
    THIS process; ``partition_num`` divides a global batch like the reference's totalBatch/partitionNum."""

    def __init__(self, batch_size: int, feature_padding=None, label_padding=None, partition_num: int | None = None,
                 drop_remainder: bool = False):
        self.batch_size = batch_size // partition_num if partition_num else batch_size
        self.feature_padding, self.label_padding, self.drop_remainder = feature_padding, label_padding, drop_remainder

    def __call__(self, prev):
        buf = []
        for s in prev:
            buf.append(s)
            if len(buf) == self.batch_size:
                yield MiniBatch.from_samples(buf, self.feature_padding, self.label_padding)
                buf = []
        if buf and not self.drop_remainder:
            yield MiniBatch.from_samples(buf, self.feature_padding, self.label_padding)


class SampleToBatch(SampleToMiniBatch):
    """The pre-0.2 spelling (S/dataset/Transformer.scala:136, deprecated there in favour of SampleToMiniBatch):
    ``SampleToBatch(totalBatch, featurePadding, labelPadding, fixedLength, partitionNum)``."""

    def __init__(self, totalBatch: int, featurePadding=None, labelPadding=None, fixedLength=None, partitionNum=None):
        fp = PaddingParam(featurePadding, fixed_length=fixedLength) if (featurePadding is not None or fixedLength) else None
        lp = PaddingParam(labelPadding, fixed_length=fixedLength) if (labelPadding is not None or fixedLength) else None
        super().__init__(totalBatch, fp, lp, partitionNum)


# ====================================================================== datasets
class AbstractDataSet:
    def data(self, train: bool) -> Iterator:  # pragma: no cover
        raise NotImplementedError

    def size(self) -> int:  # pragma: no cover
        raise NotImplementedError

    def shuffle(self):
        return self

    def transform(self, transformer: Transformer) -> "AbstractDataSet":
        """Chain a dataset ``Transformer`` (iterator -> iterator) or a vision ``FeatureTransformer`` (applied to every
        ImageFeature; P/dataset/dataset.py ``DataSet.transform``)."""
        if not isinstance(transformer, Transformer) and hasattr(transformer, "transform") and callable(transformer.transform):
            ft = transformer
            # every pass starts from the stored feature: transformers overwrite ``mat`` in place, so without the copy the
            # second epoch would transform the first epoch's output
            transformer = _FnTransformer(lambda it: (ft.transform(f.clone() if hasattr(f, "clone") else f) for f in it))
        return TransformedDataSet(self, transformer)

    def get_image_frame(self):
        """The ImageFrame this dataset was made from (``DataSet.image_frame``), or None."""
        d = self
        while d is not None:
            fr = d.__dict__.get("_image_frame")
            if fr is not None:
                return fr
            d = d.__dict__.get("base")
        return None

    def __rshift__(self, transformer: Transformer):
        return self.transform(transformer)

    def toLocal(self):
        return self

    def toDistributed(self):
        return self

    def is_distributed(self) -> bool:
        return False


class _FnTransformer(Transformer):
    """A Transformer from a plain iterator -> iterator function."""

    def __init__(self, fn):
        self.fn = fn

    def __call__(self, prev):
        return self.fn(prev)


class TransformedDataSet(AbstractDataSet):
    def __init__(self, base: AbstractDataSet, transformer: Transformer):
        self.base, self.transformer = base, transformer

    def data(self, train: bool):
        return self.transformer(self.base.data(train))

    def size(self):
        return self.base.size()

    def shuffle(self):
        self.base.shuffle()
        return self

    def is_distributed(self):
        return self.base.is_distributed()

    def originRDD(self):
        return self.base


LocalDataSet = AbstractDataSet          # the reference's trait names (DataSet.scala:110-165); one class serves both here


class LocalArrayDataSet(AbstractDataSet):
    """In-memory records; the training iterator is infinite and walks a shuffled index (DataSet.scala:135-169)."""

    def __init__(self, buffer: Sequence):
        self.buffer = list(buffer)
        self.index = list(range(len(self.buffer)))

    def shuffle(self):
        RNG.shuffle(self.index)
        return self

    def size(self):
        return len(self.buffer)

    def data(self, train: bool):
        n = len(self.buffer)
        if not train:
            return iter(self.buffer)

        def gen():
            i = 0
            while True:
                yield self.buffer[self.index[i % n]]
                i += 1

        return gen()


class DistributedDataSet(LocalArrayDataSet):
    """Records sharded over ranks: rank r keeps records r, r+W, r+2W, … (the analogue of ``coalesce(nodeNumber)``);
    ``size()`` is the GLOBAL record count; the training iterator is infinite with a random start offset
    (CachedDistriDataSet, DataSet.scala:268-283)."""

    def __init__(self, buffer: Sequence, rank: int | None = None, world_size: int | None = None):
        from ..utils.engine import Engine

        self.rank = Engine.ensure().rank if rank is None else rank
        self.world_size = Engine.world_size if world_size is None else world_size
        self.global_size = len(buffer)
        super().__init__([b for i, b in enumerate(buffer) if i % self.world_size == self.rank])

    def size(self):
        return self.global_size

    def local_size(self):
        return len(self.buffer)

    def is_distributed(self):
        return True

    def data(self, train: bool):
        n = len(self.buffer)
        if not train:
            return iter(self.buffer)
        start = int(RNG.uniform(0, n)) if n > 0 else 0

        def gen():
            i = start
            while True:
                yield self.buffer[self.index[i % n]]
                i += 1

        return gen()


class DataSet:
    """Factory namespace (S/dataset/DataSet.scala:324-669). The instance methods of the Python API's ``DataSet``
    (P/dataset/dataset.py) live on the datasets the factories return; they are reachable here as
    ``DataSet.transform(ds, t)`` / ``DataSet.get_image_frame(ds)`` too."""

    transform = AbstractDataSet.transform
    get_image_frame = AbstractDataSet.get_image_frame

    @staticmethod
    def array(data: Sequence, distributed: bool = False) -> AbstractDataSet:
        return DistributedDataSet(data) if distributed else LocalArrayDataSet(data)

    @staticmethod
    def rdd(data: Iterable, partition_num: int | None = None) -> AbstractDataSet:
        """A Python iterable stands in for an RDD; it is materialised and sharded across ranks."""
        return DistributedDataSet(list(data))

    @staticmethod
    def from_ndarrays(x: np.ndarray, y: np.ndarray | None = None, distributed: bool = False) -> AbstractDataSet:
        samples = [Sample(x[i], None if y is None else y[i]) for i in range(len(x))]
        return DataSet.array(samples, distributed)

    @staticmethod
    def imageFrame(frame):
        """DataSet of the frame's ImageFeatures (DataSet.scala ``imageFrame``: a DataSet[ImageFeature], to be piped
        through image transformers); use ``frame.toDataSet()`` for the Samples written by ImageFrameToSample."""
        feats = getattr(frame, "features", None)
        feats = list(feats() if callable(feats) else feats) if feats is not None else list(frame)
        return DataSet.array(feats, bool(getattr(frame, "is_distributed", lambda: False)()))

    @staticmethod
    def image_frame(frame):
        """Python API spelling (P/dataset/dataset.py ``DataSet.image_frame``); the result remembers its frame."""
        ds = DataSet.imageFrame(frame)
        ds.__dict__["_image_frame"] = frame
        return ds

    @staticmethod
    def ImageFolder(path: str, scale_to: int = 256, distributed: bool = False):
        from .image import image_folder_dataset

        return image_folder_dataset(path, scale_to, distributed)

    @staticmethod
    def SeqFileFolder(path: str, class_num: int = 1000, distributed: bool = True):
        from .seqfile import seq_file_dataset

        return seq_file_dataset(path, class_num, distributed)


class FixedMiniBatchDataSet(AbstractDataSet):
    """One fixed MiniBatch served forever — the synthetic-data mode of the perf harnesses
    (S/models/utils/DistriOptimizerPerf.scala:100-120)."""

    def __init__(self, batch: MiniBatch, epoch_size: int | None = None, distributed: bool = True):
        self.batch, self._size, self._dist = batch, epoch_size or batch.size(), distributed

    def size(self):
        return self._size

    def is_distributed(self):
        return self._dist

    def data(self, train: bool):
        if train:
            def gen():
                while True:
                    yield self.batch
            return gen()
        n = max(1, math.ceil(self._size / max(self.batch.size(), 1)))
        return iter([self.batch] * n)


# ---- reference class names kept as aliases (S/dataset/Sample.scala ArraySample:138 / TensorSample:446)
ArraySample = Sample
TensorSample = Sample


class SparseMiniBatch(MiniBatch):
    """MiniBatch whose features may be ``bigdl_b200.tensor.SparseTensor``s (S/dataset/MiniBatch.scala SparseMiniBatch):
    sparse features of the samples are concatenated along the batch dimension, dense ones stacked as usual."""

    @staticmethod
    def from_samples(samples, feature_padding=None, label_padding=None):
        from ..tensor.sparse import SparseTensor

        first = samples[0]
        feats = first.features if hasattr(first, "features") else [first.feature()]
        out = []
        for i in range(len(feats)):
            col = [(s.features[i] if hasattr(s, "features") else s.feature()) for s in samples]
            if isinstance(col[0], SparseTensor):
                rows = [c if c.nDimension() == 2 else SparseTensor(torch.cat([torch.zeros(1, c.sp._nnz(), dtype=torch.long),
                                                                                 c.sp.indices()]), c.sp.values(),
                                                                      (1,) + tuple(c.sp.shape)) for c in col]
                out.append(SparseTensor.concat(1, rows))
            else:
                out.append(torch.stack([torch.as_tensor(c) for c in col]))
        labels = [s.label() for s in samples]
        tgt = torch.stack([torch.as_tensor(l) for l in labels]) if labels and labels[0] is not None else None
        return SparseMiniBatch(out if len(out) > 1 else out[0], tgt)


def _sort_by_length(samples):
    return sorted(samples, key=lambda s: int(torch.as_tensor(s.feature()).shape[0]))


DataSet.sortRDD = staticmethod(lambda data, is_in_order=True, group_size=1: DistributedDataSet(_sort_by_length(list(data))))
DataSet.sort_rdd = DataSet.sortRDD


# The reference's concrete distributed data set (one cached array per partition + a shuffled index, DataSet.scala:247-322)
CachedDistriDataSet = DistributedDataSet


# the reference's default MiniBatch implementation class (S/dataset/MiniBatch.scala:112 ArrayTensorMiniBatch)
ArrayTensorMiniBatch = MiniBatch


class Utils:
    """S/dataset/Utils.scala:22-45."""

    @staticmethod
    def getBatchSize(batchSize: int, totalPartition: int | None = None) -> int:
        """The share of a global batch one partition (= one rank / GPU here) processes; the batch must divide
        evenly."""
        import torch.distributed as dist

        nodes = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
        parts = totalPartition if totalPartition is not None else nodes
        if parts <= 0:
            raise ValueError(f"Utils.getBatchSize: partitionNum should be larger than 0, but get {parts}")
        if batchSize % parts:
            raise ValueError(f"Utils.getBatchSize: total batch size {batchSize} should be divided by partitionNum {parts}")
        return batchSize // parts


class Label:
    """Something that carries a label (S/dataset/Types.scala:60-66)."""

    def setLabel(self, label):
        self._label = label
        return self

    def label(self):
        return self._label


class Image:
    """Marker base of the in-memory image types (S/dataset/Types.scala:27-31)."""
