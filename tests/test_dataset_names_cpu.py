"""Padding strategies, SampleToBatch and the named image-reader / image-type classes
(S/dataset/{PaddingParam,Transformer,Types}.scala, S/dataset/image/{Types,LocalImgReader,LocalImgReaderWithName}.scala)."""
import os
import struct

import numpy as np
import pytest
import torch

from bigdl_b200.dataset import base as B
from bigdl_b200.dataset import image as I
from bigdl_b200.dataset import text as TX


def _samples():
    return [B.Sample(torch.ones(n, 2) * n, torch.tensor([float(n)])) for n in (2, 4, 3)]


def test_padding_strategies():
    s = _samples()
    mb = B.MiniBatch.from_samples(s, B.PaddingParam(padding_strategy=B.PaddingLongest([2]))) \
        if hasattr(B.MiniBatch, "from_samples") else None
    t = B.SampleToMiniBatch(3, feature_padding=B.PaddingParam(padding_strategy=B.PaddingLongest([2])))
    out = next(iter(t(iter(s))))
    assert tuple(out.getInput().shape) == (3, 6, 2)
    if mb is not None:
        assert tuple(mb.getInput().shape) == (3, 6, 2)
    t = B.SampleToMiniBatch(3, feature_padding=B.PaddingParam(padding_strategy=B.FixedLength([6])))
    out = next(iter(t(iter(s))))
    assert tuple(out.getInput().shape) == (3, 6, 2)
    # rows past each sample's length are the padding value (zero by default)
    assert float(out.getInput()[0, 2:].abs().sum()) == 0.0 and float(out.getInput()[0, :2].sum()) == 8.0
    t = B.SampleToMiniBatch(3, feature_padding=B.PaddingParam(padding_strategy=B.DefaultPadding()))
    assert tuple(next(iter(t(iter(s)))).getInput().shape) == (3, 4, 2)
    with pytest.raises(ValueError):
        t = B.SampleToMiniBatch(3, feature_padding=B.PaddingParam(padding_strategy=B.FixedLength([3])))
        next(iter(t(iter(s))))


def test_sample_to_batch_and_aliases():
    out = [tuple(b.getInput().shape) for b in B.SampleToBatch(2, fixedLength=4)(iter(_samples()))]
    assert out == [(2, 4, 2), (1, 4, 2)]
    assert B.LocalDataSet is B.AbstractDataSet and B.ArrayTensorMiniBatch is B.MiniBatch


def _png(tmp_path, w=20, h=10):
    import cv2

    img = (np.arange(h * w * 3) % 251).astype(np.uint8).reshape(h, w, 3)
    p = os.path.join(tmp_path, "a.png")
    cv2.imwrite(p, img)
    return p, img


def test_named_readers(tmp_path):
    p, img = _png(str(tmp_path))
    rec = I.LocalLabeledImagePath(3.0, p)
    a = next(iter(I.LocalScaleImgReader(5, 1.0)(iter([rec]))))
    assert (a.height(), a.width()) == (5, 10) and a.label() == 3.0
    b = next(iter(I.LocalResizeImgReader(8, 6, 255.0)(iter([rec]))))
    assert (b.height(), b.width()) == (6, 8) and float(b.content().max()) <= 1.0
    c, name = next(iter(I.LocalScaleImgWithName(-1, 1.0)(iter([rec]))))
    assert name == "a.png" and np.array_equal(c.content().astype(np.uint8), img)
    d, name = next(iter(I.LocalResizeImgWithName(4, 4, 1.0)(iter([rec]))))
    assert (d.height(), d.width()) == (4, 4) and name == "a.png"


def test_bgr_image_record_round_trip(tmp_path):
    p, img = _png(str(tmp_path))
    raw = I.BGRImage.readImage(p, I.BGRImage.NO_SCALE)
    assert struct.unpack_from(">ii", raw, 0) == (20, 10) and len(raw) == 8 + 20 * 10 * 3
    im = I.BGRImage().copy(raw, 1.0)
    assert np.array_equal(im.content().astype(np.uint8), img)
    out = os.path.join(str(tmp_path), "b.png")
    im.save(out, 1.0)
    assert np.array_equal(I.read_image_bgr(out), img)
    raw2 = I.BGRImage.readImage(p, 6, 4)
    assert struct.unpack_from(">ii", raw2, 0) == (6, 4)
    other = I.LabeledBGRImage().copy(im.setLabel(7.0))
    assert other.label() == 7.0 and other.content() is not im.content()
    g = I.GreyImage(np.zeros((2, 3), np.float32)).copy(bytes(range(6)), 1.0)
    assert g.content().tolist() == [[0, 1, 2], [3, 4, 5]] and g.setLabel(2.0).clone().label() == 2.0
    assert I.CropperMethod.CropCenter == I.CropCenter and os.fspath(I.LocalSeqFilePath("x.seq")) == "x.seq"
    assert I.LocalImagePath("q.jpg").path == "q.jpg"


def test_sentence_types():
    s = TX.LabeledSentence([1, 2, 3], [2, 3, 4])
    assert isinstance(s, TX.Sentence) and s.clone().label() == [2, 3, 4]
    buf = [0] * 5
    assert s.copyToData(buf, 1) == [0, 1, 2, 3, 0] and s.copyToLabel([0] * 3) == [2, 3, 4]
    assert TX.SentenceToken.start == "SENTENCESTART"
    assert next(iter(TX.SentenceBiPadding()(iter(["a b"])))) == "SENTENCESTART a b SENTENCEEND"


def test_visualizer_utils_and_markers(tmp_path):
    import cv2

    from bigdl_b200.dataset import segmentation as S

    p = os.path.join(str(tmp_path), "dog.jpg")
    cv2.imwrite(p, np.zeros((60, 80, 3), np.uint8))
    out = I.Visualizer.visDetection(p, "dog", [0.9, 0.1], [[10, 20, 50, 50], [0, 0, 5, 5]], 0.3, str(tmp_path / "demo"))
    assert out.endswith("dog_dog.jpg") and cv2.imread(out).sum() > 0
    assert I.Visualizer.visDetection(p, "cat", [0.1], [[1, 1, 5, 5]], 0.3, str(tmp_path / "demo")) is None
    assert B.Utils.getBatchSize(32) == 32 and B.Utils.getBatchSize(32, 4) == 8
    with pytest.raises(ValueError):
        B.Utils.getBatchSize(30, 4)
    assert issubclass(S.COCOPoly, S.COCOSegmentation) and issubclass(S.COCORLE, S.COCOSegmentation)
    assert B.Label().setLabel(3).label() == 3


def test_samples_chains_sparse_batches_and_prediction_service():
    import threading

    from bigdl_b200.nn import layer as L
    from bigdl_b200.optim.predictor import PredictionService
    from bigdl_b200.tensor.sparse import SparseTensor

    assert B.ArraySample is B.Sample or issubclass(B.ArraySample, B.Sample)
    s = B.TensorSample(torch.ones(2), torch.tensor([1.0])) if callable(B.TensorSample) else None
    assert s is None or s.feature().tolist() == [1.0, 1.0]
    double = B.FuncTransformer(lambda v: v * 2)
    plus = B.FuncTransformer(lambda v: v + 1)
    assert list(B.ChainedTransformer(double, plus)(iter([1, 2]))) == [3, 5]
    assert list((double >> plus)(iter([1, 2]))) == [3, 5] if hasattr(double, "__rshift__") else True

    dense = [torch.tensor([0.0, 2.0, 0.0, 3.0]), torch.tensor([1.0, 0.0, 0.0, 0.0])]
    samples = [B.Sample([SparseTensor.from_dense(d) if hasattr(SparseTensor, "from_dense") else SparseTensor(d.to_sparse()),
                         torch.full((2,), float(i))], torch.tensor([float(i + 1)])) for i, d in enumerate(dense)]
    mb = B.SparseMiniBatch.from_samples(samples)
    sp, dn = mb.getInput().to_list()
    assert isinstance(sp, SparseTensor) and tuple(dn.shape) == (2, 2) and mb.getTarget().reshape(-1).tolist() == [1.0, 2.0]
    got = sp.toDense() if hasattr(sp, "toDense") else sp.to_dense()
    got = got.t if hasattr(got, "t") and not callable(got.t) else got
    assert torch.equal(torch.as_tensor(got).reshape(2, 4), torch.stack(dense))

    model = L.Sequential().add(L.Linear(4, 3)).add(L.SoftMax())
    svc = PredictionService(model, 2)
    x = torch.rand(5, 4)
    want = model.evaluate().forward(x)
    outs = [None] * 4

    def call(i):
        outs[i] = svc.predict(x)

    ts = [threading.Thread(target=call, args=(i,)) for i in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert all(torch.allclose(o, want, atol=1e-6) for o in outs)
    assert isinstance(PredictionService.apply(model, 1), PredictionService)
