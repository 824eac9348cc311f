"""BoundingBox / BboxUtil / BatchSampler / OpenCVMat / detection batchers against element-wise restatements of the
reference arithmetic (S/transform/vision/image/util/{BoundingBox,BboxUtil}.scala, label/roi/BatchSampler.scala,
MTImageFeatureToBatch.scala)."""
import math
import pickle

import numpy as np
import pytest
import torch

from bigdl_b200.transform.vision import image as V
from bigdl_b200.transform.vision.bbox import BatchSampler, BboxUtil, BoundingBox, OpenCVMat
from bigdl_b200.utils.random_generator import RNG


def test_bounding_box_geometry():
    a, b = BoundingBox(0.1, 0.1, 0.5, 0.5), BoundingBox(0.3, 0.3, 0.7, 0.9)
    assert a.width() == pytest.approx(0.4) and a.area() == pytest.approx(0.16) and a.centerX() == pytest.approx(0.3)
    inter = 0.2 * 0.2
    assert a.jaccardOverlap(b) == pytest.approx(inter / (0.16 + 0.4 * 0.6 - inter))
    assert a.isOverlap(b) and not a.isOverlap(BoundingBox(0.5, 0.5, 0.6, 0.6))
    assert a.jaccardOverlap(BoundingBox(0.6, 0.6, 0.7, 0.7)) == 0.0
    assert a.meetEmitCenterConstraint(BoundingBox(0.2, 0.2, 0.4, 0.4)) and not a.meetEmitCenterConstraint(b)
    p = BoundingBox()
    assert a.projectBbox(b, p)
    assert p.toTuple() == pytest.approx((0.5, 0.5, 1.0, 1.0))
    back = BoundingBox()
    a.locateBBox(BoundingBox(0.5, 0.5, 1.0, 1.0), back)
    assert back.toTuple() == pytest.approx((0.3, 0.3, 0.5, 0.5))
    assert not a.projectBbox(BoundingBox(0.8, 0.8, 0.9, 0.9), p)
    s = BoundingBox()
    a.scaleBox(100, 200, s)
    assert s.toTuple() == pytest.approx((20, 10, 100, 50))
    c = BoundingBox(-0.2, 0.3, 1.4, 0.9)
    c.clipBox(c)
    assert c.toTuple() == pytest.approx((0.0, 0.3, 1.0, 0.9))
    assert BoundingBox(a) == a and hash(BoundingBox(a)) == hash(a) and a != b
    assert a.setLabel(3).label == 3.0 and a.setDifficult(1).difficult == 1.0 and "BoundingBox" in repr(a)


def test_bbox_transform_inv_and_clip():
    g = torch.Generator().manual_seed(1)
    boxes = torch.rand(7, 4, generator=g) * 50
    boxes[:, 2:] += boxes[:, :2] + 5
    deltas = torch.randn(7, 8, generator=g) * 0.3
    out = BboxUtil.bboxTransformInv(boxes, deltas)
    for i in range(7):
        x1, y1, x2, y2 = boxes[i].tolist()
        w, h = x2 - x1 + 1, y2 - y1 + 1
        for r in range(2):
            dx, dy, dw, dh = deltas[i, 4 * r:4 * r + 4].tolist()
            cx, cy = dx * w + x1 + w / 2, dy * h + y1 + h / 2
            pw, ph = math.exp(dw) * w / 2, math.exp(dh) * h / 2
            assert out[i, 4 * r:4 * r + 4].tolist() == pytest.approx([cx - pw, cy - ph, cx + pw, cy + ph], rel=1e-5)
    n = BboxUtil.bboxTransformInv(boxes / 100, deltas[:, :4], normalized=True)
    w = (boxes[0, 2] - boxes[0, 0]) / 100
    assert float(n[0, 2] - n[0, 0]) == pytest.approx(float(torch.exp(deltas[0, 2]) * w), rel=1e-5)

    b = torch.tensor([[-5.0, 3.0, 120.0, 40.0], [10.0, 10.0, 11.0, 30.0], [20.0, 20.0, 60.0, 21.0]])
    scores = torch.ones(3)
    kept = BboxUtil.clipBoxes(b, 50, 100, minH=4, minW=4, scores=scores)
    assert b[0].tolist() == [0.0, 3.0, 99.0, 40.0] and kept == 1 and scores.tolist() == [1.0, 0.0, 0.0]
    assert BboxUtil.clipBoxes(torch.tensor([[-0.1, 0.2, 1.3, 0.9]])).tolist() == [[0.0, pytest.approx(0.2), 1.0,
                                                                                   pytest.approx(0.9)]]


def test_ssd_decode_and_layout_helpers():
    g = torch.Generator().manual_seed(2)
    n = 6
    prior = torch.rand(n, 4, generator=g) * 0.4
    prior[:, 2:] += prior[:, :2] + 0.1
    var = torch.tensor([0.1, 0.1, 0.2, 0.2]).repeat(n, 1)
    loc = torch.randn(n, 4, generator=g)
    dec = BboxUtil.decodeBoxes(prior, var, False, loc, False)
    for i in range(n):
        x1, y1, x2, y2 = prior[i].tolist()
        pw, ph, pcx, pcy = x2 - x1, y2 - y1, (x1 + x2) / 2, (y1 + y2) / 2
        cx, cy = 0.1 * float(loc[i, 0]) * pw + pcx, 0.1 * float(loc[i, 1]) * ph + pcy
        w, h = math.exp(0.2 * float(loc[i, 2])) * pw, math.exp(0.2 * float(loc[i, 3])) * ph
        assert dec[i].tolist() == pytest.approx([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], rel=1e-5, abs=1e-6)
    enc = BboxUtil.decodeBoxes(prior, var, True, loc, True)
    assert float(enc.min()) >= 0 and float(enc.max()) <= 1
    pb, pv = BboxUtil.getPriorBboxes(torch.cat([prior.reshape(-1), var.reshape(-1)]), n)
    assert torch.equal(pb, prior) and torch.equal(pv, var)

    batch, P, C = 2, 3, 4
    locs = torch.arange(batch * P * C * 4, dtype=torch.float32).reshape(batch, -1)
    per = BboxUtil.getLocPredictions(locs, P, C, False)
    assert per[1][2][1].tolist() == locs[1].reshape(P, C, 4)[1, 2].tolist()
    shared = BboxUtil.getLocPredictions(locs[:, :P * 4], P, 1, True)
    assert shared[0][-1].tolist() == locs[0, :P * 4].reshape(P, 4).tolist()
    conf = torch.arange(batch * P * C, dtype=torch.float32).reshape(batch, -1)
    cs = BboxUtil.getConfidenceScores(conf, P, C)
    assert cs[1][3].tolist() == conf[1].reshape(P, C)[:, 3].tolist()
    allb = BboxUtil.decodeBboxesAll([[loc] * 3], prior, var, 3, 0, False, False, False)
    assert allb[0][0] is None and torch.allclose(allb[0][2], dec)


def test_decode_rois_ground_truths_areas_vote():
    rows = torch.tensor([[1, 0.9, 1, 2, 3, 4], [2, 0.8, 5, 6, 7, 8]], dtype=torch.float32)
    vec = torch.cat([torch.tensor([2.0]), rows.reshape(-1), torch.zeros(5)])
    assert torch.equal(BboxUtil.decodeRois(vec), rows) and BboxUtil.decodeRois(rows) is rows
    assert BboxUtil.decodeRois(torch.zeros(13)).numel() == 0
    gt = torch.tensor([[0, 1, 0, 1, 1, 5, 5], [0, 2, 0, 2, 2, 6, 6], [1, -1, 0, 0, 0, 0, 0], [3, 4, 0, 1, 1, 2, 2]],
                      dtype=torch.float32)
    assert BboxUtil.getGroundTruthIndices(gt) == {0: (0, 2), 1: (2, 1), 3: (3, 1)}
    gts = BboxUtil.getGroundTruths(gt)
    assert sorted(gts) == [0, 2] and gts[0].shape[0] == 2 and float(gts[2][0, 1]) == 4
    b = torch.tensor([[0.0, 0.0, 9.0, 4.0]])
    assert BboxUtil.getAreas(b).tolist() == [50.0] and BboxUtil.getAreas(b, normalized=True).tolist() == [36.0]
    assert BboxUtil.getAreas(torch.tensor([[7.0, 0.0, 0.0, 9.0, 4.0]]), startInd=2).tolist() == [50.0]
    s = torch.tensor([[1.0, 2.0, 3.0, 4.0]])
    BboxUtil.scaleBBox(s, 10, 100)
    assert s.tolist() == [[100.0, 20.0, 300.0, 40.0]]

    box_all = torch.tensor([[0.0, 0, 10, 10], [1.0, 1, 11, 11], [50.0, 50, 60, 60]])
    sc_all = torch.tensor([0.9, 0.6, 0.8])
    lab = BboxUtil.bboxVote(torch.tensor([0.9, 0.8]), box_all[[0, 2]].clone(), sc_all, box_all)
    want0 = (0.9 * box_all[0] + 0.6 * box_all[1]) / 1.5
    assert lab.bboxes[0].tolist() == pytest.approx(want0.tolist()) and lab.bboxes[1].tolist() == [50, 50, 60, 60]


def test_decode_with_weight():
    bb = torch.tensor([[10.0, 20.0, 50.0, 80.0]])
    enc = torch.tensor([[1.0, -2.0, 0.5, 1000.0]])
    out = BboxUtil.decodeWithWeight(enc, bb, (10.0, 10.0, 5.0, 5.0))
    w, h = 41.0, 61.0
    cx, cy = 10 + w / 2 + 0.1 * w, 20 + h / 2 - 0.2 * h
    pw, ph = math.exp(0.1) * w / 2, math.exp(62.5) * h / 2
    assert out[0].tolist() == pytest.approx([cx - pw, cy - ph, cx + pw - 1, cy + ph - 1], rel=1e-5)


def test_batch_sampler_and_random_sampler():
    RNG.setSeed(7)
    lab = V.RoiLabel([1.0, 2.0], [[0.1, 0.1, 0.5, 0.6], [0.55, 0.5, 0.95, 0.9]])
    s = BatchSampler(maxSample=3, maxTrials=200, minScale=0.3, minAspectRatio=0.5, maxAspectRatio=2, minOverlap=0.5)
    got: list = []
    s.sample(BoundingBox(0, 0, 1, 1), lab, got)
    assert len(got) == 3
    for b in got:
        assert 0 <= b.x1 < b.x2 <= 1 + 1e-6 and 0.5 - 1e-6 <= b.width() / b.height() <= 2 + 1e-6
        assert max(b.jaccardOverlap(BoundingBox(*g)) for g in lab.bboxes) >= 0.5
        assert s.satisfySampleConstraint(b, lab)
    whole: list = []
    BatchSampler.generateBatchSamples(lab, [BatchSampler(maxTrials=1)], whole)
    assert [b.toTuple() for b in whole] == [(0.0, 0.0, 1.0, 1.0)]
    assert BatchSampler(maxOverlap=0.0).satisfySampleConstraint(BoundingBox(0.96, 0.96, 1, 1), lab)
    with pytest.raises(ValueError):
        BatchSampler(minScale=0.9, maxScale=0.5)

    sampler = V.RandomSampler()
    assert len(sampler.batchSamplers) == 7 and isinstance(sampler, V.Crop)
    sizes = set()
    for _ in range(12):
        f = V.ImageFeature(np.zeros((60, 80, 3), np.uint8))
        f["label"] = V.RoiLabel([1.0, 2.0], [[0.1, 0.1, 0.5, 0.6], [0.55, 0.5, 0.95, 0.9]])
        f = sampler.transform(f)
        h, w = f["mat"].shape[:2]
        sizes.add((h, w))
        bb = f["label"].bboxes
        assert f["label"].classes.shape[-1] == bb.shape[0]
        assert bb.size == 0 or (bb.min() >= 0 and bb.max() <= 1)
    assert len(sizes) > 1            # some crops, not always the whole image


def test_opencv_mat(tmp_path):
    img = (np.arange(6 * 8 * 3) % 255).astype(np.uint8).reshape(6, 8, 3)
    m = OpenCVMat(img)
    assert m.shape() == (6, 8, 3) and (m.height(), m.width(), m.channels()) == (6, 8, 3)
    png = OpenCVMat.imencode(m)
    assert np.array_equal(OpenCVMat.fromImageBytes(png).mat, img)
    p = tmp_path / "x.png"
    p.write_bytes(png)
    assert np.array_equal(OpenCVMat.read(str(p)).mat, img)
    fl, h, w, c = OpenCVMat.toFloatPixels(m)
    assert (h, w, c) == (6, 8, 3) and fl.dtype == np.float32
    assert np.array_equal(OpenCVMat.fromFloats(fl, h, w, c).mat, img.astype(np.float32))
    by, h, w, c = OpenCVMat.toBytePixels(OpenCVMat(img.astype(np.float32)))
    assert np.array_equal(OpenCVMat.fromPixelsBytes(by, h, w, c).mat, img)
    chw = torch.from_numpy(img.astype(np.float32)).permute(2, 0, 1)
    assert np.array_equal(OpenCVMat.fromTensor(chw, "CHW").mat, img.astype(np.float32))
    with pytest.raises(ValueError):
        OpenCVMat.fromFloats(fl, 6, 8, 5)
    m2 = pickle.loads(pickle.dumps(m))
    assert np.array_equal(m2.mat, img) and m2.mat.dtype == np.uint8
    drawn = OpenCVMat(np.zeros((40, 40, 3), np.uint8)).drawBoundingBox(BoundingBox(5, 10, 30, 35, False), "a")
    assert drawn.mat.sum() > 0
    m.release()
    assert m.isReleased and m.mat is None


def _roi_feature(h, w, n):
    f = V.ImageFeature(np.full((h, w, 3), 7, np.uint8))
    f["label"] = V.RoiLabel(np.arange(1, n + 1, dtype=np.float32), np.tile([[1.0, 2.0, 3.0, 4.0]], (n, 1)))
    f[V.RoiLabel.ISCROWD] = np.zeros(n, np.float32)
    return f


def test_roi_batchers():
    feats = [_roi_feature(10, 12, 2), _roi_feature(10, 12, 1), _roi_feature(10, 12, 3)]
    out = list(V.RoiMTImageFeatureToBatch(12, 10, 2, toRGB=False)(iter(feats)))
    assert [b.size() for b in out] == [2, 1] and isinstance(out[0], V.RoiMiniBatch)
    assert tuple(out[0].getInput().shape) == (2, 3, 10, 12)
    tgt = out[0].getTarget()
    assert tgt[1][V.RoiLabel.CLASSES].tolist() == [1.0, 2.0] and tuple(tgt[2][V.RoiLabel.BBOXES].shape) == (1, 4)
    assert tgt[1][V.RoiLabel.ISCROWD].tolist() == [0.0, 0.0] and tuple(tgt[1][V.RoiLabel.ORIGSIZE]) == (10, 12, 3)
    assert V.RoiLabel.getBBoxes(tgt[1]) is tgt[1]["bboxes"]
    sl = out[0].slice(2, 1)
    assert sl.size() == 1 and sl.target[0].size() == 1
    with pytest.raises(NotImplementedError):
        out[0].set([])

    mixed = [_roi_feature(10, 12, 1), _roi_feature(17, 9, 2)]
    rb = next(iter(V.MTImageFeatureToBatchWithResize(8, 2, None, False)(iter(mixed))))
    assert tuple(rb.getInput().shape) == (2, 3, 24, 16)
    x = rb.getInput()
    assert float(x[0, :, :10, :12].min()) == 7 and float(x[0, :, 10:].abs().sum()) == 0 and float(x[1, :, :, 9:].abs().sum()) == 0
    assert rb.originalSizes == [(10, 12, 3), (17, 9, 3)]
    rb2 = next(iter(V.RoiImageFeatureToBatchWithResize(-1, 2)(iter(mixed))))
    assert tuple(rb2.getInput().shape) == (2, 3, 17, 12)

    bad = _roi_feature(10, 12, 2)
    bad[V.RoiLabel.ISCROWD] = np.zeros(3, np.float32)
    with pytest.raises(Exception):
        next(iter(V.RoiMTImageFeatureToBatch(12, 10, 1)(iter([bad]))))

    cls = [V.ImageFeature(np.zeros((10, 12, 3), np.uint8), [float(i + 1)]) for i in range(3)]
    cb = list(V.ClassificationMTImageFeatureToBatch(12, 10, 3)(iter(cls)))
    assert tuple(cb[0].getInput().shape) == (3, 3, 10, 12) and cb[0].getTarget().reshape(-1).tolist() == [1.0, 2.0, 3.0]


def test_roi_label_from_tensor_and_chained():
    lab = V.RoiLabel.fromTensor(torch.tensor([[3, 0, 1, 2, 3, 4], [5, 1, 5, 6, 7, 8]], dtype=torch.float32))
    assert lab.classes.tolist() == [[3.0, 5.0], [0.0, 1.0]] and lab.bboxes.shape == (2, 4)
    with pytest.raises(ValueError):
        V.RoiLabel([1.0], [[0, 0, 1, 1], [0, 0, 1, 1]])
    ch = V.ChainedFeatureTransformer(V.Resize(6, 8), V.HFlip())
    f = ch.transform(V.ImageFeature(np.arange(4 * 4 * 3, dtype=np.uint8).reshape(4, 4, 3)))
    assert f["mat"].shape[:2] == (6, 8) and ch.enableIgnoreException() is ch

    from bigdl_b200.dataset.base import Sample
    feats = []
    for i in range(3):
        g = V.ImageFeature(np.zeros((2, 2, 3), np.uint8))
        g["sample"] = Sample(torch.full((3, 2, 2), float(i)), torch.tensor([float(i + 1)]))
        feats.append(g)
    bs = list(V.ImageFeatureToMiniBatch(2)(iter(feats)))
    assert [b.size() for b in bs] == [2, 1] and float(bs[1].getInput().mean()) == 2.0


def test_untested_ops_and_transformers(tmp_path):
    from bigdl_b200.nn import layer as L
    from bigdl_b200.nn import ops as O
    from bigdl_b200.utils.table import T

    assert O.Any().forward(T(torch.tensor([[0.0, 1.0], [0.0, 0.0]]), torch.tensor([1]))).tolist() == [True, False]
    assert O.Substr().forward(T(np.array([b"hello", b"world"], dtype=object), torch.tensor(1), torch.tensor(3))).tolist() == [b"ell", b"orl"]
    assert O.Substr().forward(T("abcdef", torch.tensor(2), torch.tensor(2))) == "cd"
    op = O.ModuleToOperation(L.Linear(3, 2))
    assert tuple(op.forward(torch.rand(4, 3)).shape) == (4, 2)

    class Greater(O.Compare):
        def compareFloat(self, a, b):
            return a > b

    assert Greater().forward(T(torch.tensor([1.0, 3.0]), torch.tensor([2.0, 2.0]))).tolist() == [False, True]

    f = V.ImageFeature(np.full((4, 6, 3), 10, np.uint8))
    V.PixelNormalize(np.full((4, 6, 3), 4.0)).transform(f)
    assert f["mat"].dtype == np.float32 and float(f["mat"].mean()) == 6.0
    g = V.ImageFeature(np.zeros((20, 10, 3), np.uint8))
    g["label"] = V.RoiLabel([1.0], [[2.0, 4.0, 6.0, 8.0]])
    pipe = V.Pipeline([V.Resize(40, 30), V.RoiResize(normalized=False)])           # Resize(height, width)
    pipe.transform(g)
    assert g["mat"].shape[:2] == (40, 30) and g["label"].bboxes[0].tolist() == [6.0, 8.0, 18.0, 16.0]
    raw = V.ImageFeature()
    raw["bytes"], raw["originalSize"] = bytes(range(24)), (2, 4, 3)
    V.PixelBytesToMat().transform(raw)
    assert raw["mat"].shape == (2, 4, 3) and int(raw["mat"][1, 3, 2]) == 23
    frame = V.DistributedImageFrame([np.zeros((2, 2, 3), np.uint8)] * 3, [1.0, 2.0, 3.0])
    assert len(frame.features) == 3 and float(np.asarray(frame.features[2]["label"]).reshape(-1)[0]) == 3.0

    from bigdl_b200.nn.keras import layers as K
    from bigdl_b200.optim.optim_method import SGD, Default
    assert issubclass(K.MaxPooling1D, K.Pooling1D) and issubclass(K.MaxPooling3D, K.Pooling3D)
    assert issubclass(K.GlobalAveragePooling2D, K.GlobalPooling2D) and issubclass(K.GlobalMaxPooling3D, K.GlobalPooling3D)
    sgd = SGD(learningrate=0.1, learningrate_decay=0.5, learningrate_schedule=Default())
    x, g2 = torch.ones(2), torch.ones(2)
    sgd.update(x, g2)
    sgd.update(x, g2)
    assert sgd.getLearningRate() == pytest.approx(-0.1 / (1 + 1 * 0.5))          # negative rate, decayed by neval


def test_reference_bounding_box_spec_values():
    """Numbers asserted by T/transform/vision/image/util/BoundingBoxSpec.scala."""
    scaled = BoundingBox()
    BoundingBox(1, 4, 5, 6, False).scaleBox(1.0 / 4, 1.0 / 2, scaled)
    assert scaled == BoundingBox(0.5, 1, 2.5, 1.5)
    assert BoundingBox(0, 0, 5, 3, False).meetEmitCenterConstraint(BoundingBox(1, 0, 7, 4, False))
    assert not BoundingBox(0, 0, 5, 3, False).meetEmitCenterConstraint(BoundingBox(4, 0, 7, 4, False))
    b1 = BoundingBox(0.2, 0.3, 0.3, 0.5)
    assert abs(b1.jaccardOverlap(BoundingBox(0.1, 0.1, 0.3, 0.4)) - 1.0 / 7) < 1e-6
    assert abs(b1.jaccardOverlap(BoundingBox(0.1, 0.1, 0.4, 0.6)) - 2.0 / 15) < 1e-6
    assert abs(b1.jaccardOverlap(BoundingBox(0, 0, 0.1, 0.1))) < 1e-6
    box1, box2 = BoundingBox(0.222159, 0.427017, 0.606492, 0.679355), BoundingBox(0.418, 0.396396, 0.55, 0.666667)
    proj = BoundingBox()
    assert box1.projectBbox(box2, proj)
    assert np.allclose(proj.toTuple(), (0.509561, 0.0, 0.853014, 0.949717), atol=1e-5)
    assert box1.meetEmitCenterConstraint(box2)
    assert not BoundingBox(0.0268208, 0.388175, 0.394421, 0.916685).meetEmitCenterConstraint(box2)

    # getPriorBboxes
    n = 4
    prior = torch.zeros(2 * n * 4)
    for i in range(n):
        prior[i * 4:i * 4 + 4] = torch.tensor([i * 0.1, i * 0.1, i * 0.1 + 0.2, i * 0.1 + 0.1])
    prior[n * 4:] = 0.1
    boxes, variances = BboxUtil.getPriorBboxes(prior.reshape(1, 2, n * 4), n)
    assert tuple(boxes.shape) == (n, 4) and torch.allclose(variances, torch.full((n, 4), 0.1))
    assert torch.allclose(boxes[2], torch.tensor([0.2, 0.2, 0.4, 0.3]))

    # getGroundTruths
    inp = torch.tensor([[0.0, 1.0, 0.14285715, 0.1904762, 0.23809524, 0.2857143, 0.33333334],
                        [0.0, 1.0, 0.47619048, 0.52380955, 0.5714286, 0.61904764, 0.6666667],
                        [1.0, 3.0, 0.8095238, 0.85714287, 0.9047619, 0.95238096, 1.0]])
    gts = BboxUtil.getGroundTruths(inp)
    assert torch.equal(gts[0], inp[:2]) and torch.equal(gts[1], inp[2:])
    assert torch.equal(BboxUtil.getGroundTruths(inp[2:])[0], inp[2:])
    label = torch.tensor([[3.0, 8.0, 0.0, 0.241746, 0.322738, 0.447184, 0.478388],
                          [3.0, 8.0, 0.0, 0.318659, 0.336546, 0.661729, 0.675461],
                          [3.0, 8.0, 0.0, 0.56154, 0.300144, 0.699173, 0.708098],
                          [3.0, 8.0, 0.0, 0.220494, 0.327759, 0.327767, 0.396797],
                          [3.0, 8.0, 0.0, 0.194182, 0.317717, 0.279191, 0.389266],
                          [4.0, -1.0, -1.0, -1.0, -1.0, -1.0, -1.0],
                          [5.0, -1.0, -1.0, -1.0, -1.0, -1.0, -1.0],
                          [6.0, 10.0, 0.0, 0.67894, 0.471823, 0.929308, 0.632044],
                          [6.0, 10.0, 0.0, 0.381443, 0.572376, 0.892489, 0.691713],
                          [7.0, 9.0, 0.0, 0.0, 0.0620616, 0.667269, 1.0]])
    lg = BboxUtil.getGroundTruths(label)
    assert len(lg) == 3 and lg[0].shape[0] == 5 and float(lg[0][0, 0]) == 3
    assert lg[3].shape[0] == 2 and float(lg[3][0, 0]) == 6 and lg[4].shape[0] == 1 and float(lg[4][0, 0]) == 7


def test_reference_loc_predictions_and_decode_values():
    """"getLocPredictions shared" and "decodeBoxes" of BoundingBoxSpec.scala."""
    num, per = 2, 2
    loc = torch.zeros(num, per * 4, 1, 1)
    for i in range(num):
        for j in range(per):
            v = i * per * 0.1 + j * 0.1
            loc[i, j * 4:j * 4 + 4, 0, 0] = torch.tensor([v, v, v + 0.2, v + 0.2])
    out = BboxUtil.getLocPredictions(loc, per, 1, True)
    assert len(out) == num and all(len(o) == 1 for o in out)
    for i in range(num):
        for j in range(per):
            v = i * per * 0.1 + j * 0.1
            assert torch.allclose(out[i][0][j], torch.tensor([v, v, v + 0.2, v + 0.2]), atol=1e-6)

    prior = torch.tensor([[0.1 * i, 0.1 * i, 0.1 * i + 0.2, 0.1 * i + 0.2] for i in range(1, 5)])
    var = torch.tensor([[0.1, 0.1, 0.2, 0.2]] * 4)
    boxes = torch.tensor([[0.0, 0.75, math.log(2), math.log(1.5)]] * 4)
    dec = BboxUtil.decodeBoxes(prior, var, False, boxes, True)
    for i in range(4):
        assert torch.allclose(dec[i], torch.tensor([0.0, 0.2, 0.4, 0.5]) + i * 0.1, atol=1e-5)
