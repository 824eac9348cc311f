"""Data pipeline: SequenceFile container, legacy BGR image transformers, C++ batcher, text, COCO masks, RowTransformer,
vision-2.0 ImageFrame transformers (mirrors T/dataset/*Spec.scala and T/transform/vision/image/*Spec.scala)."""
import json
import os

import numpy as np
import pytest
import torch

from bigdl_b200.dataset import image as I
from bigdl_b200.dataset import seqfile as SF
from bigdl_b200.dataset import text as TX
from bigdl_b200.dataset.datamining import RowTransformer
from bigdl_b200.dataset.segmentation import COCODataset, MaskUtils, PolyMasks, RLEMasks
from bigdl_b200.transform.vision import image as V
from bigdl_b200.utils.random_generator import RNG


def _imgs(n=5, h=40, w=50):
    rs = np.random.RandomState(0)
    return [rs.randint(0, 256, size=(h + i, w + 2 * i, 3), dtype=np.uint8) for i in range(n)]


def test_vint_and_sequence_file_roundtrip(tmp_path):
    for v in (0, 1, 127, 128, 255, 256, 70000, -1, -112, -113, -70000, 2 ** 31 - 1):
        b = SF._write_vint(v)
        assert SF._read_vint(b, 0) == (v, len(b))
    imgs = _imgs(7)
    files = SF.write_image_seq_files(((im, i + 1, f"img{i}.jpg") for i, im in enumerate(imgs)), str(tmp_path / "part"),
                                     block_size=3, has_name=True)
    assert len(files) == 3
    recs = [r for f in files for r in SF.read_sequence_file(f)]
    assert len(recs) == 7
    for i, (k, v) in enumerate(recs):
        assert SF.read_label(k) == str(i + 1) and SF.read_name(k) == f"img{i}.jpg"
        assert np.array_equal(SF.decode_image_value(v), imgs[i])
    assert open(files[0], "rb").read(4) == b"SEQ\x06"
    ds = SF.seq_file_dataset(str(tmp_path), class_num=5, distributed=False)
    assert ds.size() == 5                             # labels 6,7 filtered


def test_legacy_bgr_pipeline(tmp_path):
    import cv2

    for cls in ("cat", "dog"):
        os.makedirs(tmp_path / cls)
        for i in range(3):
            cv2.imwrite(str(tmp_path / cls / f"{i}.png"), _imgs(1, 60, 80)[0])
    paths = I.LocalImageFiles.readPaths(str(tmp_path))
    assert [p.label for p in paths] == [1.0] * 3 + [2.0] * 3
    RNG.setSeed(1)
    pipe = I.LocalImgReader(48, normalize=1.0) >> I.BGRImgCropper(32, 32, I.CropCenter) >> I.HFlip(0.5) >> \
        I.BGRImgNormalizer(100, 110, 120, 50, 60, 70) >> I.BGRImgToBatch(4)
    batches = list(pipe(iter(paths)))
    assert [b.size() for b in batches] == [4, 2]
    assert tuple(batches[0].getInput().shape) == (4, 3, 32, 32)
    img = next(I.LocalImgReader(48, normalize=1.0)(iter(paths[:1])))
    assert min(img.height(), img.width()) == 48
    out = next(I.ColorJitter()(iter([img.clone()])))
    assert out.data.shape == img.data.shape and np.isfinite(out.data).all()
    lit = next(I.Lighting()(iter([img.clone()])))
    assert not np.array_equal(lit.data, img.data)
    norm = I.BGRImgNormalizer.from_dataset([img])
    z = next(norm(iter([img.clone()])))
    assert abs(z.data.mean()) < 1e-3


def test_mt_batcher_matches_python_path():
    imgs = _imgs(6, 36, 36)
    recs = [I.ByteRecord(SF.encode_image_value(im), float(i + 1)) for i, im in enumerate(imgs)]
    mean, std = (120.0, 110.0, 100.0), (60.0, 55.0, 50.0)          # R,G,B
    mt = I.MTLabeledBGRImgToBatch(32, 32, 4, mean, std, random_crop=False, random_flip=False, to_rgb=True)
    got = list(mt(iter(recs)))
    ref_pipe = I.BytesToBGRImg(normalize=1.0) >> I.BGRImgCropper(32, 32, I.CropCenter) >> \
        I.BGRImgNormalizer(*mean, *std) >> I.BGRImgToBatch(4, to_rgb=True)
    ref = list(ref_pipe(iter(recs)))
    assert len(got) == len(ref) == 2
    for g, r in zip(got, ref):
        assert torch.allclose(g.getInput(), r.getInput(), atol=1e-4)
        assert torch.equal(g.getTarget(), r.getTarget())


def test_text_pipeline(tmp_path):
    sents = ["the cat sat on the mat .", "the dog sat ."]
    padded = list(TX.SentenceBiPadding()(iter(sents)))
    assert padded[0].startswith("SENTENCESTART") and padded[0].endswith("SENTENCEEND")
    toks = list(TX.SentenceTokenizer()(iter(padded)))
    d = TX.Dictionary(toks, vocab_size=6)
    assert d.getVocabSize() == 6 and d.getDiscardSize() > 0
    assert d.getIndex("the") < 6 and d.getIndex("zzz") == 6
    ls = list(TX.TextToLabeledSentence(d)(iter(toks)))
    assert ls[0].dataLength() == len(toks[0]) - 1 and ls[0].getLabel(0) == ls[0].getData(1)
    samples = list(TX.LabeledSentenceToSample(7, fix_data_length=12, fix_label_length=12)(iter(ls)))
    f, l = samples[1].feature(), samples[1].label()
    assert tuple(f.shape) == (12, 7) and torch.all(f.sum(1) == 1)
    assert l.min() >= 1                                     # labels are 1-based
    d.save(str(tmp_path))
    d2 = TX.Dictionary(str(tmp_path))
    assert d2.word2Index() == d.word2Index() and d2.discardVocab() == d.discardVocab()
    assert TX.SentenceSplitter()(iter(["Hi there. How are you? Fine!"])).__next__() == ["Hi there.", "How are you?", "Fine!"]


def test_rle_masks_and_coco(tmp_path):
    m = np.zeros((5, 6), dtype=np.uint8)
    m[1:4, 2:5] = 1
    rle = MaskUtils.binaryToRLE(m)
    assert rle.counts.sum() == 30 and MaskUtils.rleArea(rle) == 9
    assert np.array_equal(MaskUtils.RLE2Binary(rle), m)
    s = MaskUtils.RLE2String(rle)
    back = MaskUtils.string2RLE(s, 5, 6)
    assert back.counts.tolist() == rle.counts.tolist()
    m2 = np.zeros((5, 6), dtype=np.uint8)
    m2[2:5, 3:6] = 1
    assert abs(MaskUtils.rleIOU(rle, MaskUtils.binaryToRLE(m2)) - 4 / 14) < 1e-6
    poly = PolyMasks([[2, 1, 4, 1, 4, 3, 2, 3]], 5, 6).toRLE()
    assert MaskUtils.rleArea(poly) == 4                     # COCO rule: pixel centres inside [2,4) x [1,3)
    ann = {"images": [{"id": 7, "height": 5, "width": 6, "file_name": "a.jpg"}],
           "categories": [{"id": 18, "name": "dog"}, {"id": 3, "name": "car"}],
           "annotations": [{"id": 1, "image_id": 7, "category_id": 18, "bbox": [2, 1, 3, 3], "area": 9, "iscrowd": 0,
                            "segmentation": [[2, 1, 4, 1, 4, 3, 2, 3]]},
                           {"id": 2, "image_id": 7, "category_id": 3, "bbox": [2, 1, 3, 3], "area": 9, "iscrowd": 1,
                            "segmentation": {"counts": s, "size": [5, 6]}}]}
    p = tmp_path / "ann.json"
    p.write_text(json.dumps(ann))
    ds = COCODataset.load(str(p), "/imgs")
    assert ds.categoryId2Idx == {18: 1, 3: 2} and len(ds.images[0].annotations) == 2        # json order, 1-based
    assert isinstance(ds.images[0].annotations[1].segmentation, RLEMasks)


def test_row_transformer():
    rows = [{"a": 1, "b": 2.5, "c": 3}, {"a": 4, "b": 5.5, "c": 6}]
    out = list(RowTransformer.numeric("feat", ["a", "c"])(iter(rows)))
    assert out[1]["feat"].tolist() == [4.0, 6.0]
    at = list(RowTransformer.atomic(["b"])(iter(rows)))
    assert at[0]["b"].tolist() == [2.5]
    with pytest.raises(ValueError):
        from bigdl_b200.dataset.datamining import ColToTensor
        RowTransformer([ColToTensor("k", "a"), ColToTensor("k", "b")])


def test_vision_frame_transformers(tmp_path):
    import cv2

    imgs = _imgs(4, 60, 80)
    for i, im in enumerate(imgs):
        cv2.imwrite(str(tmp_path / f"{i}.png"), im)
    frame = V.ImageFrame.read(str(tmp_path))
    assert frame.is_local() and len(frame) == 4 and np.array_equal(frame.features[0]["mat"], imgs[0])
    RNG.setSeed(3)
    pipe = V.Resize(48, 64) >> V.CenterCrop(32, 32) >> V.HFlip() >> V.ChannelNormalize(100, 110, 120, 2, 2, 2) >> \
        V.MatToTensor(to_rgb=True) >> V.ImageFrameToSample()
    frame.set_label([[1.0], [2.0], [1.0], [2.0]])
    frame = frame.transform(pipe)
    t = frame.features[0]["imageTensor"]
    assert tuple(t.shape) == (3, 32, 32)
    ref = cv2.resize(imgs[0], (64, 48), interpolation=1)[8:40, 16:48][:, ::-1].astype(np.float32)
    ref = ((ref - np.array([120, 110, 100], np.float32)) / 2)[..., ::-1].transpose(2, 0, 1)
    assert np.allclose(t.numpy(), ref, atol=1e-4)
    assert frame.toDataSet().size() == 4
    # a failing transformer invalidates one feature, not the frame
    bad = V.LocalImageFrame([V.ImageFeature(b"not an image")]).transform(V.BytesToMat())
    assert not bad.features[0].isValid
    # photometric + geometric augmentations keep shape/dtype
    f = V.ImageFeature(imgs[1].copy())
    for tr in (V.Brightness(-20, 20), V.Contrast(0.8, 1.2), V.Saturation(0.8, 1.2), V.Hue(-10, 10), V.ChannelOrder(),
               V.ColorJitter(shuffle=True), V.Expand(max_expand_ratio=2.0), V.RandomCrop(30, 30),
               V.RandomAlterAspect(crop_length=24), V.Filler(0, 0, 0.5, 0.5, 0), V.AspectScale(40, max_size=50),
               V.RandomResize(30, 40), V.FixExpand(64, 64)):
        f = tr.transform(f)
        assert f.isValid and f["mat"].ndim == 3, type(tr).__name__
    # roi label bookkeeping
    g = V.ImageFeature(imgs[0].copy(), None)
    g["label"] = V.RoiLabel([1, 2], [[8, 6, 40, 30], [16, 12, 72, 54]])
    g = (V.RoiNormalize() >> V.HFlip() >> V.RoiHFlip()).transform(g)
    assert np.allclose(g["label"].bboxes[0], [1 - 40 / 80, 6 / 60, 1 - 8 / 80, 30 / 60])
    a, b = frame.random_split([0.5, 0.5])
    assert len(a) + len(b) == 4
    batches = list(V.MTImageFeatureToBatch(32, 32, 3, to_rgb=False)(
        V.LocalImageFrame([V.ImageFeature(im[:32, :32].copy(), [float(i + 1)]) for i, im in enumerate(imgs)])))
    assert [bt.size() for bt in batches] == [3, 1] and batches[0].getTarget().tolist() == [1.0, 2.0, 3.0]


def test_gpu_image_batcher_cpu_path_is_deterministic_and_normalised():
    import torch
    from bigdl_b200.ops.data import GPUImageBatcher, image_batch
    from bigdl_b200.utils.random_generator import RNG

    frames = torch.randint(0, 256, (6, 36, 40, 3), dtype=torch.uint8)
    RNG.setSeed(3)
    a = GPUImageBatcher(32, 32, (100, 110, 120), (50, 55, 60), train=True, dtype=torch.float32)(frames)
    RNG.setSeed(3)
    b = GPUImageBatcher(32, 32, (100, 110, 120), (50, 55, 60), train=True, dtype=torch.float32)(frames)
    assert a.shape == (6, 3, 32, 32) and torch.equal(a, b)
    c = image_batch(frames, 32, 32, (100, 110, 120), (50, 55, 60), dtype=torch.float32)
    ref = (frames[:, 2:34, 4:36].float() - torch.tensor([100., 110., 120.])) / torch.tensor([50., 55., 60.])
    assert torch.allclose(c, ref.permute(0, 3, 1, 2))


def test_scale_resize_rule_and_aliases():
    import numpy as np
    from bigdl_b200.dataset.base import CachedDistriDataSet, DistributedDataSet
    from bigdl_b200.transform.vision.image import ScaleResize

    t = ScaleResize(800, 1333)
    assert t._size(480, 640) == (800, 1066)        # shorter side to 800
    assert t._size(300, 1200) == (333, 1332)       # capped by the longer side
    assert t._size(800, 1000) == (800, 1000)       # already there
    f = {"mat": np.zeros((48, 64, 3), np.uint8)}
    ScaleResize(24).transformMat(f)
    assert f["mat"].shape[:2] == (24, 32)
    assert CachedDistriDataSet is DistributedDataSet
