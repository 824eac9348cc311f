"""The reference's own DataSetSpec expectations on its own fixtures (T/dataset/DataSetSpec.scala:49-200):
COCO json, COCO → SequenceFile → back, CIFAR image folders, the ImageNet folder, MNIST labels."""
import glob
import os

import numpy as np
import pytest

RES = "/root/reference/spark/dl/src/test/resources"
pytestmark = pytest.mark.skipif(not os.path.isdir(RES), reason="reference tree not available")


def test_coco_dataset_loads_like_the_reference():
    from bigdl_b200.dataset.segmentation import COCODataset, COCOPoly, COCORLE

    ds = COCODataset.load(f"{RES}/coco/cocomini.json")
    assert len(ds.images) == 5 and len(ds.annotations) == 6
    idx = ds.categoryId2Idx
    look = (lambda c: idx[c]) if isinstance(idx, dict) else idx
    assert [look(a.categoryId) for a in ds.annotations] == [53, 53, 53, 1, 19, 1]
    assert [(i.height, i.width) for i in ds.images] == [(428, 640), (480, 640), (427, 640), (480, 640), (427, 640)]
    for a in ds.annotations:
        assert a.image.id == a.imageId
        if a.isCrowd:
            assert isinstance(a.segmentation, COCORLE)
        else:
            assert isinstance(a.segmentation, COCOPoly)
            assert (a.segmentation.height, a.segmentation.width) == (a.image.height, a.image.width)
    for i in range(1, len(ds.categories) + 1):
        assert look(ds.getCategoryByIdx(i).id) == i


def test_coco_sequence_files_round_trip(tmp_path):
    import cv2

    from bigdl_b200.cli.coco_seqfile_generator import main, read_coco_seqfile
    from bigdl_b200.dataset.segmentation import COCODataset, PolyMasks, RLEMasks

    ds = COCODataset.load(f"{RES}/coco/cocomini.json")
    index = {im.fileName: im for im in ds.images}
    main(["-f", f"{RES}/coco/", "-o", str(tmp_path), "-p", "4", "-b", "2", "-m", f"{RES}/coco/cocomini.json"])
    files = sorted(glob.glob(str(tmp_path / "*.seq")))
    assert len(files) == 3                                  # 5 images in blocks of 2
    seen = 0
    for f in files:
        for name, h, w, anns, data in read_coco_seqfile(f):
            img = index[name]
            seen += 1
            assert (h, w) == (img.height, img.width) and len(anns) == len(img.annotations)
            assert [1.0 if a.isCrowd else 0.0 for a in anns] == [1.0 if a.isCrowd else 0.0 for a in img.annotations]
            for got, want in zip(anns, img.annotations):
                gm, wm = got.masks if hasattr(got, "masks") else got.segmentation, want.segmentation
                if isinstance(wm, RLEMasks):
                    assert list(gm.counts) == list(wm.counts)
                else:
                    assert isinstance(gm, PolyMasks)
                    assert np.allclose(np.concatenate([np.ravel(p) for p in gm.poly]),
                                       np.concatenate([np.ravel(p) for p in wm.poly]))
                gb = [got.bbox1, got.bbox2, got.bbox3, got.bbox4] if hasattr(got, "bbox1") else list(got.bbox)
                assert np.allclose(gb, list(want.bbox))
            decoded = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
            assert decoded.shape == (img.height, img.width, 3)
            assert data == open(f"{RES}/coco/{name}", "rb").read()
    assert seen == 5


def test_cifar_folder_pixels_match_the_reference_spec():
    from bigdl_b200.dataset.image import BGRImage, LocalImageFiles, LocalImgReader

    labels = LocalImageFiles.readLabels(f"{RES}/cifar")
    assert labels["airplane"] == 1 and labels["deer"] == 2
    paths = LocalImageFiles.readPaths(f"{RES}/cifar")
    assert len(paths) == 7
    imgs = {os.path.basename(p.path): im for p, im in zip(paths, LocalImgReader(BGRImage.NO_SCALE, normalize=255.0)(iter(paths)))}
    assert sorted(im.label() for im in imgs.values()) == [1.0] * 3 + [2.0] * 4
    px = lambda im, x, y: [round(float(v) * 255) for v in im.content()[y, x]]  # noqa: E731  (B, G, R)
    # "img1" of the spec: first pixel (59, 125, 234), pixel (22, 4) = (31, 148, 253)
    first = [n for n, im in imgs.items() if px(im, 0, 0) == [59, 125, 234]]
    assert len(first) == 1 and imgs[first[0]].label() == 1.0 and px(imgs[first[0]], 22, 4) == [31, 148, 253]
    # "img4": a deer whose pixel (9, 8) is (37, 51, 40)
    deer = [n for n, im in imgs.items() if px(im, 9, 8) == [37, 51, 40]]
    assert len(deer) == 1 and imgs[deer[0]].label() == 2.0
    assert all(im.height() == 32 and im.width() == 32 for im in imgs.values())


def test_imagenet_folder_and_mnist_labels():
    from bigdl_b200.dataset.image import LocalImageFiles, LocalImgReader
    from bigdl_b200.dataset import mnist

    labels = LocalImageFiles.readLabels(f"{RES}/imagenet")
    assert len(labels) == 4 and sorted(labels.values()) == [1, 2, 3, 4]
    paths = LocalImageFiles.readPaths(f"{RES}/imagenet")
    assert len(paths) == 11
    imgs = list(LocalImgReader(256, normalize=1.0)(iter(paths)))
    assert all(min(i.height(), i.width()) == 256 for i in imgs)          # shorter side scaled to 256

    with open(f"{RES}/mnist/t10k-labels.idx1-ubyte", "rb") as f:
        y = mnist.extract_labels(f)
    assert len(y) == 10000 and int(y.min()) == 0 and int(y.max()) == 9   # +1 shift happens in the record loader


def test_coco_sequence_files_to_roi_image_frame(tmp_path):
    from bigdl_b200.cli.coco_seqfile_generator import main
    from bigdl_b200.dataset.segmentation import COCODataset
    from bigdl_b200.transform.vision import image as V

    ds = COCODataset.load(f"{RES}/coco/cocomini.json")
    index = {im.fileName: im for im in ds.images}
    main(["-f", f"{RES}/coco/", "-o", str(tmp_path), "-b", "2", "-m", f"{RES}/coco/cocomini.json"])
    frame = V.SeqFileFolder.filesToRoiImageFrame(str(tmp_path))
    assert len(frame.features) == 5
    for f in frame.features:
        img = index[f["uri"]]
        assert tuple(f["originalSize"]) == (img.height, img.width, 3)
        lab = f["label"]
        assert len(lab.masks) == len(img.annotations) and lab.bboxes.shape == (len(img.annotations), 4)
        assert f[V.RoiLabel.ISCROWD].tolist() == [1.0 if a.isCrowd else 0.0 for a in img.annotations]
        assert len(f["bytes"]) == img.height * img.width * 3
        for row, a in zip(lab.bboxes, img.annotations):
            assert np.allclose(row, list(a.bbox))
    # the frame feeds the detection batcher directly
    batch = next(iter(V.RoiImageFeatureToBatchWithResize(32, 2, V.Pipeline([V.PixelBytesToMat(), V.MatToTensor()]))(frame)))
    assert batch.getInput().shape[0] == 2 and batch.getInput().shape[2] % 32 == 0 and len(batch.target) == 2


def _pascal():
    from bigdl_b200.transform.vision import image as V

    frame = V.ImageFrame.read(f"{RES}/pascal")
    V.BytesToMat().transform(frame.features[0])
    return V, frame.features[0]


def test_vision_specs_on_the_pascal_image():
    """Sizes asserted by T/transform/vision/image/augmentation/{Resize,Crop,Expand,RandomCropper}Spec.scala."""
    from bigdl_b200.transform.vision.bbox import OpenCVMat

    V, f = _pascal()
    assert (f.getHeight(), f.getWidth()) == (375, 500)
    size = lambda t: (lambda g: (g.getHeight(), g.getWidth()))(t.transform(f.clone()))  # noqa: E731
    assert size(V.Resize(300, 300)) == (300, 300)
    assert size(V.AspectScale(750, max_size=3000)) == (750, 1000)
    assert size(V.RandomAspectScale([750], max_size=3000)) == (750, 1000)
    img = OpenCVMat.read(f"{RES}/pascal/000025.jpg")
    assert V.AspectScale.getHeightWidthAfterRatioScale(img, 600, 1000, 1) == (600, 800)
    assert size(V.ScaleResize(100, 120, False)) == (90, 120)
    assert size(V.CenterCrop(50, 50)) == (50, 50) and size(V.RandomCrop(200, 200)) == (200, 200)
    assert size(V.FixedCrop(0, 0, 50, 50, False)) == (50, 50)
    assert size(V.FixedCrop(0, 0, 50 / 500.0, 50 / 375.0, True)) == (50, 50)
    assert size(V.FixedCrop(0, 0, 600.0, 700.0, True)) == (375, 500)                    # clipped to the image
    g = f.clone()
    g["roi"] = np.array([1, 1, 0.2, 0, 0, 0.5, 0.5], dtype=np.float32)
    V.DetectionCrop("roi").transform(g)
    assert (g.getHeight(), g.getWidth()) == (187, 250)
    assert size(V.Expand(min_expand_ratio=2, max_expand_ratio=2)) == (750, 1000)
    assert size(V.FixExpand(600, 600)) == (600, 600)
    assert size(V.RandomCropper(224, 224, False, "Random")) == (224, 224)
    assert size(V.RandomResize(375, 375)) == (375, 500)                                  # shorter side already 375


def test_image_frame_specs_read_grey_and_globs():
    """T/transform/vision/image/{ImageFrameSpec,ConvertorSpec}.scala."""
    import torch

    from bigdl_b200.transform.vision import image as V

    assert len(V.ImageFrame.read(f"{RES}/pascal/").features) == 1
    assert len(V.ImageFrame.read(f"{RES}/pascal/*.jpg").features) == 1
    one = V.ImageFrame.read(f"{RES}/pascal/000025.jpg").features[0]
    assert one["uri"].endswith("000025.jpg") and len(one["bytes"]) == 95959 and one.getSize() == (375, 500, 3)
    assert len(V.ImageFrame.read(f"{RES}/pascal/0000251.jpg").features) == 0

    g = V.ImageFrame.read(f"{RES}/gray/gray.bmp").features[0]
    assert g.getOriginalSize() == (50, 50, 1)
    V.Pipeline([V.Resize(28, 28), V.MatToTensor()]).transform(g)
    assert g.getOriginalSize() == (50, 50, 1) and g.getSize() == (28, 28, 1) and tuple(g["imageTensor"].shape) == (1, 28, 28)

    grey = V.ImageFrame.read(f"{RES}/grey").features[0]
    V.MatToTensor(greyToRGB=True).transform(grey)
    t = grey["imageTensor"]
    assert grey.getSize() == (500, 357, 1) and tuple(t.shape) == (3, 500, 357)
    assert torch.equal(t[0], t[1]) and torch.equal(t[0], t[2])
    forced = V.ImageFeature(open(f"{RES}/gray/gray.bmp", "rb").read())
    V.BytesToMat(image_codec=1).transform(forced)
    assert forced.getSize() == (50, 50, 3)

    # MatToFloats + toTensor equals MatToTensor (ConvertorSpec "toTensor")
    a = V.ImageFrame.read(f"{RES}/pascal").features[0]
    b = a.clone()
    V.MatToFloats().transform(a)
    V.MatToTensor().transform(b)
    assert np.array_equal(np.asarray(a["floats"]).reshape(375, 500, 3).transpose(2, 0, 1), b["imageTensor"].numpy())


_BOXES = [2.0, 84.0, 59.0, 248.0, 68.0, 115.0, 233.0, 279.0, 64.0, 173.0, 377.0, 373.0, 320.0, 2.0, 496.0, 375.0,
          221.0, 4.0, 341.0, 374.0, 135.0, 14.0, 220.0, 148.0, 69.0, 43.0, 156.0, 177.0, 58.0, 54.0, 104.0, 139.0,
          279.0, 1.0, 331.0, 86.0, 320.0, 22.0, 344.0, 96.0, 337.0, 1.0, 390.0, 107.0]
_CLASSES = [11.0, 11.0, 11.0, 16.0, 16.0, 16.0, 11.0, 16.0, 16.0, 16.0, 16.0] + [0.0] * 10 + [1.0]


def _roi_feature():
    from bigdl_b200.transform.vision import image as V

    f = V.ImageFrame.read(f"{RES}/pascal").features[0]
    f["label"] = V.RoiLabel(np.array(_CLASSES, np.float32).reshape(2, 11), np.array(_BOXES, np.float32).reshape(11, 4))
    return V, f


def test_roi_transformers_match_the_reference_spec():
    """Expected tensors of T/transform/vision/image/label/roi/RoiTransformerSpec.scala."""
    V, f = _roi_feature()
    V.RoiNormalize().transform(f)
    norm = f["label"].bboxes
    assert np.allclose(norm[0], [0.004, 0.224, 0.118, 0.6613333], atol=1e-6)
    assert np.allclose(norm[3], [0.64, 0.0053333, 0.992, 1.0], atol=1e-6)
    assert np.allclose(norm[10], [0.674, 0.0026667, 0.78, 0.28533334], atol=1e-6)
    assert f["label"].classes.reshape(-1).tolist() == _CLASSES

    V, f = _roi_feature()
    V.Pipeline([V.HFlip(), V.RoiHFlip(False)]).transform(f)
    want = [441.0, 84.0, 498.0, 248.0, 267.0, 115.0, 432.0, 279.0, 123.0, 173.0, 436.0, 373.0, 4.0, 2.0, 180.0, 375.0,
            159.0, 4.0, 279.0, 374.0, 280.0, 14.0, 365.0, 148.0, 344.0, 43.0, 431.0, 177.0, 396.0, 54.0, 442.0, 139.0,
            169.0, 1.0, 221.0, 86.0, 156.0, 22.0, 180.0, 96.0, 110.0, 1.0, 163.0, 107.0]
    assert f["label"].bboxes.reshape(-1).tolist() == want

    V, f = _roi_feature()
    V.Pipeline([V.RoiNormalize(), V.RoiHFlip()]).transform(f)
    assert np.allclose(f["label"].bboxes[0], [0.882, 0.224, 0.996, 0.6613333], atol=1e-6)
    assert np.allclose(f["label"].bboxes[3], [0.008, 0.0053333, 0.36, 1.0], atol=1e-6)

    V, f = _roi_feature()
    V.Resize(300, 300).transform(f)
    # Scala's RoiResize() defaults to normalized = false (the Python API's default is True): boxes / 300 · 300 / 500
    V.Pipeline([V.RoiNormalize(), V.RoiResize(False)]).transform(f)
    assert np.allclose(f["label"].bboxes, norm, atol=1e-6)

    V, f = _roi_feature()
    V.Resize(300, 300).transform(f)
    V.RoiResize(False).transform(f)
    assert np.allclose(f["label"].bboxes[0], [1.2, 67.2, 35.4, 198.4], atol=1e-4)
    assert np.allclose(f["label"].bboxes[10], [202.2, 0.8, 234.0, 85.6], atol=1e-4)

    V, f = _roi_feature()
    V.Pipeline([V.CenterCrop(300, 300), V.RoiNormalize(), V.RoiProject()]).transform(f)
    expected = np.array([0.0, 0.25833336, 0.44333336, 0.805, 0.0, 0.45166665, 0.9233333, 1.0, 0.40333334, 0.0, 0.8033333,
                         1.0, 0.116666675, 0.0, 0.4, 0.36833334, 0.0, 0.01833333, 0.1866667, 0.46500003, 0.5966667, 0.0,
                         0.77, 0.16166666, 0.73333335, 0.0, 0.8133333, 0.195, 0.78999996, 0.0, 0.9666667, 0.23166668],
                        np.float32).reshape(8, 4)
    assert f["label"].bboxes.shape == (8, 4) and np.allclose(f["label"].bboxes, expected, atol=1e-5)
    assert f["label"].classes.shape == (2, 8) and f["mat"].shape == (300, 300, 3)
