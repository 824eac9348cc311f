"""RLE / polygon mask utilities and the COCO record format (S/dataset/segmentation/{MaskUtils,COCODataset}.scala;
the reference's checks are in spark/dl/src/test/scala/.../dataset/segmentation/{MaskUtilsSpec,COCOSpec}.scala)."""
import json
import struct

import numpy as np
import pytest

from bigdl_b200.dataset.segmentation import (COCODataset, COCODeserializer, COCOSerializeContext, MaskUtils, PolyMasks,
                                             RLEMasks)


def _rand_mask(rng, h, w, p=0.4):
    return (rng.random((h, w)) < p).astype(np.uint8)


def test_set_operations_match_dense_masks():
    rng = np.random.default_rng(0)
    for _ in range(20):
        h, w = int(rng.integers(1, 12)), int(rng.integers(1, 12))
        ms = [_rand_mask(rng, h, w) for _ in range(3)]
        rs = [MaskUtils.binaryToRLE(m) for m in ms]
        for m, r in zip(ms, rs):
            assert r.counts.sum() == h * w and np.array_equal(MaskUtils.RLE2Binary(r), m)
            assert MaskUtils.rleArea(r) == int(m.sum())
        union = MaskUtils.mergeRLEs(rs, False)
        inter = MaskUtils.mergeRLEs(rs, True)
        assert np.array_equal(MaskUtils.RLE2Binary(union), ms[0] | ms[1] | ms[2])
        assert np.array_equal(MaskUtils.RLE2Binary(inter), ms[0] & ms[1] & ms[2])
        assert union == MaskUtils.binaryToRLE(ms[0] | ms[1] | ms[2])          # canonical counts, not just the same pixels
        i = float((ms[0] & ms[1]).sum())
        u = float((ms[0] | ms[1]).sum())
        assert MaskUtils.rleIOU(rs[0], rs[1], False) == pytest.approx(i / u if i else 0.0, abs=1e-6)
        assert MaskUtils.rleIOU(rs[0], rs[1], True) == pytest.approx(i / ms[0].sum() if i else 0.0, abs=1e-6)


def test_bbox_of_rle_and_box_iou():
    m = np.zeros((7, 9), dtype=np.uint8)
    m[2:5, 3:8] = 1
    m[3, 5] = 0
    r = MaskUtils.binaryToRLE(m)
    assert r.bbox == (3.0, 2.0, 7.0, 4.0)
    assert MaskUtils.binaryToRLE(np.zeros((4, 4))).bbox == (0.0, 0.0, 0.0, 0.0)
    full = MaskUtils.binaryToRLE(np.ones((4, 5)))
    assert full.bbox == (0.0, 0.0, 4.0, 3.0)
    assert MaskUtils.bboxIOU((0, 0, 9, 9), (5, 5, 14, 14), False) == pytest.approx(25 / 175)
    assert MaskUtils.bboxIOU((0, 0, 9, 9), (5, 5, 14, 14), True) == pytest.approx(25 / 100)
    assert MaskUtils.bboxIOU(0, 0, 9, 9, 20, 20, 29, 29, False) == 0.0


def test_polygon_rasterisation_follows_the_coco_rule():
    # axis-aligned rectangle with corners on pixel centres' edges: x in [2, 4], y in [1, 3] -> COCO fills 2 x 2 ... the
    # reference rule: a pixel column x is inside when its centre lies within the 5x-upsampled boundary walk
    r = PolyMasks([[1.5, 0.5, 4.5, 0.5, 4.5, 3.5, 1.5, 3.5]], 6, 7).toRLE()
    m = MaskUtils.RLE2Binary(r)
    want = np.zeros((6, 7), dtype=np.uint8)
    want[1:4, 2:5] = 1
    assert np.array_equal(m, want)
    # triangle: compare against a dense even-odd fill sampled at pixel centres (area within one boundary pixel row)
    tri = [1.0, 1.0, 18.0, 3.0, 6.0, 17.0]
    rt = PolyMasks([tri], 20, 20).toRLE()
    area = 0.5 * abs((18 - 1) * (17 - 1) - (6 - 1) * (3 - 1))
    assert abs(MaskUtils.rleArea(rt) - area) < 20
    bx = rt.bbox
    assert bx[0] >= 1 and bx[2] <= 18 and bx[1] >= 1 and bx[3] <= 17
    # two parts of one object are merged into a single RLE
    two = PolyMasks([[0.5, 0.5, 2.5, 0.5, 2.5, 2.5, 0.5, 2.5], [4.5, 4.5, 6.5, 4.5, 6.5, 6.5, 4.5, 6.5]], 8, 8)
    per = MaskUtils.poly2RLE(two, 8, 8)
    assert len(per) == 2 and MaskUtils.rleArea(two.toRLE()) == sum(MaskUtils.rleArea(p) for p in per)


def test_string_codec_round_trip_and_known_vector():
    rng = np.random.default_rng(1)
    for _ in range(10):
        r = MaskUtils.binaryToRLE(_rand_mask(rng, 30, 40, 0.5))
        back = MaskUtils.string2RLE(MaskUtils.RLE2String(r), 30, 40)
        assert back == r
    # long runs need several 5-bit groups, negative deltas set the sign bit
    r = RLEMasks([1000, 3, 2, 700, 295], 40, 50)
    assert MaskUtils.string2RLE(MaskUtils.RLE2String(r), 40, 50) == r


def test_coco_json_and_sequence_file_key(tmp_path):
    s = MaskUtils.RLE2String(RLEMasks([7, 9, 14], 5, 6))
    ann = {"info": {"year": 2017, "version": "1.0", "description": "toy"},
           "licenses": [{"id": 1, "name": "cc", "url": "u"}],
           "images": [{"id": 7, "height": 5, "width": 6, "file_name": "a.jpg", "license": 1}],
           "categories": [{"id": 18, "name": "dog", "supercategory": "animal"}, {"id": 3, "name": "car"}],
           "annotations": [{"id": 1, "image_id": 7, "category_id": 3, "bbox": [2, 1, 3, 3], "area": 9.0, "iscrowd": 0,
                            "segmentation": [[2, 1, 4, 1, 4, 3, 2, 3]]},
                           {"id": 2, "image_id": 7, "category_id": 18, "bbox": [1, 1, 2, 4], "area": 9.0, "iscrowd": 1,
                            "segmentation": {"counts": [7, 9, 14], "size": [5, 6]}},
                           {"id": 3, "image_id": 7, "category_id": 18, "bbox": [1, 1, 2, 4], "area": 9.0, "iscrowd": 1,
                            "segmentation": {"counts": s, "size": [5, 6]}}]}
    p = tmp_path / "ann.json"
    p.write_text(json.dumps(ann))
    (tmp_path / "a.jpg").write_bytes(b"\xff\xd8fake")
    ds = COCODataset.load(str(p), str(tmp_path))
    assert ds.info.year == 2017 and ds.licenses[0].name == "cc"
    assert ds.categoryId2Idx == {18: 1, 3: 2} and ds.getCategoryByIdx(1).superCategory == "animal"
    im = ds.getImageById(7)
    assert im.data == b"\xff\xd8fake" and len(im.annotations) == 3 and im.annotations[0].image is im
    assert im.annotations[0].bbox == (2, 1, 4, 3)                      # (x, y, w, h) -> inclusive corners
    assert im.annotations[1].segmentation == im.annotations[2].segmentation
    key = ds.imageKey(im)
    # layout: int32 name length, name, height, width, #annotations, ..., magic
    assert struct.unpack(">i", key[:4])[0] == 5 and key[4:9] == b"a.jpg"
    assert struct.unpack(">i", key[-4:])[0] == COCODataset.MAGIC_NUM
    name, h, w, anns = COCODataset.parseKey(key)
    assert (name, h, w, len(anns)) == ("a.jpg", 5, 6, 3)
    assert anns[0].categoryId == 2 and not anns[0].isCrowd and isinstance(anns[0].masks, PolyMasks)
    assert anns[0].masks.poly[0].tolist() == [2, 1, 4, 1, 4, 3, 2, 3]
    assert anns[1].isCrowd and anns[1].masks.counts.tolist() == [7, 9, 14]
    assert (anns[0].bbox1, anns[0].bbox2, anns[0].bbox3, anns[0].bbox4) == (2.0, 1.0, 4.0, 3.0)
    with pytest.raises(ValueError):
        COCODataset.parseKey(key[:-4] + b"\0\0\0\0")


def test_serialize_context_types():
    c = COCOSerializeContext()
    c.dump(3).dump(1.5).dump(True).dump("héllo").dumpLong(1 << 40).dumpDouble(2.25)
    d = COCODeserializer(c.toByteArray())
    assert d.getInt() == 3 and d.getFloat() == 1.5 and d.getBoolean() is True and d.getString() == "héllo"
    assert d.getLong() == 1 << 40 and d.getDouble() == 2.25


def test_coco_seqfile_generator_round_trip(tmp_path):
    from bigdl_b200.cli import coco_seqfile_generator as gen
    ann = {"images": [{"id": 1, "height": 4, "width": 4, "file_name": "x.jpg"},
                      {"id": 2, "height": 4, "width": 4, "file_name": "missing.jpg"}],
           "categories": [{"id": 5, "name": "cat"}],
           "annotations": [{"id": 1, "image_id": 1, "category_id": 5, "bbox": [0, 0, 2, 2], "area": 4.0, "iscrowd": 0,
                            "segmentation": [[0.5, 0.5, 2.5, 0.5, 2.5, 2.5, 0.5, 2.5]]}]}
    (tmp_path / "ann.json").write_text(json.dumps(ann))
    (tmp_path / "x.jpg").write_bytes(b"JPEGDATA")
    out = tmp_path / "seq"
    gen.main(["-f", str(tmp_path), "-m", str(tmp_path / "ann.json"), "-o", str(out), "-b", "10"])
    recs = list(gen.read_coco_seqfile(str(out / "coco-seq-0.seq")))
    assert len(recs) == 1                                       # the image without a file is skipped with a warning
    name, h, w, anns, data = recs[0]
    assert (name, h, w, data) == ("x.jpg", 4, 4, b"JPEGDATA")
    assert MaskUtils.rleArea(anns[0].masks.toRLE()) == 4
