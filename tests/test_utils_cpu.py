import time

import pytest
import torch

from bigdl_b200.utils.directed_graph import Node
from bigdl_b200.utils.file import File
from bigdl_b200.utils.shape import MultiShape, Shape, SingleShape
from bigdl_b200.utils.thread_pool import ThreadPool
from bigdl_b200.utils.util import Crc32, HashFunc, LayerException, kthLargest, shift


def test_shape():
    s = Shape([2, 3])
    assert isinstance(s, SingleShape) and s.toSingle() == [2, 3]
    assert s.copyAndUpdate(-1, 5).toSingle() == [2, 5]
    m = Shape([Shape([1, 2]), [3]])
    assert isinstance(m, MultiShape) and m.toMulti()[1] == SingleShape([3])
    with pytest.raises(ValueError):
        s.toMulti()


def test_file_roundtrip(tmp_path):
    p = str(tmp_path / "a" / "obj.bin")
    File.save({"w": torch.ones(3)}, p)
    assert torch.equal(File.load(p)["w"], torch.ones(3))
    with pytest.raises(FileExistsError):
        File.save(1, p)
    File.save(2, p, isOverwrite=True)
    assert File.load("file://" + p) == 2
    File.saveBytes(b"xyz", str(tmp_path / "b.bin"))
    assert File.readBytes(str(tmp_path / "b.bin")) == b"xyz"
    with pytest.raises(IOError):
        File.load("hdfs://nn/path")


def test_misc_utils():
    assert kthLargest([3, 1, 2], 1) == 3
    assert shift([1, 2, 3, 4], 0, 2) == [2, 3, 1, 4]
    assert HashFunc.murmurhash3_32(b"", 0) == 0
    assert HashFunc.murmurhash3_32(b"hello", 0) == 0x248BFA47          # public MurmurHash3 test vector
    assert HashFunc.murmurhash3_32("The quick brown fox jumps over the lazy dog", 0x9747B28C) == 0x2FA826CD
    assert Crc32.crc32c(b"123456789") == 0xE3069283
    e = LayerException("Sequential/Linear[fc1]", ValueError("bad"))
    assert "fc1" in str(e) and isinstance(e.error, ValueError)


def test_thread_pool_timeout_marks_stragglers():
    tp = ThreadPool(4)
    assert tp.invokeAndWait([lambda i=i: i * i for i in range(5)]) == [0, 1, 4, 9, 16]
    futs = tp.invokeAndWait2([lambda: 1, lambda: time.sleep(0.5) or 2], timeout=0.1)
    assert [f.abandoned for f in futs] == [False, True]
    tp.sync(tp.invoke([lambda: None]))
    tp.shutdown()


def test_directed_graph():
    a, b, c, d = Node("a"), Node("b"), Node("c"), Node("d")
    a >> b >> d
    a >> c >> d
    g = a.graph()
    order = [n.element for n in g.topologySort()]
    assert order[0] == "a" and order[-1] == "d" and g.size() == 4 and g.edges() == 4
    assert [n.element for n in d.graph(reverse=True).topologySort()][0] == "d"
    d >> a
    with pytest.raises(ValueError):
        g.topologySort()
    d.delete(a)
    assert g.cloneGraph().size() == 4


def test_nn_shape_helpers_follow_the_reference_rules():
    from bigdl_b200.nn import utils as U
    from bigdl_b200.nn.layer import Linear, Sequential

    # Inception pool1: 112 -> 56 with 3x3/2 ceil mode, no padding
    assert U.getOutSizeAndPadding(112, 112, 2, 2, 3, 3, 0, 0, True) == [0, 0, 0, 0, 56, 56]
    assert U.getOutSizeAndPadding(112, 112, 2, 2, 3, 3, 0, 0, False)[-2:] == [55, 55]
    # padded ceil mode drops a window that would start in the padding
    assert U.getOutSizeAndPadding(4, 4, 2, 2, 2, 2, 1, 1, True)[-2:] == [3, 3]
    assert U.getSAMEOutSizeAndPadding(224, 224, 2, 2, 7, 7) == [2, 3, 2, 3, 112, 112]
    assert U.getSAMEOutSizeAndPadding(5, 5, 1, 1, 3, 3, 4, 1, 3)[-3:] == [4, 5, 5]
    assert U.getPaddingAndOutputSize(112, 112, 2, 2, 3, 3, 0, 0, True) == (0, 1, 0, 1, 56, 56)
    assert U.getOutputShape(7, 9, 16) == [16, 7, 9]
    assert U.getOutputShape(7, 9, 16, 4, "NHWC") == [4, 7, 9, 16]
    m = Sequential().add(Linear(4, 4)).add(Linear(4, 2))
    assert not U.isLayerwiseScaled(m)
    m.modules[1].setScaleW(0.5)
    assert U.isLayerwiseScaled(m)
    assert "Reshape" in U.ErrorInfo.constrainInputAsVectorOrBatch


def test_reflection_utils_roundtrip_and_dnn_tensor():
    import torch
    from bigdl_b200.nn.layer import SpatialConvolution
    from bigdl_b200.tensor import DnnTensor, Tensor
    from bigdl_b200.utils.reflection import ReflectionUtils as R

    conv = SpatialConvolution(3, 8, 5, 5, 2, 2, 1, 1)
    op, args = R.reflectToIR(conv)
    assert op == "SpatialConvolution" and args["n_output_plane"] == 8
    assert R.getPrimCtorParams(SpatialConvolution)[:2] == ["n_input_plane", "n_output_plane"]
    twin = R.reflectFromIR(op, args)
    assert twin.weight.shape == conv.weight.shape and (twin.stride_h, twin.pad_w) == (2, 1)

    before = DnnTensor.liveCount()
    a, b = DnnTensor(2, 3, device="cpu"), DnnTensor([2, 3], device="cpu")
    a.copy(Tensor(torch.arange(6.).reshape(2, 3)))
    b.copy(a).axpby(2.0, 0.5, a)          # b = 2a + 0.5b = 2.5a
    assert torch.allclose(b.dense().t, 2.5 * torch.arange(6.).reshape(2, 3))
    assert a.size() == [2, 3] and a.size(2) == 3 and a.nElement() == 6 and DnnTensor.liveCount() == before + 2
    a.release(); a.release()
    assert a.isReleased() and DnnTensor.liveCount() == before + 1
    try:
        a.zero()
        raise AssertionError("released tensor must refuse work")
    except RuntimeError:
        pass


def test_detection_map_voc_and_coco():
    import torch
    from bigdl_b200.optim.validation import (MAPCOCO, MAPPascalVoc2007, MeanAveragePrecision,
                                              MeanAveragePrecisionObjectDetection)

    # image 0: two class-1 objects (second one "difficult"), one class-2 object; image 1: one class-1 object
    target = [
        {"bboxes": torch.tensor([[10., 10, 50, 50], [100, 100, 150, 150], [200, 200, 260, 260]]),
         "classes": torch.tensor([[1., 1, 2], [0, 1, 0]]), "is_crowd": torch.zeros(3)},
        {"bboxes": torch.tensor([[20., 20, 80, 80]]), "classes": torch.tensor([1.]), "is_crowd": torch.zeros(1)},
    ]
    output = [
        {"bboxes": torch.tensor([[12., 11, 50, 52], [11, 10, 49, 50], [100, 100, 150, 150], [205, 200, 260, 255],
                                 [300, 300, 340, 340]]),
         "classes": torch.tensor([1., 1, 1, 2, 2]), "scores": torch.tensor([0.9, 0.8, 0.7, 0.95, 0.3])},
        {"bboxes": torch.tensor([[300., 300, 350, 350]]), "classes": torch.tensor([1.]), "scores": torch.tensor([0.85])},
    ]
    m = MeanAveragePrecisionObjectDetection(3, skipClass=0)
    r = m(output, target)
    # class 1: 2 countable GTs; ranked hits: 0.9 TP, 0.85 FP, 0.8 FP (duplicate), 0.7 ignored (difficult)
    assert r.gt_cnt == [0, 2, 1]
    assert sorted(r.predict[1], reverse=True) == [(0.9, True), (0.85, False), (0.8, False)] or \
        [round(s, 2) for s, _ in sorted(r.predict[1], reverse=True)] == [0.9, 0.85, 0.8]
    ap1 = r.class_ap(1)           # recall 0.5 at precision 1.0, recall 1.0 never reached -> (1.0 + 0) / 2
    assert abs(ap1 - 0.5) < 1e-6
    assert abs(r.class_ap(2) - 1.0) < 1e-6      # 0.95 TP first, then one FP
    assert abs(r.result()[0] - 0.75) < 1e-6
    # merging two halves equals one pass
    a = m(output[:1], target[:1]) + m(output[1:], target[1:])
    assert abs(a.result()[0] - 0.75) < 1e-6
    v07 = MeanAveragePrecisionObjectDetection(3, theType=MAPPascalVoc2007, skipClass=0)(output, target)
    assert abs(v07.class_ap(1) - 6 / 11) < 1e-6  # precision 1 for the recall points 0 .. 0.5
    coco = MeanAveragePrecision.cocoBBox(3)(output, target)
    assert 0.0 < coco.result()[0] <= 1.0 and len(coco.impl) == 10
    # tensor form of the detections: [n, (label, score, x1, y1, x2, y2) * n]
    row0 = [5.] + sum(([float(c), float(s)] + [float(v) for v in b] for b, c, s in
                       zip(output[0]["bboxes"], output[0]["classes"], output[0]["scores"])), [])
    row1 = [1., 1., 0.85, 300, 300, 350, 350] + [0.] * 24
    t = MeanAveragePrecisionObjectDetection(3, skipClass=0)(torch.tensor([row0, row1]), target)
    assert abs(t.result()[0] - 0.75) < 1e-6
    assert MeanAveragePrecisionObjectDetection(3, theType=MAPCOCO, skipClass=0)(output, target).class_ap(0) == -1.0


def test_profile_layers_reports_every_leaf_once():
    import torch
    from bigdl_b200.nn.layer import Linear, ReLU, Sequential
    from bigdl_b200.utils.profiling import enable_nvtx, format_profile, profile_layers

    m = Sequential().add(Linear(8, 16).setName("fc1")).add(ReLU().setName("act")).add(Linear(16, 4).setName("fc2"))
    rows = profile_layers(m, torch.randn(5, 8), warmup=1, iters=2)
    assert [r[0] for r in rows] == ["fc1", "act", "fc2"] and all(r[2] >= 0 for r in rows)
    assert "fc2" in format_profile(rows)
    assert "updateOutput" not in m.modules[0].__dict__          # wrappers removed
    enable_nvtx(True)                                           # no GPU here: stays off, forward still works
    assert m.forward(torch.randn(2, 8)).shape == (2, 4)
    enable_nvtx(False)


def test_thread_pool_file_bytes_and_dlmodel_feature_size(tmp_path):
    from bigdl_b200.utils.file import File
    from bigdl_b200.utils.thread_pool import ThreadPool

    pool = ThreadPool(2)
    assert pool.getPoolSize() == 2
    pool.setPoolSize(3)
    assert pool.getPoolSize() == 3
    pool.setMKLThread(1)
    assert sorted(pool.invokeAndWait([lambda i=i: i * i for i in range(5)])) == [0, 1, 4, 9, 16]
    p = str(tmp_path / "blob.bin")
    File.save_bytes(b"\x00\x01payload", p, True)
    assert File.read_bytes(p) == b"\x00\x01payload"

    import pandas as pd
    from bigdl_b200.dlframes import DLModel
    from bigdl_b200.nn.layer import Linear

    m = DLModel(Linear(3, 2), [3])
    assert m.getFeatureSize() == [3]
    m.setFeatureSize([3]).setBatchSize(2)
    out = m.transform(pd.DataFrame({"features": [[1.0, 2.0, 3.0], [0.0, 0.0, 0.0]]}))
    assert len(out["prediction"]) == 2 and len(out["prediction"][0]) == 2


def test_file_remote_schemes_through_arrow_filesystems(tmp_path):
    """hdfs:// s3a:// ... resolve to Arrow filesystem clients (S/utils/File.scala:112-193); exercised offline with
    Arrow's in-memory filesystem registered for a scheme, and a sub-tree of the local disk standing in for a bucket."""
    import pyarrow.fs as pfs
    from datetime import datetime, timezone
    mem = pfs._MockFileSystem(datetime.now(timezone.utc))
    File.register_filesystem("mem", mem)
    File.save({"a": [1, 2, 3]}, "mem://models/run1/model.obj")
    assert File.load("mem://models/run1/model.obj") == {"a": [1, 2, 3]}
    assert File.exists("mem://models/run1/model.obj") and not File.exists("mem://models/none")
    with pytest.raises(FileExistsError):
        File.saveBytes(b"x", "mem://models/run1/model.obj")
    File.saveBytes(b"x", "mem://models/run1/model.obj", isOverwrite=True)
    assert File.readBytes("mem://models/run1/model.obj") == b"x"
    bucket = pfs.SubTreeFileSystem(str(tmp_path), pfs.LocalFileSystem())
    File.register_filesystem("s3a", bucket)
    File.saveBytes(b"weights", "s3a://ckpt/w.bin")
    assert (tmp_path / "ckpt" / "w.bin").read_bytes() == b"weights"
    with pytest.raises(ValueError):
        File.saveToHdfs(1, "s3a://ckpt/x", True)
    from bigdl_b200.utils import file as F
    F._ARROW_FS.pop("s3a"), F._ARROW_FS.pop("mem")


def test_module_save_load_on_a_remote_scheme(tmp_path):
    import pyarrow.fs as pfs
    from bigdl_b200.nn import layer as nn
    from bigdl_b200.nn.module import Module
    from bigdl_b200.utils import file as F
    File.register_filesystem("hdfs", pfs.SubTreeFileSystem(str(tmp_path), pfs.LocalFileSystem()))
    try:
        m = nn.Sequential().add(nn.Linear(4, 3)).add(nn.ReLU())
        m.saveModule("hdfs://models/m.bigdl", "hdfs://models/m.bin", overWrite=True)
        assert (tmp_path / "models" / "m.bigdl").exists() and (tmp_path / "models" / "m.bin").exists()
        m2 = Module.loadModule("hdfs://models/m.bigdl", "hdfs://models/m.bin")
        x = torch.randn(2, 4)
        assert torch.allclose(m.forward(x), m2.forward(x))
    finally:
        F._ARROW_FS.pop("hdfs", None)
