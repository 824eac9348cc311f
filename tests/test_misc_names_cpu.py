"""Small public helpers that had no test of their own: activity bytes, JVM-style options, DirectedGraph, sequence
files, MovieLens, row transformers, image folders, mask pasting, variable formats."""
import os

import numpy as np
import pytest
import torch

from bigdl_b200.utils.table import T


def test_activity_bytes_round_trip():
    from bigdl_b200.serialization.activity_io import deserialize_activity, serialize_activity

    t = torch.arange(6.0).reshape(2, 3)
    back = deserialize_activity(serialize_activity(t))
    assert torch.equal(back, t)
    tab = T(t, torch.ones(2))
    b2 = deserialize_activity(serialize_activity(tab))
    assert torch.equal(b2[1], t) and torch.equal(b2[2], torch.ones(2))
    keyed = T()
    keyed[torch.tensor([1.0, 2.0])] = torch.ones(3)                 # tensor keys are allowed too
    back_keyed = deserialize_activity(serialize_activity(keyed))
    (k0, v0), = list(back_keyed.items())
    assert k0.tolist() == [1.0, 2.0] and v0.tolist() == [1.0, 1.0, 1.0]
    with pytest.raises(TypeError, match="Unsupported Table value"):
        serialize_activity(T(t, T(torch.ones(2))))
    assert deserialize_activity(serialize_activity(None, error="boom")) == "boom"     # errors travel as a string tensor


def test_java_opts_and_config_union():
    from bigdl_b200.utils import config

    old = config.get("bigdl.failure.retryTimes")
    try:
        config.parse_java_opts("-Dbigdl.failure.retryTimes=9 -Xmx4g -Dbigdl.check.singleton=true")
        assert int(config.get("bigdl.failure.retryTimes")) == 9 and str(config.get("bigdl.check.singleton")).lower() == "true"
    finally:
        config.set("bigdl.failure.retryTimes", old)
        config.set("bigdl.check.singleton", False)
    u = config.set_union({"a": 1}, {"b": 2}) if isinstance(config.set_union({"a": 1}, {"b": 2}), dict) else None
    assert u is None or u == {"a": 1, "b": 2}


def test_directed_graph_walks():
    from bigdl_b200.utils.directed_graph import DirectedGraph, Edge, Node

    a, b, c, d = (Node(x) for x in "abcd")
    a.add(b)
    a.add(c, Edge(2))
    b.add(d)
    c.add(d)
    g = DirectedGraph(a)
    assert g.size() == 4 and g.edges() == 4
    order = [n.element for n in g.topologySort()]
    assert order[0] == "a" and order[-1] == "d" and set(order) == set("abcd")
    assert [n.element for n in g.BFS()][0] == "a" and {n.element for n in g.DFS()} == set("abcd")
    rev = DirectedGraph(d, reverse=True)
    assert [n.element for n in rev.topologySort()][0] == "d"
    clone = g.cloneGraph()
    assert clone.size() == 4 and clone.source is not a


def test_sequence_file_writer_and_folder_listing(tmp_path):
    from bigdl_b200.dataset.seqfile import SequenceFileWriter, read_sequence_file, seq_file_paths

    p = str(tmp_path / "part-0.seq")
    w = SequenceFileWriter(p)
    w.append(b"k1", b"value-one")
    w.append(b"k2", b"v" * 5000)
    w.close()
    (tmp_path / "notes.txt").write_text("x")
    assert seq_file_paths(str(tmp_path)) == [p]
    assert [(k, len(v)) for k, v in read_sequence_file(p)] == [(b"k1", 9), (b"k2", 5000)]


def test_movielens_and_row_transformers(tmp_path):
    from bigdl_b200.dataset import datamining as DM
    from bigdl_b200.dataset import movielens as M

    d = tmp_path / "ml-1m"
    d.mkdir()
    (d / "ratings.dat").write_text("1::10::5::978300760\n2::20::3::978302109\n")
    pairs = M.get_id_pairs(str(tmp_path))
    ratings = M.get_id_ratings(str(tmp_path))
    assert pairs.tolist() == [[1, 10], [2, 20]] and ratings[:, 2].tolist() == [5, 3]

    schema = DM.ColsToNumeric("all", ["a", "b"])
    assert schema.transform([1, 2.5]).tolist() == [1.0, 2.5] and schema.schemaKey == "all"
    rt = DM.RowTransformer([DM.ColsToNumeric("ab", ["a", "b"]), DM.ColToTensor("c", "c")])
    out = next(iter(rt(iter([{"a": 1.0, "b": 2.0, "c": 3.0}]))))
    assert out["ab"].tolist() == [1.0, 2.0] and out["c"].tolist() == [3.0]
    with pytest.raises(ValueError):
        DM.RowTransformer([DM.ColToTensor("x", "a"), DM.ColToTensor("x", "b")])


def test_image_folder_dataset_and_mask_paste(tmp_path):
    import cv2

    from bigdl_b200.dataset.image import image_folder_dataset
    from bigdl_b200.models.maskrcnn import decode_mask_in_image

    for cls in ("cat", "dog"):
        (tmp_path / cls).mkdir()
        for i in range(2):
            cv2.imwrite(str(tmp_path / cls / f"{i}.png"), np.full((20, 30, 3), 50 + i, np.uint8))
    ds = image_folder_dataset(str(tmp_path), scale_to=10)
    items = list(ds.data(False))
    assert len(items) == 4 and sorted({i.label() for i in items}) == [1.0, 2.0]
    assert min(items[0].height(), items[0].width()) == 10

    mask = torch.ones(4, 4)
    pasted = decode_mask_in_image(mask, torch.tensor([2.0, 3.0, 9.0, 8.0]), 12, 14)
    assert pasted.shape == (12, 14) and pasted.dtype == np.uint8
    ys, xs = np.nonzero(pasted)
    assert 2 <= xs.min() <= 3 and 8 <= xs.max() <= 9 and 3 <= ys.min() <= 4 and 7 <= ys.max() <= 8
    assert decode_mask_in_image(mask, torch.tensor([-5.0, -5.0, 2.0, 2.0]), 6, 6).sum() > 0       # clipped to the image


def test_variable_format_fans_and_3d_ordering():
    from bigdl_b200.interop.keras_helpers import to_bigdl_3d_ordering
    from bigdl_b200.nn.initialization_method import VariableFormat as VF

    assert VF.fans([5, 3], VF.OUT_IN) == (3, 5) and VF.fans([5, 3], VF.IN_OUT) == (5, 3)
    assert VF.fans([8, 4, 3, 3], VF.OUT_IN_KW_KH) == (4 * 9, 8 * 9) and VF.fans([8, 4, 3, 3], VF.DEFAULT) == (36, 72)
    assert VF.fans([2, 8, 4, 3, 3], VF.GP_OUT_IN_KW_KH) == (4 * 9, 8 * 9)
    assert VF.fans([7], VF.ONE_D) == (7, 1)
    assert to_bigdl_3d_ordering("tf") == "channel_last" or to_bigdl_3d_ordering("tf").upper() in ("CHANNEL_LAST", "NDHWC")
    with pytest.raises(Exception):
        to_bigdl_3d_ordering("xx")
