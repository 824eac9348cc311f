"""Event-file writer/reader primitives, the numeric row transformer and the fixed-step sentence splitter
(S/visualization/tensorboard/{FileWriter,FileReader,Summary}.scala, S/dataset/datamining/RowTransformer.scala,
S/dataset/text/TextToLabeledSentence.scala)."""
import os

import numpy as np
import torch

from bigdl_b200.dataset import datamining as DM
from bigdl_b200.dataset import text as Tx
from bigdl_b200.visualization import summary as S


def test_file_writer_reader_roundtrip_with_histogram(tmp_path):
    log = str(tmp_path / "ev")
    w = S.FileWriter(log, flush_millis=50)
    for step in range(1, 6):
        w.addSummary(S.Summary.scalar("loss", 1.0 / step), step)
    w.addSummary(S.Summary.histogram("weights", torch.randn(256)), 5)
    w.flush()
    w.close()
    files = S.FileReader.list_files(log)
    assert len(files) == 1 and os.path.getsize(files[0]) > 0
    recs = S.FileReader.readScalar(log, "loss")
    assert [r[0] for r in recs] == [1, 2, 3, 4, 5] and abs(recs[2][1] - 1 / 3) < 1e-6
    assert S.FileReader.readScalar(log, "missing") == []
    h = S.histogram_proto(np.array([0.0, 1.0, 1.0, 5.0]))
    assert isinstance(h, (bytes, bytearray)) and len(h) > 0


def test_row_transformer_numeric_and_atomic():
    rows = [{"age": 30, "height": 1.8, "name": "a"}, {"age": 41, "height": 1.6, "name": "b"}]
    numeric = DM.RowTransformer.numeric("num", ["age", "height"])
    out = list(numeric(iter(rows)))
    assert torch.allclose(torch.as_tensor(out[0]["num"]).float(), torch.tensor([30.0, 1.8]))
    atomic = DM.RowTransformer.atomic(["age", "height"])
    out = list(atomic(iter(rows)))
    assert float(torch.as_tensor(out[1]["age"]).reshape(-1)[0]) == 41.0 and set(out[0].keys()) == {"age", "height"}


def test_text_to_sentence_with_steps():
    sents = [list(range(10, 16)), list(range(20, 30))]
    out = list(Tx.TextToSentenceWithSteps(5)(iter(sents)))
    assert out[0].data() == [10, 11, 12, 13, 14] and out[0].label() == [11, 12, 13, 14, 15]
    assert out[1].dataLength() == 5 and out[1].labelLength() == 5 and out[1].getLabel(0) == 21
    try:
        list(Tx.TextToSentenceWithSteps(8)(iter([list(range(5))])))
        raise AssertionError("a sentence shorter than numSteps + 1 must be rejected")
    except ValueError:
        pass
