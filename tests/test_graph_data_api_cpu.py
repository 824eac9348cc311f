"""Graph / Recurrent / DataSet builders that no other test exercises: stopGradient, graph topology export, hidden-state
get/set on Recurrent, ImageFolder / SeqFileFolder / imageFrame dataset builders, Node edge removal."""
import os

import numpy as np
import torch

from bigdl_b200.nn import layer as L
from bigdl_b200.nn.module import Graph, Input
from bigdl_b200.utils.random_generator import RNG


def test_stop_gradient_blocks_upstream_parameter_gradients(tmp_path):
    RNG.setSeed(2)
    inp = Input()
    a = L.Linear(4, 5).setName("a").inputs(inp)
    b = L.Tanh().setName("act").inputs(a)
    c = L.Linear(5, 2).setName("c").inputs(b)
    g = Graph(inp, c)
    x = torch.randn(3, 4)
    g.forward(x)
    g.backward(x, torch.ones(3, 2))
    assert float(g.find("a").gradWeight.abs().sum()) > 0
    g.zeroGradParameters()
    g.stopGradient(["act"])
    g.forward(x)
    g.backward(x, torch.ones(3, 2))
    assert float(g.find("a").gradWeight.abs().sum()) == 0.0          # nothing flows back through "act"
    assert float(g.find("c").gradWeight.abs().sum()) > 0
    g.stop_gradient(["act"])
    log = str(tmp_path / "topo")
    g.saveGraphTopology(log)
    g.save_graph_topology(log)
    assert any(f.startswith("events") or "tfevents" in f for _, _, fs in os.walk(log) for f in fs)
    node = c
    assert len(node.prev) == 1
    node.removePrevEdges()
    assert len(node.prev) == 0 and all(node not in p.next for p in [b])


def test_recurrent_hidden_state_roundtrip():
    RNG.setSeed(3)
    rec = L.Recurrent().add(L.LSTM(4, 5))
    x = torch.randn(2, 6, 4)
    full = rec.forward(x)
    first = rec.forward(x[:, :3])
    state = rec.getHiddenState()
    assert rec.get_hidden_state() is not None
    rec.setHiddenState(state)
    second = rec.forward(x[:, 3:])
    assert torch.allclose(torch.cat([first, second], dim=1), full, atol=1e-5)     # carried state continues the sequence
    rec.set_hidden_state(state)


def test_dataset_builders_from_folders_and_frames(tmp_path):
    import cv2

    from bigdl_b200.dataset.base import DataSet
    from bigdl_b200.transform.vision.image import ImageFrame

    for cls in ("cat", "dog"):
        os.makedirs(tmp_path / "imgs" / cls)
        for i in range(3):
            cv2.imwrite(str(tmp_path / "imgs" / cls / f"{i}.png"),
                        np.random.RandomState(i).randint(0, 255, (20, 24, 3)).astype(np.uint8))
    ds = DataSet.ImageFolder(str(tmp_path / "imgs"), scale_to=16)
    items = list(ds.data(train=False))
    assert ds.size() == 6 and len(items) == 6
    labels = sorted({float(getattr(it, "label", it[1] if isinstance(it, tuple) else 0) if not hasattr(it, "label")
                           else it.label() if callable(it.label) else it.label) for it in items})
    assert labels == [1.0, 2.0]
    frame = ImageFrame.array([np.zeros((8, 8, 3), np.uint8)] * 4, labels=[np.array([1.0])] * 4)
    assert DataSet.imageFrame(frame).size() == 4
