"""The model-zoo Train / Test mains on tiny on-disk datasets in the reference's file formats
(S/models/{lenet,autoencoder,vgg,resnet,rnn}/{Train,Test}.scala)."""
import os
import struct

import numpy as np
import pytest
import torch

from bigdl_b200.models import datasets as D
from bigdl_b200.models import test as zoo_test
from bigdl_b200.models import train as zoo_train


def _write_mnist(folder, n_train=96, n_test=48):
    protos = np.random.RandomState(3).randint(0, 256, (10, 28, 28))
    for kind, n, seed in (("train", n_train, 0), ("t10k", n_test, 1)):
        rs = np.random.RandomState(seed)
        labels = rs.randint(0, 10, n).astype(np.uint8)
        imgs = np.clip(protos[labels] + rs.randint(-20, 20, (n, 28, 28)), 0, 255).astype(np.uint8)
        with open(os.path.join(folder, f"{kind}-images-idx3-ubyte"), "wb") as f:
            f.write(struct.pack(">iiii", 2051, n, 28, 28) + imgs.tobytes())
        with open(os.path.join(folder, f"{kind}-labels-idx1-ubyte"), "wb") as f:
            f.write(struct.pack(">ii", 2049, n) + labels.tobytes())


def _write_cifar(folder, per_file=8, n_test=8):
    rs = np.random.RandomState(0)
    protos = rs.randint(0, 256, (10, 3072))
    names = [f"data_batch_{i}.bin" for i in range(1, 6)] + ["test_batch.bin"]
    for name in names:
        n = n_test if name.startswith("test") else per_file
        labels = rs.randint(0, 10, (n, 1))
        px = np.clip(protos[labels[:, 0]] + rs.randint(-10, 10, (n, 3072)), 0, 255)
        np.concatenate([labels, px], 1).astype(np.uint8).tofile(os.path.join(folder, name))


def test_mnist_and_cifar_loaders(tmp_path):
    d = str(tmp_path)
    _write_mnist(d, 10, 4)
    recs = D.load_mnist(os.path.join(d, "train-images-idx3-ubyte"), os.path.join(d, "train-labels-idx1-ubyte"))
    assert len(recs) == 10 and len(recs[0].data) == 784 and 1 <= recs[0].label <= 10
    b = next(iter(D.mnist_batches(d, "test", 4).data(False)))
    assert tuple(b.getInput().shape) == (4, 28, 28)
    x = b.getInput()
    ab = next(iter(D.toAutoencoderBatch()(iter([b]))))
    assert ab.getTarget() is ab.getInput()
    with pytest.raises(ValueError):
        D.load_mnist(os.path.join(d, "train-labels-idx1-ubyte"), os.path.join(d, "train-labels-idx1-ubyte"))

    _write_cifar(d, 2, 2)
    rec = D.load_cifar10_test(d)[0]
    raw = np.fromfile(os.path.join(d, "test_batch.bin"), dtype=np.uint8)[:3073]
    assert struct.unpack_from(">ii", rec.data, 0) == (32, 32) and rec.label == raw[0] + 1.0
    px = np.frombuffer(rec.data, np.uint8, offset=8).reshape(32, 32, 3)
    planes = raw[1:].reshape(3, 32, 32)                    # R, G, B planes in the file
    assert np.array_equal(px[..., 2], planes[0]) and np.array_equal(px[..., 0], planes[2])
    assert len(D.load_cifar10_train(d)) == 10


def test_lenet_train_then_test(tmp_path, capsys):
    d, ck = str(tmp_path / "mnist"), str(tmp_path / "ck")
    os.makedirs(d)
    _write_mnist(d, 256, 48)
    opt = zoo_train.main(["lenet", "-f", d, "-b", "16", "-e", "4", "-r", "0.1", "--checkpoint", ck, "--overWrite"])
    assert opt == 0
    files = sorted(os.listdir(ck))
    assert any(f.startswith("model") for f in files) and any(f.startswith("optimMethod") or f.startswith("state")
                                                               for f in files), files
    model_file = os.path.join(ck, [f for f in files if f.startswith("model")][-1])
    res = zoo_test.lenet(["-f", d, "--model", model_file, "-b", "16"])
    acc, n = res[0][0].result()
    assert n == 48 and acc > 0.5
    assert "Top1Accuracy" in capsys.readouterr().out
    # resume from the snapshots, graph form flag ignored for a loaded model
    state_file = os.path.join(ck, [f for f in files if not f.startswith("model")][-1])
    o2 = zoo_train.lenet(["-f", d, "-b", "16", "-e", "5", "--model", model_file, "--state", state_file])
    assert o2.state["epoch"] >= 5


def test_autoencoder_vgg_resnet_mains(tmp_path):
    d = str(tmp_path / "mnist")
    os.makedirs(d)
    _write_mnist(d, 32, 8)
    o = zoo_train.autoencoder(["-f", d, "-b", "16", "--maxIteration", "3", "-g"])
    assert o.state["neval"] == 4 and np.isfinite(o.state["Loss"])

    c = str(tmp_path / "cifar")
    os.makedirs(c)
    _write_cifar(c, 4, 4)
    o = zoo_train.vgg(["-f", c, "-b", "4", "--maxIteration", "2", "--summary", str(tmp_path / "tb")])
    assert o.state["neval"] == 3 and os.listdir(str(tmp_path / "tb"))
    o = zoo_train.resnet(["-f", c, "-b", "4", "--maxIteration", "2", "--depth", "8", "--shortcutType", "A"])
    assert o.state["neval"] == 3
    path = str(tmp_path / "r8.bigdl")
    o.model.saveModule(path, overWrite=True)
    res = zoo_test._mains()["resnet"](["-f", c, "--model", path, "-b", "4"])
    assert res[0][0].result()[1] == 4
    assert zoo_train.cifar10Decay(1) == 0.0 and zoo_train.cifar10Decay(81) == 1.0 and zoo_train.cifar10Decay(122) == 2.0
    assert zoo_train.main(["nosuch"]) == 2 and zoo_test.main([]) == 0


def test_rnn_train_test_and_sequence_preprocess(tmp_path):
    d, s = str(tmp_path / "txt"), str(tmp_path / "save")
    os.makedirs(d)
    words = ["the", "cat", "sat", "on", "mat", "dog", "ran", "far"]
    rs = np.random.RandomState(0)
    for name, n in (("train.txt", 24), ("val.txt", 6), ("test.txt", 4)):
        with open(os.path.join(d, name), "w") as f:
            for _ in range(n):
                f.write(" ".join(rs.choice(words, rs.randint(3, 6))) + ".\n")
    toks = D.SequencePreprocess(os.path.join(d, "train.txt"))
    assert toks[0][0] == "SENTENCESTART" and toks[0][-1] == "SENTENCEEND" and len(toks) == 24
    o = zoo_train.rnn(["-f", d, "-s", s, "-b", "8", "--hidden", "8", "--vocab", "20", "-e", "2", "--checkpoint",
                       str(tmp_path / "ck")])
    assert os.path.exists(os.path.join(s, "dictionary.txt")) and np.isfinite(o.state["Loss"])
    mp = os.path.join(s, "rnn.bigdl")
    o.model.saveModule(mp, overWrite=True)
    for name in ("test.txt",):
        os.replace(os.path.join(d, name), os.path.join(s, name))
    res = zoo_test.rnn(["-f", s, "--model", mp, "-b", "2"])
    assert np.isfinite(res[0][0].result()[0])
    gen = zoo_test.rnn(["-f", s, "--model", mp, "--evaluate", "false", "--words", "3"])
    assert len(gen) == 4 and all(len(g) >= 4 for g in gen)

    # the PTB form and the windowed reader
    p = str(tmp_path / "ptb")
    os.makedirs(p)
    for k in ("train", "valid", "test"):
        with open(os.path.join(p, f"ptb.{k}.txt"), "w") as f:
            f.write("a b c a b\nb c a\n")
    tr, va, te, dic = D.SequencePreprocess.ptb(p, 4)
    assert len(tr) == 10 and dic.getVocabSize() == 3 and max(tr) <= 4 and min(tr) >= 1
    win = D.SequencePreprocess.reader(list(range(10)), 3)
    assert win == [[0, 1, 2, 3], [3, 4, 5, 6], [6, 7, 8, 9]]


def test_imagenet_pipelines_and_test_mains(tmp_path):
    from bigdl_b200.dataset.seqfile import write_image_seq_files
    from bigdl_b200.nn import layer as L

    d = str(tmp_path / "seq")
    os.makedirs(d)
    rs = np.random.RandomState(0)
    items = [(rs.randint(0, 256, (260 + 10 * i, 300, 3)).astype(np.uint8), float(i % 3 + 1), f"img{i}.jpg")
             for i in range(5)]
    list(write_image_seq_files(iter(items), os.path.join(d, "part"), 3, True))
    sets = {"i2012": D.ImageNet2012(d, None, 224, 2), "i2012val": D.ImageNet2012Val(d, None, 224, 2),
            "resnet-train": D.ImageNetDataSet.trainDataSet(d, 2, 224), "resnet-val": D.ImageNetDataSet.valDataSet(d, 2, 224),
            "vgg-train": D.vgg_imagenet_train(d, None, 224, 2), "vgg-val": D.vgg_imagenet_val(d, None, 224, 2),
            "frame": D.ImageNet2012.rdd(d, 2, None, 224), "frame-val": D.ImageNet2012Val.rdd(d, 2, None, 224)}
    for name, ds in sets.items():
        batches = list(ds.data(False))
        assert [tuple(b.getInput().shape) for b in batches] == [(2, 3, 224, 224)] * 2 + [(1, 3, 224, 224)], name
        assert batches[0].getTarget().reshape(-1).tolist() == [1.0, 2.0], name
    # the two centre-crop validation pipelines are deterministic
    a = next(iter(sets["i2012val"].data(False))).getInput()
    b = next(iter(D.ImageNet2012Val(d, None, 224, 2).data(False))).getInput()
    assert torch.equal(a, b)
    # ResNet pipeline: (x − mean) / 128 keeps values inside (−1, 1.2)
    x = next(iter(sets["resnet-val"].data(False))).getInput()
    assert float(x.min()) >= -1.0 and float(x.max()) <= 1.2

    tiny = L.Sequential().add(L.SpatialAveragePooling(224, 224)).add(L.View(3)).add(L.Linear(3, 1000)).add(L.LogSoftMax())
    mp = str(tmp_path / "tiny.bigdl")
    tiny.saveModule(mp, overWrite=True)
    r1 = zoo_test.resnet_imagenet(["-f", d, "--model", mp, "-b", "2"])
    r2 = zoo_test.inception(["-f", d, "--model", mp, "-b", "2"])
    assert [r[0].result()[1] for r in r1] == [5, 5] and [r[0].result()[1] for r in r2] == [5, 5]


def test_companion_object_factories():
    from bigdl_b200.models.autoencoder import Autoencoder
    from bigdl_b200.models.lenet import LeNet5
    from bigdl_b200.models.resnet import ResNet
    from bigdl_b200.models.vgg import VggForCifar10
    from bigdl_b200.nn.layer import Graph
    from bigdl_b200.utils.table import T

    x = torch.rand(2, 28, 28)
    seq, gr = LeNet5(10), LeNet5.graph(10)
    assert isinstance(gr, Graph) and tuple(gr.forward(x).shape) == (2, 10)
    for p, q in zip(seq.parameters()[0], gr.parameters()[0]):
        q.data.copy_(p.data)
    assert torch.allclose(seq.forward(x), gr.forward(x), atol=1e-6)
    k, kg = LeNet5.keras(10), LeNet5.kerasGraph(10)
    xk = x.reshape(2, 28, 28, 1)
    assert torch.allclose(k.forward(xk).sum(1), torch.ones(2), atol=1e-5) and tuple(kg.forward(xk).shape) == (2, 10)
    assert tuple(LeNet5.dnn(2, 10).forward(x).shape) == (2, 10) and isinstance(LeNet5.dnnGraph(2, 10), Graph)
    assert isinstance(VggForCifar10.graph(10), Graph) and isinstance(Autoencoder.graph(32), Graph)
    r = ResNet.graph(10, T(depth=8, shortcutType="A"))
    assert isinstance(r, Graph) and tuple(r.forward(torch.rand(2, 3, 32, 32)).shape) == (2, 10)
    m = ResNet.apply(10, {"depth": 8, "shortcutType": "A"})
    assert ResNet.shareGradInput(m) is m
    ResNet.modelInit(m)


def _lenet_rank(rank, world, folder):
    from bigdl_b200.models import train as zt

    opt = zt.lenet(["-f", folder, "-b", "32", "--maxIteration", "6", "-r", "0.1"])
    w, _ = opt.model.getParameters()
    return float(w.double().sum()), float(w.double().abs().sum()), int(opt.state["neval"]), type(opt).__name__


def test_lenet_main_data_parallel_world2(tmp_path):
    from dist_utils import run_ranks

    d = str(tmp_path / "mnist")
    os.makedirs(d)
    _write_mnist(d, 128, 16)
    out = run_ranks(_lenet_rank, 2, d)
    assert out[0][3] == "DistriOptimizer" and out[0][2] == 7
    assert out[0][:2] == pytest.approx(out[1][:2], rel=1e-9)          # both ranks hold the same weights after training
