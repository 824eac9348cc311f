"""Interop loaders against the REFERENCE's own test fixtures (read-only under /root/reference; skipped when the tree is
not there): Caffe prototxt + caffemodel (with a customized converter for the fixture's 'Dummy' layer, as the reference's
CaffeLoaderSpec does), Torch7 .t7 tensors, TensorFlow GraphDef / TFRecord files."""
import os

import numpy as np
import pytest
import torch

RES = "/root/reference/spark/dl/src/test/resources"
pytestmark = pytest.mark.skipif(not os.path.isdir(RES), reason="reference fixtures not available")


def test_caffe_fixture_with_customized_converter():
    from bigdl_b200.interop.caffe import CaffeLoader
    from bigdl_b200.nn.layer import Identity

    seen = []

    def dummy(layer, blobs):
        seen.append(layer.name)
        return Identity().setName(layer.name)

    with pytest.raises(ValueError, match="Dummy"):          # without the converter the custom layer is reported by type
        CaffeLoader.loadCaffe(f"{RES}/caffe/test.prototxt", f"{RES}/caffe/test.caffemodel")
    g, crit = CaffeLoader.loadCaffe(f"{RES}/caffe/test.prototxt", f"{RES}/caffe/test.caffemodel",
                                    customizedConverters={"Dummy": dummy})
    assert seen and crit is not None

    # the object form: a Customizable gets the parsed net as context (CaffeLoader.scala:37-47,73-83)
    from bigdl_b200.interop.caffe import CaffeConversionException, Customizable, LayerConverter

    class DummyConv(Customizable):
        def convertor(self, layer):
            return Identity().setName(layer.name)

    dc = DummyConv()
    g2, _ = CaffeLoader.loadCaffe(f"{RES}/caffe/test.prototxt", f"{RES}/caffe/test.caffemodel",
                                  customizedConverters={"Dummy": dc})
    assert "netparam" in dc.contexts and dc.contexts["name2LayerV2"] and len(list(g2.all_modules())) == len(list(g.all_modules()))
    with pytest.raises(CaffeConversionException, match="Caffe conversion error"):
        LayerConverter().convert("NoSuchType", type("L", (), {"name": "x"})(), [])
    assert LayerConverter().registerCutomizedConverter("Dummy", dummy).custom["Dummy"] is dummy
    g.evaluate()
    params = [p for p in g.torch_parameters()]
    assert len(params) >= 5 and all(float(p.abs().sum()) > 0 for p in params if p.dim() > 1)     # trained blobs were copied
    x = torch.randn(2, 3, 5, 5)
    y = g.forward(x)
    y = y if isinstance(y, torch.Tensor) else y[1]
    assert torch.isfinite(y).all()


def test_caffe_persist_fixture():
    from bigdl_b200.interop.caffe import CaffeLoader

    g, _ = CaffeLoader.loadCaffe(f"{RES}/caffe/test_persist.prototxt", f"{RES}/caffe/test_persist.caffemodel")
    assert len(list(g.all_modules())) > 3


@pytest.mark.parametrize("name", ["n02110063_11239.t7", "n03000134_4970.t7", "n04370456_5753.t7", "n15075141_38508.t7"])
def test_torch7_fixture_tensors(name):
    from bigdl_b200.interop.torchfile import load as load_t7

    t = load_t7(f"{RES}/torch/{name}")
    t = t if isinstance(t, torch.Tensor) else torch.as_tensor(np.asarray(t))
    assert tuple(t.shape) == (3, 224, 224) and torch.isfinite(t.float()).all()


def test_tensorflow_fixture_graph_runs():
    from bigdl_b200.interop.tensorflow import load_tf

    m = load_tf(f"{RES}/tf/test.pb", ["Placeholder"], ["output"])
    m.evaluate()
    y = m.forward(torch.randn(4, 1))
    assert torch.isfinite(y if isinstance(y, torch.Tensor) else y[1]).all()


def test_tfrecord_fixture_reads():
    from bigdl_b200.interop.tfrecord import TFRecordIterator

    recs = list(TFRecordIterator(f"{RES}/tf/mnist_train.tfrecord"))
    assert len(recs) > 0 and all(len(r) > 0 for r in recs[:5])
