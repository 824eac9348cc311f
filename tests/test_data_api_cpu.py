"""Small public methods of the data / optim-method surface that nothing else calls (S/dataset/{DataSet,Sample,MiniBatch},
S/transform/vision/image/{ImageFeature,ImageFrame}, S/optim/OptimMethod)."""
import numpy as np
import torch

from bigdl_b200.dataset.base import DataSet, MiniBatch, Sample
from bigdl_b200.optim.optim_method import SGD, Adam
from bigdl_b200.transform.vision.image import ImageFeature, ImageFrame
from bigdl_b200.utils.table import T


def test_sample_minibatch_and_dataset_helpers():
    s = Sample([torch.randn(3), torch.randn(2)], torch.tensor([1.0]))
    assert s.numFeature() == 2 and s.numLabel() == 1
    samples = [Sample(torch.full((3,), float(i)), torch.tensor([float(i % 2 + 1)])) for i in range(6)]
    mb = MiniBatch.from_samples(samples)
    assert mb.size() == 6 and mb.get_input().shape == (6, 3) and mb.get_target().shape[0] == 6
    assert torch.equal(mb.slice(2, 3).get_input()[:, 0], torch.tensor([1.0, 2.0, 3.0]))       # 1-based offset
    ds = DataSet.array(samples)
    assert not ds.is_distributed() and ds.size() == 6
    dd = ds.toDistributed() if hasattr(ds, "toDistributed") else ds
    assert dd.size() == 6
    assert dd.toLocal().size() == 6 if hasattr(dd, "toLocal") else True


def test_image_feature_and_frame_getters():
    img = (np.arange(5 * 7 * 3) % 255).astype(np.uint8).reshape(5, 7, 3)
    f = ImageFeature(img, label=np.array([2.0]), path="mem://0")
    assert (f.getHeight(), f.getWidth()) == (5, 7) and (f.getOriginalHeight(), f.getOriginalWidth()) == (5, 7)
    assert f.opencvMat().shape == (5, 7, 3) and f.get_label() is not None
    frame = ImageFrame.array([img, img], labels=[np.array([1.0]), np.array([2.0])])
    assert frame.isLocal() and frame.is_local() and not frame.isDistributed() and not frame.is_distributed()
    imgs = frame.get_image(to_chw=True)
    assert len(imgs) == 2 and tuple(np.asarray(imgs[0]).shape) == (3, 5, 7)
    assert [float(np.asarray(l).reshape(-1)[0]) for l in frame.get_label()] == [1.0, 2.0]
    assert len(frame.get_uri()) == 2
    assert frame.get_predict() is not None and frame.get_sample() is not None


def test_optim_method_hyper_parameters_and_history():
    om = SGD(0.1, momentum=0.9, dampening=0.0)
    w, g = torch.ones(8), torch.ones(8)
    om.update(w, g)
    om.update(w, g)
    assert "0.1" in om.getHyperParameter() or "rate" in om.getHyperParameter().lower()
    assert abs(om.get_learning_rate() - 0.1) < 1e-9 or abs(om.get_learning_rate() + 0.1) < 1e-9
    om.clearHistory()
    w2 = torch.ones(8)
    om.clear_history()
    om.update(w2, g)                                  # momentum buffer starts again from zero
    assert torch.allclose(w2, torch.full((8,), 0.9))
    ad = Adam(1e-2)
    ad.loadFromTable(T(learningRate=0.5, beta1=0.8))
    assert ad.learningRate == 0.5 and ad.beta1 == 0.8
