"""Smoke + semantics of the AbstractModule surface that no other test touches (S/nn/abstractnn/AbstractModule.scala):
timers, names, parameter tables, init methods, scale getters, updateParameters, clearState, snake-case aliases, and
the save*/load* convenience wrappers."""
import torch

from bigdl_b200.nn import layer as L
from bigdl_b200.nn.initialization_method import ConstInitMethod, Zeros
from bigdl_b200.nn.module import Module
from bigdl_b200.utils.random_generator import RNG


def _net():
    RNG.setSeed(12)
    return L.Sequential().add(L.Linear(6, 5).setName("fc1")).add(L.ReLU()).add(L.Linear(5, 3).setName("fc2"))


def test_names_timers_and_training_flags():
    m = _net()
    assert m.modules[0].getName() == "fc1" == m.modules[0].get_name() and m.modules[0].hasName()
    assert not L.ReLU().hasName()
    assert m.apply_name("fc2") is m.modules[2] if hasattr(m, "apply_name") and m.apply_name("fc2") is not None else True
    x = torch.randn(4, 6)
    m.forward(x)
    m.backward(x, torch.randn(4, 3))
    times = m.getTimes()
    assert len(times) >= 3 and all(t[1] >= 0 and t[2] >= 0 for t in times)
    grouped = m.getTimesGroupByModuleType()
    assert {g[0] for g in grouped} >= {"Linear", "ReLU"}
    m.resetTimes()
    assert all(t[1] == 0 and t[2] == 0 for t in m.get_times())
    m.reset_times()
    assert m.isTraining() and m.is_training()
    m.evaluate()
    assert not m.isTraining() and not m.modules[0].isTraining()


def test_parameter_tables_scales_and_update():
    m = _net()
    tab = m.getParametersTable()
    assert set(tab.keys()) >= {"fc1", "fc2"} and set(tab["fc1"].keys()) >= {"weight", "bias", "gradWeight", "gradBias"}
    assert m.get_parameters_table().keys() == tab.keys()
    w, g = m.getParameters()
    w2, g2 = m.get_parameters()
    assert w.data_ptr() == w2.data_ptr() and w.numel() >= 6 * 5 + 5 + 5 * 3 + 3 and w.numel() % 8 == 0   # 8-element padding
    fc1 = m.modules[0]
    fc1.setScaleW(0.5).setScaleB(2.0)
    assert (fc1.getScaleW(), fc1.getScaleB()) == (0.5, 2.0)
    x = torch.randn(4, 6)
    m.zeroGradParameters()
    m.forward(x)
    m.backward(x, torch.ones(4, 3))
    before = w.clone()
    m.updateParameters(0.1)
    assert torch.allclose(w, before - 0.1 * g)
    m.update_parameters(0.0)
    assert fc1.own_parameters() if hasattr(fc1, "own_parameters") else True


def test_init_methods_freeze_and_clear_state():
    m = _net()
    fc2 = m.modules[2]
    fc2.setInitMethod(ConstInitMethod(0.25), Zeros())
    assert torch.all(fc2.weight == 0.25) and torch.all(fc2.bias == 0)
    fc2.set_init_method(ConstInitMethod(-0.5), ConstInitMethod(1.0))
    assert torch.all(fc2.weight == -0.5) and torch.all(fc2.bias == 1)
    m.freeze("fc1")
    x = torch.randn(4, 6)
    m.zeroGradParameters()
    m.forward(x)
    m.backward(x, torch.ones(4, 3))
    assert float(m.modules[0].gradWeight.abs().sum()) == 0.0
    m.unfreeze()
    m.forward(x)
    m.backward(x, torch.ones(4, 3))
    assert float(m.modules[0].gradWeight.abs().sum()) > 0.0
    m.clearState()
    m.clear_state()
    assert m.forward(x).shape == (4, 3)                # still usable after dropping cached activations


def test_update_output_grad_input_protocol_and_predict_class():
    m = _net()
    x = torch.randn(5, 6)
    y = m.forward(x)
    gi = m.updateGradInput(x, torch.ones_like(y))
    m.accGradParameters(x, torch.ones_like(y))
    assert gi.shape == x.shape
    cls = m.predict_class(x) if hasattr(m, "predict_class") else m.predictClass(x)
    assert len(cls) == 5 and all(1 <= c <= 3 for c in cls)
    g = m.to_graph()
    assert torch.allclose(g.forward(x), y, atol=1e-6)


def test_save_load_convenience_wrappers(tmp_path):
    m = _net()
    m.evaluate()
    x = torch.randn(3, 6)
    y = m.forward(x)
    p = str(tmp_path / "m.model")
    m.saveModel(p, over_write=True) if "over_write" in Module.saveModel.__code__.co_varnames else m.saveModel(p, None, True)
    assert torch.allclose(Module.loadModule(p).evaluate().forward(x), y, atol=1e-6)
    m.save_model(str(tmp_path / "m2.model"), None, True)
    t7 = str(tmp_path / "m.t7")
    m.saveTorch(t7, True)
    assert torch.allclose(Module.load_torch(t7).evaluate().forward(x), y, atol=1e-5)
    pb = str(tmp_path / "m.pb")
    m.save_tensorflow([("input", [3, 6])], pb)
    back_tf = Module.load_tensorflow(pb, ["input"], ["fc2"])
    assert torch.allclose(back_tf.evaluate().forward(x), y, atol=1e-5)
    proto, weights = str(tmp_path / "m.prototxt"), str(tmp_path / "m.caffemodel")
    m.save_caffe(proto, weights, True, True)
    back = Module.load_caffe_model(proto, weights)
    assert torch.allclose(back.evaluate().forward(x), y, atol=1e-5)
    again = Module.load_caffe(_net(), proto, weights, match_all=False)
    assert torch.allclose(again.evaluate().forward(x), y, atol=1e-5)


def test_bulk_tensor_storage_bytes_match_the_elementwise_encoding(monkeypatch):
    """Large float/double/bool storages are written as one raw packed record; the bytes on the wire and the values
    read back must be the same as with the element-by-element encoder."""
    import torch
    from bigdl_b200.serialization import module_io as mio

    torch.manual_seed(0)
    for t in (torch.randn(40, 20), torch.randn(300, dtype=torch.float64), torch.rand(17, 31) > 0.5,
              torch.randint(-5, 5, (400,)), torch.randn(50, 30).t()):
        fast = mio.tensor_to_proto(t, None)
        monkeypatch.setattr(mio, "_RAW_MIN", 1 << 60)
        slow = mio.tensor_to_proto(t, None)
        back_slow = mio.proto_to_tensor(fast, {})
        monkeypatch.setattr(mio, "_RAW_MIN", 256)
        assert fast.SerializeToString() == slow.SerializeToString()
        back = mio.proto_to_tensor(fast, {})
        assert back.dtype == back_slow.dtype and torch.equal(back, back_slow)
        assert torch.equal(back.to(t.dtype), t)
