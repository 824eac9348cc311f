"""Optimizer configuration surface (S/optim/Optimizer.scala:47-474 and P/optim/optimizer.py BaseOptimizer): gradient
clipping, per-submodule optim methods, end trigger / state / model / criterion / data setters, checkpoint path helpers
and summaries — the knobs DistriOptimizerSpec exercises (:494-586, :658-716, :782-957)."""
import os

import torch

from bigdl_b200.dataset.base import FixedMiniBatchDataSet, MiniBatch
from bigdl_b200.nn.criterion import ClassNLLCriterion, MSECriterion
from bigdl_b200.nn.layer import Linear, LogSoftMax, Sequential
from bigdl_b200.optim.optim_method import SGD
from bigdl_b200.optim.optimizer import LocalOptimizer, Optimizer
from bigdl_b200.optim.trigger import MaxIteration, SeveralIteration
from bigdl_b200.utils.random_generator import RNG
from bigdl_b200.utils.table import T
from bigdl_b200.visualization.summary import TrainSummary, ValidationSummary


def _problem(scale=1.0, seed=0):
    RNG.setSeed(31)
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(16, 4, generator=g) * scale
    y = torch.randn(16, 2, generator=g) * scale
    model = Sequential().add(Linear(4, 3).setName("a")).add(Linear(3, 2).setName("b"))
    ds = FixedMiniBatchDataSet(MiniBatch(x, y), 16, distributed=False)
    return model, ds


def _one_step_delta(configure, scale=100.0):
    model, ds = _problem(scale)
    w, _ = model.getParameters()
    before = w.clone()
    opt = LocalOptimizer(model, ds, MSECriterion(), MaxIteration(1), 16, SGD(1.0))
    configure(opt)
    opt.optimize()
    return (model.getParameters()[0] - before)


def test_constant_and_l2_gradient_clipping():
    free = _one_step_delta(lambda o: None)
    assert float(free.abs().max()) > 1.0                         # large inputs: unclipped gradients are huge
    const = _one_step_delta(lambda o: o.setConstantGradientClipping(-0.01, 0.01))
    assert float(const.abs().max()) <= 0.01 + 1e-6               # lr 1: |delta| = |clipped grad|
    l2 = _one_step_delta(lambda o: o.set_gradclip_l2norm(0.5))
    assert abs(float(l2.norm()) - 0.5) < 1e-3
    off = _one_step_delta(lambda o: (o.set_gradclip_const(-0.01, 0.01), o.disableGradientClipping()))
    assert torch.allclose(off, free)
    off2 = _one_step_delta(lambda o: (o.setGradientClippingByl2Norm(0.5), o.disable_gradclip()))
    assert torch.allclose(off2, free)


def test_per_submodule_optim_methods():
    model, ds = _problem()
    a0, b0 = model.modules[0].weight.clone(), model.modules[1].weight.clone()
    opt = LocalOptimizer(model, ds, MSECriterion(), MaxIteration(3), 16, SGD(0.1))
    opt.setOptimMethods({"a": SGD(0.0), "b": SGD(0.1)})
    opt.optimize()
    assert torch.equal(model.modules[0].weight, a0)              # learning rate 0 on sub-module "a"
    assert not torch.allclose(model.modules[1].weight, b0)
    opt2 = LocalOptimizer(*_problem()[:1], ds, MSECriterion(), MaxIteration(1), 16, SGD(0.1))
    opt2.set_optim_methods({"a": SGD(0.1), "b": SGD(0.1)})
    opt2.set_optim_method(SGD(0.2))
    assert opt2.optimize() is not None


def test_setters_end_trigger_state_model_criterion_data(tmp_path):
    model, ds = _problem()
    opt = Optimizer(model, ds, MSECriterion(), MaxIteration(50), 16, SGD(0.05))
    opt.setEndWhen(MaxIteration(4))
    opt.optimize()
    assert opt.state["neval"] == 5
    opt.set_end_when(MaxIteration(6))
    opt.optimize()                                               # continues from the optimizer's own state
    assert opt.state["neval"] == 7
    opt.setState(T(epoch=1, neval=1))
    opt.set_end_when(MaxIteration(2))
    opt.optimize()
    assert opt.state["neval"] == 3
    other, other_ds = _problem(seed=5)
    opt.setModel(other).set_model(other)
    opt.setCriterion(MSECriterion()).set_criterion(MSECriterion())
    opt.setTrainData(other_ds, 16)
    opt.set_traindata(other_ds, 16)
    opt.setState(T(epoch=1, neval=1))
    assert opt.optimize() is other
    opt.prepareInput()
    opt.prepare_input()
    opt.setDropModuleProperty(0.1, 0.4)


def test_checkpoint_helpers_and_summaries(tmp_path):
    model, ds = _problem()
    opt = Optimizer(model, ds, MSECriterion(), MaxIteration(6), 16, SGD(0.05))
    ck = str(tmp_path / "ck")
    os.makedirs(ck)
    opt.set_checkpoint(SeveralIteration(2), ck)
    opt.overWriteCheckpoint()
    assert opt.getCheckpointPath() == ck
    ts = TrainSummary(str(tmp_path / "log"), "app")
    ts.set_summary_trigger("Parameters", SeveralIteration(2))
    assert ts.getSummaryTrigger("Parameters") is not None
    vs = ValidationSummary(str(tmp_path / "log"), "app")
    opt.set_train_summary(ts)
    opt.set_val_summary(vs)
    opt.optimize()
    losses = ts.read_scalar("Loss")
    assert len(losses) == 6 and [r[0] for r in losses] == list(range(1, 7))
    assert len(ts.read_scalar("Throughput")) == 6 and len(ts.read_scalar("LearningRate")) == 6
    saved = []
    for root, _, files in os.walk(ck):
        saved += files
    # single-process runs write model.N + state.N (LocalOptimizer.scala:216-227), distributed ones optimMethod-<name>.N
    assert any(f.startswith("model.") for f in saved) and any(f.startswith(("state.", "optimMethod")) for f in saved)
    ts.add_scalar("custom", 1.5, 7)
    ts.add_histogram("hist", torch.randn(100), 7)
    ts.close() if hasattr(ts, "close") else None
    vs.close() if hasattr(vs, "close") else None


def test_checkpointing_does_not_disturb_training(tmp_path):
    """Saving a snapshot clones the model; the live model's flat parameter buffer (which the optimizer updates) must
    stay in place — trajectories with and without checkpoints are identical."""
    def run(ck):
        model, ds = _problem()
        opt = LocalOptimizer(model, ds, MSECriterion(), MaxIteration(12), 16, SGD(0.1, momentum=0.9))
        if ck:
            opt.setCheckpoint(str(tmp_path / "c"), SeveralIteration(2))
        w_before = model.getParameters()[0]
        opt.optimize()
        assert model.getParameters()[0].data_ptr() == w_before.data_ptr()
        return model.getParameters()[0].clone(), float(opt.state["Loss"])

    (wa, la), (wb, lb) = run(False), run(True)
    assert torch.equal(wa, wb) and la == lb
    # and a clone never aliases or re-homes its source
    m, _ = _problem()
    w, _ = m.getParameters()
    c = m.cloneModule()
    c.modules[0].weight.data.add_(1.0)
    assert m.getParameters()[0].data_ptr() == w.data_ptr()
    assert not torch.equal(c.modules[0].weight, m.modules[0].weight)


def test_epoch_boundaries_schedule_survives_a_checkpoint(tmp_path):
    """EpochDecayWithWarmUp with the picklable EpochBoundaries decay: values, save/load, and the error for a local
    function (which pickle cannot write) names the fix and leaves no partial file behind."""
    import pytest
    from bigdl_b200.optim.optim_method import SGD, EpochBoundaries, EpochDecayWithWarmUp, OptimMethod

    d = EpochBoundaries(60, 30, 80)
    assert [d(e) for e in (1, 29, 30, 59, 60, 80, 200)] == [0, 0, 1, 1, 2, 3, 3]
    om = SGD(0.1, learningrate_schedule=EpochDecayWithWarmUp(0, 0.0, d))
    p = str(tmp_path / "om")
    om.save(p)
    back = OptimMethod.load(p)
    assert back.learningRateSchedule.decay_type.boundaries == (30, 60, 80)

    bad = SGD(0.1, learningrate_schedule=EpochDecayWithWarmUp(0, 0.0, lambda e: 0))
    q = str(tmp_path / "bad")
    with pytest.raises(TypeError, match="EpochBoundaries"):
        bad.save(q)
    assert not (tmp_path / "bad").exists() and not (tmp_path / "bad.tmp").exists()
