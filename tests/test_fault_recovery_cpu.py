"""The retry-from-snapshot loop of DistriOptimizer (reference: DistriOptimizer.scala:881-963, exercised there with the
ExceptionTest layer): an injected failure mid-training is survived when a checkpoint exists, configuration errors and
exhausted budgets are not."""
import time

import pytest
import torch


def _setup(tmp_path, fail_at, iters=12):
    from bigdl_b200.dataset.base import FixedMiniBatchDataSet, MiniBatch
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.nn.layer import Linear, LogSoftMax, Sequential
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.optim.optimizer import DistriOptimizer
    from bigdl_b200.optim.trigger import MaxIteration, SeveralIteration
    from bigdl_b200.utils.fault import FaultInjector
    from bigdl_b200.utils.random_generator import RNG

    RNG.setSeed(3)
    FaultInjector.reset()
    m = Sequential().add(Linear(6, 8)).add(FaultInjector(fail_at)).add(Linear(8, 3)).add(LogSoftMax())
    g = torch.Generator().manual_seed(0)
    x = torch.randn(16, 6, generator=g)
    y = torch.randint(1, 4, (16,), generator=g).float()
    ds = FixedMiniBatchDataSet(MiniBatch(x, y), 16, distributed=False)
    opt = DistriOptimizer(m, ds, ClassNLLCriterion(), MaxIteration(iters), 16, SGD(0.1))
    if tmp_path is not None:
        opt.setCheckpoint(str(tmp_path), SeveralIteration(2))
    return opt, m


def test_injected_fault_is_survived_from_latest_snapshot(tmp_path):
    from bigdl_b200.utils.fault import FaultInjector

    opt, m = _setup(tmp_path, fail_at={6})
    trained = opt.optimize()
    assert opt.state["neval"] == 13                      # reached the end trigger despite the failure
    assert FaultInjector.calls >= 13                     # the failed step was re-run after the reload
    assert torch.isfinite(trained.getParameters()[0]).all()


def test_fault_without_checkpoint_propagates():
    from bigdl_b200.utils.fault import InjectedFault

    opt, _ = _setup(None, fail_at={3})
    with pytest.raises(InjectedFault):
        opt.optimize()


def test_retry_budget_is_bounded(tmp_path):
    from bigdl_b200.utils import config
    from bigdl_b200.utils.fault import InjectedFault

    config.set("bigdl.failure.retryTimes", 2)
    try:
        opt, _ = _setup(tmp_path, fail_at=set(range(4, 40)))       # fails on every step from the 4th on
        with pytest.raises(InjectedFault):
            opt.optimize()
    finally:
        config.unset("bigdl.failure.retryTimes")


def test_step_watchdog_expires_and_can_be_kicked():
    from bigdl_b200.utils.fault import StepWatchdog

    with StepWatchdog(0.2) as w:
        for _ in range(5):
            time.sleep(0.05)
            w.kick()
        w.check()
        time.sleep(0.5)
        with pytest.raises(TimeoutError):
            w.check()
