"""DistriOptimizer on gloo ranks vs a straight-line oracle with the reference's truncation points
(port of T/optim/RefDistriOptimizer.scala:32-110; asserted bit-exact like T/optim/DistriOptimizerSpec.scala:378-448)."""
import numpy as np
import pytest
import torch

from dist_utils import run_ranks


def _model():
    from bigdl_b200.nn.layer import Linear, Sequential, Sigmoid, Tanh
    from bigdl_b200.utils.random_generator import RNG

    RNG.setSeed(10)
    return Sequential().add(Linear(4, 8)).add(Tanh()).add(Linear(8, 2)).add(Sigmoid())


def _batches(world):
    rs = np.random.RandomState(5)
    xs = rs.randn(world, 16, 4).astype(np.float32)
    ys = (rs.rand(world, 16, 2) > 0.5).astype(np.float32)
    return xs, ys


def _trunc(t):
    return ((t.contiguous().view(torch.int32) >> 16) << 16).view(torch.float32)


def _worker(rank, world, iters, compress, optim_name):
    from bigdl_b200.dataset.base import FixedMiniBatchDataSet, MiniBatch
    from bigdl_b200.nn.criterion import MSECriterion
    from bigdl_b200.optim.optim_method import SGD, Adam
    from bigdl_b200.optim.optimizer import DistriOptimizer, Optimizer
    from bigdl_b200.optim.trigger import MaxIteration
    from bigdl_b200.utils import config

    config.set("bigdl.parameter.compress", compress)
    xs, ys = _batches(world)
    model = _model()
    ds = FixedMiniBatchDataSet(MiniBatch(torch.from_numpy(xs[rank]), torch.from_numpy(ys[rank])), epoch_size=10 ** 6)
    om = SGD(0.5, momentum=0.9, dampening=0.0) if optim_name == "sgd" else Adam(0.05)
    opt = Optimizer(model, ds, MSECriterion(), MaxIteration(iters), 16 * world, om)
    assert isinstance(opt, DistriOptimizer)
    opt.optimize()
    w, _ = model.getParameters()
    return w.clone().numpy(), float(opt.state["Loss"])


def _oracle(world, iters, compress, optim_name):
    from bigdl_b200.nn.criterion import MSECriterion
    from bigdl_b200.optim.optim_method import SGD, Adam

    xs, ys = _batches(world)
    model = _model()
    w, g = model.getParameters()
    master = w.clone()
    crit = MSECriterion()
    om = SGD(0.5, momentum=0.9, dampening=0.0) if optim_name == "sgd" else Adam(0.05)
    rnd = _trunc if compress == "bf16_trunc" else (lambda t: t.to(torch.bfloat16).float())
    for _ in range(iters):
        with torch.no_grad():
            w.copy_(rnd(master))
        grads = []
        for r in range(world):
            model.zeroGradParameters()
            x, y = torch.from_numpy(xs[r]), torch.from_numpy(ys[r])
            out = model.forward(x)
            crit.forward(out, y)
            model.backward(x, crit.backward(out, y))
            grads.append(rnd(g.clone()))
        acc = grads[0]
        for r in range(1, world):
            acc = rnd(acc + grads[r]) if compress == "bf16_trunc" else acc + grads[r]
        om.update(master, acc / world)
    return master.numpy()


@pytest.mark.parametrize("world", [2, 4])
def test_distri_matches_reference_oracle_bit_exact(world):
    res = run_ranks(_worker, world, 6, "bf16_trunc", "sgd")
    ref = _oracle(world, 6, "bf16_trunc", "sgd")
    for w, _ in res:
        assert np.array_equal(w, ref), np.abs(w - ref).max()


def test_distri_rn_mode_close_to_fp32():
    res = run_ranks(_worker, 2, 8, "bf16_rn", "adam")
    ref = _oracle(2, 8, "bf16_rn", "adam")
    assert np.allclose(res[0][0], ref, atol=2e-3)
    assert np.array_equal(res[0][0], res[1][0])


def test_partition_rule():
    from bigdl_b200.nn.flat import partition_range
    from bigdl_b200.parallel.all_reduce_parameter import aligned_partition

    # size/N, first size%N slices get +1 (S/parameters/AllReduceParameter.scala:103-104,138-142)
    assert [partition_range(10, 4, r) for r in range(4)] == [(0, 3), (3, 3), (6, 2), (8, 2)]
    parts = [aligned_partition(22280, 8, r) for r in range(8)]
    assert parts[0][0] == 0 and sum(l for _, l in parts) == 22280
    assert all(s % 8 == 0 and l % 8 == 0 for s, l in parts)
    assert all(parts[i][0] + parts[i][1] == parts[i + 1][0] for i in range(7))


def _drop_worker(rank, world):
    """Rank 1 is made slow after the warm-up; the drop policy must leave its gradient out (and keep the ranks' weights
    identical), and a policy that tolerates no drops must discard those iterations altogether."""
    import time as _time

    import torch
    from bigdl_b200.dataset.base import FixedMiniBatchDataSet, MiniBatch
    from bigdl_b200.nn.criterion import MSECriterion
    from bigdl_b200.nn.layer import Linear, Sequential
    from bigdl_b200.nn.module import Module
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import MaxIteration
    from bigdl_b200.utils import config
    from bigdl_b200.utils.random_generator import RNG

    class Slow(Module):
        calls = 0

        def updateOutput(self, input):
            Slow.calls += 1
            if rank == 1 and Slow.calls > 6:
                _time.sleep(0.15)
            return input

    config.set("bigdl.parameter.compress", "fp32")
    out = {}
    for name, (drop, max_drop) in {"drop": (0.4, 0.5), "discard": (0.4, 0.4)}.items():
        Slow.calls = 0
        RNG.setSeed(3)
        m = Sequential().add(Slow()).add(Linear(8, 4))
        g = torch.Generator().manual_seed(10 + rank)
        x, y = torch.randn(16, 8, generator=g), torch.randn(16, 4, generator=g)
        opt = Optimizer(m, FixedMiniBatchDataSet(MiniBatch(x, y), 10 ** 6, distributed=True), MSECriterion(),
                        MaxIteration(12), 16 * world, SGD(0.05))
        opt.setDropModuleProperty(drop, max_drop, batchsize=3, warmupIteration=2)
        w0 = m.getParameters()[0].clone()
        opt.optimize()
        out[name] = (m.getParameters()[0].clone(), opt._drop_policy.dropped_total, w0)
    return out


def test_straggler_drop_policy_is_applied():
    res = run_ranks(_drop_worker, 2)
    for name in ("drop", "discard"):
        w0, w1 = res[0][name][0], res[1][name][0]
        assert torch.equal(w0, w1), name                              # every rank applied the same decisions
        assert res[0][name][1] == res[1][name][1] > 0, (name, res[0][name][1])
    # a discarded iteration changes nothing, so "discard" moved the weights less than "drop" did
    moved = {n: float((res[0][n][0] - res[0][n][2]).norm()) for n in ("drop", "discard")}
    assert 0 < moved["discard"] < moved["drop"], moved
