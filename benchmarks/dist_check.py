"""Multi-GPU correctness check for the fused NVLink parameter sync: trains a small conv net for a few iterations with
the p2p backend and with the NCCL baseline backend from the same seed and compares the resulting weights across ranks
and across backends. Launch with torchrun (one process per GPU)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist


def train(backend, iters, rank, world, dev):
    from bigdl_b200.dataset.base import FixedMiniBatchDataSet, MiniBatch
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.nn.layer import (Linear, LogSoftMax, ReLU, Sequential, SpatialConvolution, SpatialMaxPooling, View)
    from bigdl_b200.optim.optim_method import SGD
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import MaxIteration
    from bigdl_b200.utils import config
    from bigdl_b200.utils.random_generator import RNG

    config.set("bigdl.parameter.backend", {"parallel": "nccl", "parallel_p2p": "p2p"}.get(backend, backend))
    RNG.setSeed(7)
    m = Sequential()
    m.add(SpatialConvolution(8, 32, 3, 3, 1, 1, 1, 1)).add(ReLU()).add(SpatialMaxPooling(2, 2, 2, 2))
    m.add(SpatialConvolution(32, 64, 3, 3, 1, 1, 1, 1)).add(ReLU()).add(SpatialMaxPooling(2, 2, 2, 2))
    m.add(View(64 * 8 * 8).setNumInputDims(3)).add(Linear(64 * 8 * 8, 16)).add(LogSoftMax())
    m.cuda()
    g = torch.Generator().manual_seed(100 + rank)
    x = torch.randn(32, 8, 32, 32, generator=g).pin_memory()
    y = torch.randint(1, 17, (32,), generator=g).float().pin_memory()
    ds = FixedMiniBatchDataSet(MiniBatch(x, y), epoch_size=10 ** 9, distributed=True)
    om = SGD(0.05, momentum=0.9, dampening=0.0)
    if backend.startswith("parallel"):
        from bigdl_b200.optim.parallel_optimizer import ParallelOptimizer

        opt = ParallelOptimizer(m, ds, ClassNLLCriterion(), MaxIteration(iters), 32 * world, om, bucket_mb=0.05)
    else:
        opt = Optimizer(m, ds, ClassNLLCriterion(), MaxIteration(iters), 32 * world, om)
    opt.optimize()
    torch.cuda.synchronize()
    w = m._flat.compute.float().clone()
    return w, float(opt.state["Loss"])


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    from bigdl_b200.utils.engine import Engine

    Engine.init()
    w_p2p, l_p2p = train("p2p", 8, rank, world, dev)
    w_nccl, l_nccl = train("nccl", 8, rank, world, dev)
    w_par, l_par = train("parallel", 8, rank, world, dev)       # ParallelOptimizer: bucketed async all-reduce
    from bigdl_b200.parallel import distri_parameter_synchronizer as dps

    before = dps.P2P_ALLREDUCE_CALLS
    w_pp, l_pp = train("parallel_p2p", 8, rank, world, dev)     # same, buckets reduced by the one-shot NVLink kernel
    p2p_calls = dps.P2P_ALLREDUCE_CALLS - before
    diff_par = max(float((w_par - w_nccl).abs().max()), float((w_pp - w_nccl).abs().max()))
    # all ranks must hold identical compute weights
    ref = w_p2p.clone()
    dist.broadcast(ref, 0)
    same = bool(torch.equal(ref, w_p2p))
    flags = torch.tensor([int(same)], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    diff = float((w_p2p - w_nccl).abs().max())
    scale = float(w_nccl.abs().max())
    if rank == 0:
        print(json.dumps({"world": world, "ranks_identical": bool(flags.item()), "p2p_vs_nccl_max_abs_diff": diff,
                          "parallel_vs_nccl_max_abs_diff": diff_par, "weight_scale": scale, "loss_p2p": l_p2p,
                          "loss_nccl": l_nccl, "loss_parallel": l_par, "loss_parallel_p2p": l_pp,
                          "p2p_bucket_allreduces": p2p_calls}))
    ok = bool(flags.item()) and diff <= 2e-2 * scale + 1e-3 and diff_par <= 2e-2 * scale + 1e-3 and p2p_calls > 0
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
