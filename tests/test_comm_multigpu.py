"""The cross-GPU protocol of the parameter plane on REAL peers: flag barriers over CUDA-IPC mappings, the per-bucket
reduce-scatter + update launched from inside backward, the pulled all-gather, graph replay of the generation counter,
device-side global-norm clipping and per-submodule OptimMethods. One process per GPU; skipped on boxes with one GPU.
Mirrors T/optim/DistriOptimizerSpec.scala:378-448 (same result with and without compression / across partitions)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _model():
    from bigdl_b200.nn.layer import (Linear, LogSoftMax, ReLU, Sequential, SpatialConvolution, SpatialMaxPooling, View)

    m = Sequential()
    m.add(SpatialConvolution(8, 32, 3, 3, 1, 1, 1, 1)).add(ReLU()).add(SpatialMaxPooling(2, 2, 2, 2))
    m.add(SpatialConvolution(32, 64, 3, 3, 1, 1, 1, 1)).add(ReLU()).add(SpatialMaxPooling(2, 2, 2, 2))
    m.add(SpatialConvolution(64, 64, 3, 3, 1, 1, 1, 1)).add(ReLU())
    head = Sequential().setName("head")
    # 4096 x 512: big enough (with bigdl.parameter.gatherGemmMB lowered below) for the all-gather-fused forward GEMM
    head.add(View(64 * 8 * 8).setNumInputDims(3)).add(Linear(64 * 8 * 8, 512)).add(ReLU()).add(Linear(512, 16))
    m.add(head).add(LogSoftMax())
    return m


def _train(rank, world, iters=6, **cfg):
    from bigdl_b200.dataset.base import FixedMiniBatchDataSet, MiniBatch
    from bigdl_b200.nn.criterion import ClassNLLCriterion
    from bigdl_b200.optim.optim_method import SGD, Adagrad
    from bigdl_b200.optim.optimizer import Optimizer
    from bigdl_b200.optim.trigger import MaxIteration
    from bigdl_b200.utils import config
    from bigdl_b200.utils.random_generator import RNG

    defaults = {"bigdl.parameter.backend": "p2p", "bigdl.parameter.p2p": "bucketed", "bigdl.parameter.overlap": True,
                "bigdl.cudaGraph": True, "bigdl.parameter.compress": "bf16_rn", "bigdl.parameter.bucketMB": 0.05,
                "bigdl.parameter.firstBucketMB": 0.01, "bigdl.parameter.gatherGemmMB": 0.5}
    clip = cfg.pop("clip", None)
    multi = cfg.pop("multi", False)
    defaults.update(cfg)
    for k, v in defaults.items():
        config.set(k, v)
    RNG.setSeed(7)
    m = _model().cuda()
    g = torch.Generator().manual_seed(100 + rank)
    x = torch.randn(32, 8, 32, 32, generator=g).pin_memory()
    y = torch.randint(1, 17, (32,), generator=g).float().pin_memory()
    ds = FixedMiniBatchDataSet(MiniBatch(x, y), epoch_size=10 ** 9, distributed=True)
    opt = Optimizer(m, ds, ClassNLLCriterion(), MaxIteration(iters), 32 * world, SGD(0.05, momentum=0.9, dampening=0.0))
    if multi:
        # the reference's wide&deep-style set-up: a different OptimMethod per sub-module
        body = [c for c in m.modules if c.getName() != "head"]
        for i, c in enumerate(body):
            c.setName(f"body{i}")
        methods = {c.getName(): SGD(0.05, momentum=0.9, dampening=0.0) for c in body if c.parameters()[0]}
        methods["head"] = Adagrad(0.002)
        opt.setOptimMethods(methods)
    if clip is not None:
        opt.setGradientClippingByl2Norm(clip)
    opt.optimize()
    torch.cuda.synchronize()
    sync = opt._param_sync
    stats = dict(sync._plane.stats) if getattr(sync, "_plane", None) is not None else {}
    stats["buckets"] = sync._plane.nb if getattr(sync, "_plane", None) is not None else 0
    return m._flat.compute.float().cpu(), m._flat.weights.float().cpu(), float(opt.state["Loss"]), stats


def _worker(rank, world):
    import torch.distributed as dist

    out = {}
    w_graph, m_graph, loss_g, st = _train(rank, world)
    out["buckets"], out["overlapped"], out["gathered_gemms"] = st["buckets"], st["overlapped"], st["gathered_gemms"]
    # the same with the weight rows pushed by their owners instead of gathered by the consumer GEMM
    w_nogather, _, _, _ = _train(rank, world, **{"bigdl.parameter.gatherGemmMB": 0.0})
    out["gather_vs_push"] = float((w_graph - w_nogather).abs().max())
    # every rank holds the same compute weights and the same gathered fp32 master weights
    ref = w_graph.cuda()
    dist.broadcast(ref, 0)
    out["ranks_identical"] = bool(torch.equal(ref.cpu(), w_graph))
    refm = m_graph.cuda()
    dist.broadcast(refm, 0)
    out["masters_identical"] = bool(torch.equal(refm.cpu(), m_graph))
    out["master_matches_compute"] = float((m_graph.bfloat16().float() - w_graph).abs().max())
    # eager launches of the same kernels vs the graph replay (generation counter, hooks, pulls)
    w_eager, _, loss_e, _ = _train(rank, world, **{"bigdl.cudaGraph": False})
    out["graph_vs_eager"] = float((w_graph - w_eager).abs().max())
    w_graph2, _, _, _ = _train(rank, world)
    w_eager2, _, _, _ = _train(rank, world, **{"bigdl.cudaGraph": False})
    out["graph_rerun"] = float((w_graph - w_graph2).abs().max())
    out["eager_rerun"] = float((w_eager - w_eager2).abs().max())
    # NCCL reduce-scatter / all-gather baseline
    w_nccl, _, loss_n, _ = _train(rank, world, **{"bigdl.parameter.backend": "nccl"})
    out["p2p_vs_nccl"] = float((w_graph - w_nccl).abs().max())
    out["scale"] = float(w_nccl.abs().max())
    # one bucket (no overlap) and the round-1 monolithic kernel
    w_one, _, _, st1 = _train(rank, world, **{"bigdl.parameter.overlap": False})
    out["one_bucket"] = st1["buckets"]
    out["overlap_vs_one"] = float((w_graph - w_one).abs().max())
    w_mono, _, _, _ = _train(rank, world, **{"bigdl.parameter.p2p": "monolithic"})
    out["bucketed_vs_monolithic"] = float((w_graph - w_mono).abs().max())
    # reference-exact arithmetic (truncating codec, adds in partition order) on the bucketed plane vs one bucket
    w_pt, _, _, _ = _train(rank, world, **{"bigdl.parameter.compress": "bf16_trunc"})
    w_pt1, _, _, _ = _train(rank, world, **{"bigdl.parameter.compress": "bf16_trunc", "bigdl.parameter.overlap": False})
    out["parity_overlap_vs_one"] = float((w_pt - w_pt1).abs().max())
    # global-norm clipping: device-side norm exchange inside the captured graph vs NCCL + host norm
    w_cg, _, _, _ = _train(rank, world, clip=0.05)
    w_cn, _, _, _ = _train(rank, world, clip=0.05, **{"bigdl.parameter.backend": "nccl"})
    out["clip_p2p_vs_nccl"] = float((w_cg - w_cn).abs().max())
    out["clip_changes_result"] = float((w_cg - w_graph).abs().max())
    # per-submodule OptimMethods on the sharded vector, graph-captured, vs the all-reduce-everything baseline
    w_mg, _, _, _ = _train(rank, world, multi=True)
    w_mn, _, _, _ = _train(rank, world, multi=True, **{"bigdl.parameter.backend": "nccl"})
    out["multi_p2p_vs_nccl"] = float((w_mg - w_mn).abs().max())
    out["multi_changes_result"] = float((w_mg - w_graph).abs().max())
    ref = w_mg.cuda()
    dist.broadcast(ref, 0)
    out["multi_ranks_identical"] = bool(torch.equal(ref.cpu(), w_mg))
    out["losses"] = (loss_g, loss_e, loss_n)
    return out


@pytest.mark.parametrize("world", [2, 4, 8])
def test_bucketed_plane_on_real_peers(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    from dist_utils import run_gpu_ranks

    res = run_gpu_ranks(_worker, world, timeout=900)
    print(res[0])
    for r in res:
        tol = 2e-2 * r["scale"] + 1e-3
        assert r["buckets"] >= 3 and r["overlapped"] > 0, r           # the exchange really ran from inside backward
        # all-gather fused into the Linear forward == rows pushed by their owners (up to the arrival order of the split-K
        # red.adds of that 4096-deep GEMM, ~1e-4)
        # (a rank must own at least one whole 128-row tile of the 512-row weight for the fused gather to engage: world <= 4)
        assert (r["gathered_gemms"] > 0 or world > 4) and r["gather_vs_push"] <= 1e-3, r
        assert r["ranks_identical"] and r["masters_identical"] and r["multi_ranks_identical"], r
        assert r["master_matches_compute"] == 0.0, r
        # same kernels, same order: replaying the captured graph and issuing the step eagerly agree up to the arrival
        # order of split-K red.adds (measured 0 .. 1.2e-4) — a stale barrier word or a bucket sent before its gradients
        # were final shows up here as a ~5e-3 deviation
        assert r["graph_vs_eager"] <= 1e-3 and r["graph_rerun"] <= 1e-3 and r["eager_rerun"] <= 1e-3, r
        assert r["one_bucket"] == 1 and r["bucketed_vs_monolithic"] <= tol, r
        # who owns an element changes the order of the fp32 adds over the ranks: exact for 2 ranks, rounding beyond
        assert r["overlap_vs_one"] <= (1e-3 if world == 2 else tol), r
        assert r["p2p_vs_nccl"] <= tol, r
        assert r["parity_overlap_vs_one"] <= tol, r
        assert r["clip_p2p_vs_nccl"] <= tol and r["clip_changes_result"] > 0, r
        # the baseline all-reduces UNcompressed fp32 gradients; Adagrad's first steps are lr * g / |g|, so a bf16-rounded
        # gradient near zero moves an element by a few lr (0.002) — hence the wider bound for this comparison
        assert r["multi_p2p_vs_nccl"] <= 0.02 and r["multi_changes_result"] > 0, r
