"""Per-iteration losses of the small multi-GPU test model under different parameter-plane settings (debug aid).
torchrun --nproc-per-node N benchmarks/comm_debug.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import torch.distributed as dist


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    from bigdl_b200.utils.engine import Engine

    Engine.init()
    import test_comm_multigpu as t
    from bigdl_b200.optim import optimizer as O

    losses = {}
    orig = O.Optimizer.setIterationCallback if hasattr(O.Optimizer, "setIterationCallback") else None
    ws = {}
    multi = bool(int(os.environ.get("MULTI", "0")))
    for name, cfg in [("graph", {"multi": multi}), ("eager", {"bigdl.cudaGraph": False, "multi": multi}),
                      ("nccl", {"bigdl.parameter.backend": "nccl", "multi": multi}),
                      ("eager_one", {"bigdl.cudaGraph": False, "bigdl.parameter.overlap": False, "multi": multi})]:
        cur = []
        real_init = O.Optimizer.__init__

        def patched(self, *a, **k):
            real_init(self, *a, **k)
            self._iteration_callback = lambda state, loss: cur.append(float(loss.detach()))

        O.Optimizer.__init__ = patched
        try:
            w, m, l, st = t._train(rank, world, iters=6, **cfg)
        finally:
            O.Optimizer.__init__ = real_init
        losses[name] = (cur, float(w.double().sum()), float(w.abs().max()))
        ws[name] = w
    if rank == 0:
        for k, v in losses.items():
            print(k, [round(x, 5) for x in v[0]], v[1], v[2])
    if rank == 0:
        m = t._model()
        m.getParameters()
        for mod, n, off, numel in m._flat.entries:
            d = {k: round(float((ws[k][off:off + numel] - ws["nccl"][off:off + numel]).abs().max()), 5) for k in ws if k != "nccl"}
            print(type(mod).__name__, n, off, numel, d)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
