"""Helpers to run N CPU ranks (gloo) inside a test, the analogue of the reference's "Spark local[1] pretending to be
4 nodes" trick (T/optim/DistriOptimizerSpec.scala:44-49)."""
import os
import socket
import sys
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, args, ret):
    os.environ.update({"RANK": str(rank), "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank),
                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    torch.set_num_threads(1)
    try:
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
        from bigdl_b200.utils.engine import Engine

        Engine.init(device="cpu")
        ret[rank] = fn(rank, world, *args)
    except Exception:  # noqa: BLE001
        ret[rank] = ("__error__", traceback.format_exc())
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def run_ranks(fn, world: int, *args):
    """Run ``fn(rank, world, *args)`` on ``world`` gloo processes; returns the list of per-rank results. The port is
    picked by bind-and-release, so under ``pytest -n`` another worker can take it first: a rendezvous failure is
    retried once on a fresh port."""
    try:
        return _run_ranks_once(fn, world, *args)
    except RuntimeError as e:
        msg = str(e)
        if not any(k in msg for k in ("Address already in use", "EADDRINUSE", "Connection refused", "connect()",
                                      "timed out", "Connection reset")):
            raise
        return _run_ranks_once(fn, world, *args)


def _run_ranks_once(fn, world: int, *args):
    ctx = mp.get_context("spawn")
    mgr = ctx.Manager()
    ret = mgr.dict()
    port = free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    for p in procs:
        if p.is_alive():
            p.kill()
            raise RuntimeError("rank timed out")
    out = [ret.get(r) for r in range(world)]
    for r, o in enumerate(out):
        if isinstance(o, tuple) and len(o) == 2 and isinstance(o[0], str) and o[0] == "__error__":
            raise RuntimeError(f"rank {r} failed:\n{o[1]}")
    return out


def _gpu_entry(rank, world, port, fn, args, ret):
    os.environ.update({"RANK": str(rank), "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank),
                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    try:
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                                device_id=torch.device("cuda", rank))
        from bigdl_b200.utils.engine import Engine

        Engine.init()
        ret[rank] = fn(rank, world, *args)
        torch.cuda.synchronize()
    except Exception:  # noqa: BLE001
        ret[rank] = ("__error__", traceback.format_exc())
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def run_gpu_ranks(fn, world: int, *args, timeout: int = 600):
    """Run ``fn(rank, world, *args)`` on ``world`` processes, one per GPU of this box, NCCL as the control plane."""
    ctx = mp.get_context("spawn")
    mgr = ctx.Manager()
    ret = mgr.dict()
    port = free_port()
    procs = [ctx.Process(target=_gpu_entry, args=(r, world, port, fn, args, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=timeout)
    for p in procs:
        if p.is_alive():
            for q in procs:
                if q.is_alive():
                    q.kill()
            raise RuntimeError("a GPU rank timed out")
    out = [ret.get(r) for r in range(world)]
    for r, o in enumerate(out):
        if isinstance(o, tuple) and len(o) == 2 and isinstance(o[0], str) and o[0] == "__error__":
            raise RuntimeError(f"rank {r} failed:\n{o[1]}")
    return out
