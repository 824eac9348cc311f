"""All-reduce of a bf16 bucket: in-switch (NVLS, csrc/nvls.cu) vs NCCL, device-timed, max over ranks.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/nvls_bench.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
from bigdl_b200.parallel import nvls


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


ok = nvls.multicast_supported()
for mb in (1, 8, 64, 256):
    n = mb * (1 << 20) // 2
    x = torch.randn(n, device="cuda").bfloat16()
    out = torch.empty(n, dtype=torch.float32, device="cuda")
    row = {"bucket_MB_bf16": mb, "world": world}
    row["nccl_allreduce_ms"] = round(timed(lambda: dist.all_reduce(x)), 4)
    if ok:
        buf = nvls.NvlsBuffer(n, torch.bfloat16)
        buf.local.copy_(x)
        row["nvls_oneshot_to_f32_ms"] = round(timed(lambda: buf.reduce_to_f32(out, 0, True)), 4)
        row["nvls_twoshot_inplace_ms"] = round(timed(lambda: buf.all_reduce_inplace(0, n)), 4)
        row["nvls_oneshot_GBps_received"] = round(mb / 1024 / (row["nvls_oneshot_to_f32_ms"] / 1e3), 1)
        del buf
    if rank == 0:
        print(json.dumps(row), flush=True)
dist.destroy_process_group()
