#!/bin/bash
# one bench.py run at N GPUs, compact output. usage: benchmarks/bench_one.sh <ngpus> <model> [bench.py args...]
N=$1; M=$2; shift 2
if [ "$N" = "1" ]; then
  r=$(timeout 300 python bench.py --gpus 1 --model $M --steps 30 --warmup 5 --no-e2e "$@" 2>&1 | tail -1)
else
  r=$(timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus $N --model $M --steps 30 --warmup 5 --no-e2e "$@" 2>&1 | tail -1)
fi
echo "N=$N $M [$* ${BDL_COMM_THREADS:+threads=$BDL_COMM_THREADS}] $(echo "$r" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["config"]["parameter_backend"])' 2>/dev/null || echo "FAILED: ${r:0:300}")"
