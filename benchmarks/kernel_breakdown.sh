#!/bin/bash
# Per-kernel device-time list of one eager training step (ncu gpu__time_duration only) + GEMM shape log.
# usage: benchmarks/kernel_breakdown.sh <model> <tag>      (run under gpurun; output in gpurun_out/)
set -u
M=${1:-inception_v1}; T=${2:-$M}
mkdir -p gpurun_out
timeout 300 python bench.py --model $M --steps 1 --warmup 3 --no-graph --no-e2e --no-gpu-baseline > gpurun_out/plain_$T.log 2>&1 || { echo "plain run failed"; tail -3 gpurun_out/plain_$T.log; exit 1; }
rm -f gpurun_out/shapes_$T.jsonl
BDL_LOG_GEMM=gpurun_out/shapes_$T.jsonl timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv \
  --log-file gpurun_out/launches_$T.csv python bench.py --model $M --steps 1 --warmup 3 --no-graph --no-e2e --no-gpu-baseline > gpurun_out/ncu_$T.log 2>&1
tail -1 gpurun_out/ncu_$T.log | cut -c1-120
