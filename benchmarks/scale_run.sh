#!/bin/bash
# usage: benchmarks/scale_run.sh <N> <out.jsonl> [models...]   (under gpurun --gpus N)
N=$1; OUT=$2; shift 2
MODELS=${@:-inception_v1 resnet50 vgg16}
mkdir -p gpurun_out; : > $OUT
port=29600
for m in $MODELS; do
  port=$((port+1))
  if [ "$N" = "1" ]; then
    timeout 400 python bench.py --model $m --gpus 1 --steps 30 --warmup 5 2>gpurun_out/scale_err_$m.log | tail -1 >> $OUT
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
      bench.py --model $m --gpus $N --steps 30 --warmup 5 2>gpurun_out/scale_err_$m.log | tail -1 >> $OUT
  fi
done
cut -c1-230 $OUT
