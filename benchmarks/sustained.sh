#!/bin/bash
# >= 30 s device-timed runs of the three flagship models (1 GPU, batch 128, CUDA-graph step), clocks / power sampled at 20 ms.
# usage (under gpurun): benchmarks/sustained.sh ; lines land in gpurun_out/sustained_<model>.jsonl
mkdir -p gpurun_out
for spec in "inception_v1 6000" "resnet50 2100" "vgg16 1800"; do
  set -- $spec
  timeout 400 python bench.py --gpus 1 --model $1 --steps $2 --warmup 10 --no-e2e --no-gpu-baseline 2>/dev/null | tail -1 > gpurun_out/sustained_$1.jsonl
  python - "$1" <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/sustained_{sys.argv[1]}.jsonl").read())
print(sys.argv[1], d["value"], d["unit"], "ms/step", d["ms_per_step"], "timed_s", round(d["ms_per_step"] * d["steps"] / 1e3, 1), d.get("clocks"))
PY
done
