#!/bin/bash
# ncu --set full captures of the kernels that had none (BN, LRN, optimizer, xent, softmax, fp8 GEMM, LSTM, comm bucket
# kernel on virtual ranks). One GPU. Every program is first run WITHOUT ncu and must exit 0. Reports land in gpurun_out/.
set -u
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
run() {  # name, kernel regex, skip, count, command...
  local name=$1 regex=$2 skip=$3 count=$4; shift 4
  timeout 300 "$@" > gpurun_out/plain_$name.log 2>&1 || { echo "$name: plain run failed"; tail -3 gpurun_out/plain_$name.log; return; }
  timeout 900 $NCU -k "regex:$regex" -s $skip -c $count -f -o gpurun_out/ncu_$name "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name: $(grep -c '==PROF== Profiling' gpurun_out/ncu_$name.log) kernels captured"
}
run resnet_bn_opt   'bn_|optim_update|xent_'          330 10 python bench.py --model resnet50 --steps 1 --warmup 2 --no-graph --no-e2e --no-gpu-baseline
run inception_misc  'lrn_|colsum|softmax|addn_|xent_' 120 12 python bench.py --model inception_v1 --steps 1 --warmup 2 --no-graph --no-e2e --no-gpu-baseline
run lstm            'lstm_seq'                        4   2  python benchmarks/lstm_one.py 1500
run fp8             'gemm_bf16_tcgen05|quantize|absmax' 200 6 python benchmarks/quant_infer_bench.py --batches 64 --iters 3 --warmup 1 --no-graph --modes fp8
run comm_virtual    'bucket_rs_update|pull_gather|peer_clip' 0 6 python -m pytest tests/test_comm_gpu.py -q -x -k "bucketed or clip"
