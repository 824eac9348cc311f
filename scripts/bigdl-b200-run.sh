#!/usr/bin/env bash
# Launch a bigdl_b200 program on N GPUs of this node, one process per GPU (the role spark-submit-with-bigdl.sh plays for
# the reference: inject the runtime, set the engine flags, start the workers).
#   scripts/bigdl-b200-run.sh [-n GPUS] [-p PORT] module_or_script [args...]
#   scripts/bigdl-b200-run.sh -n 8 -m bigdl_b200.cli.distri_optimizer_perf -m inception_v1 -b 128
set -euo pipefail
GPUS=1; PORT=29500; MODULE=0
while getopts "n:p:m" o; do
  case "$o" in n) GPUS="$OPTARG";; p) PORT="$OPTARG";; m) MODULE=1;; *) exit 2;; esac
done
shift $((OPTIND - 1))
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export PYTHONPATH="$ROOT${PYTHONPATH:+:$PYTHONPATH}"
# properties of conf/bigdl-b200.conf become environment overrides (bigdl.foo.bar -> BIGDL_FOO_BAR), like spark-bigdl.conf
if [ -f "$ROOT/conf/bigdl-b200.conf" ]; then
  while IFS='=' read -r k v; do
    case "$k" in ''|\#*) continue;; esac
    k="$(echo "$k" | tr '[:lower:].' '[:upper:]_' | tr -d ' ')"; v="$(echo "$v" | sed 's/^ *//;s/ *$//')"
    [ -z "${!k:-}" ] && export "$k=$v"
  done < "$ROOT/conf/bigdl-b200.conf"
fi
python -c "import bigdl_b200.build as b; b.build_all()" >/dev/null
if [ "$GPUS" -le 1 ]; then
  if [ "$MODULE" = 1 ]; then exec python -m "$@"; else exec python "$@"; fi
fi
ARGS=(--nnodes=1 --nproc-per-node "$GPUS" --master-addr 127.0.0.1 --master-port "$PORT")
if [ "$MODULE" = 1 ]; then exec python -m torch.distributed.run "${ARGS[@]}" -m "$@"; else exec python -m torch.distributed.run "${ARGS[@]}" "$@"; fi
