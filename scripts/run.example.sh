#!/usr/bin/env bash
# Run one of the bundled examples end to end on synthetic data (reference: scripts/run.example.sh).
#   scripts/run.example.sh lenet|inception|resnet|vgg|rnn|textclassifier|int8|keras [extra args]
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
EX="${1:-lenet}"; shift || true
case "$EX" in
  lenet) exec "$ROOT/scripts/bigdl-b200-run.sh" "$ROOT/examples/lenet_local.py" "$@";;
  inception|resnet|vgg) exec "$ROOT/scripts/bigdl-b200-run.sh" "$ROOT/examples/imagenet_train.py" --model "$EX" "$@";;
  rnn) exec "$ROOT/scripts/bigdl-b200-run.sh" "$ROOT/examples/language_model.py" "$@";;
  textclassifier) exec "$ROOT/scripts/bigdl-b200-run.sh" "$ROOT/examples/text_classification.py" "$@";;
  int8) exec "$ROOT/scripts/bigdl-b200-run.sh" "$ROOT/examples/int8_inference.py" "$@";;
  keras) exec "$ROOT/scripts/bigdl-b200-run.sh" "$ROOT/examples/keras_mnist_cnn.py" "$@";;
  *) echo "unknown example $EX" >&2; exit 2;;
esac
