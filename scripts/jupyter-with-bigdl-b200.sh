#!/usr/bin/env bash
# Start a Jupyter notebook server whose kernels can import the framework (reference: scripts/jupyter-with-bigdl.sh).
#   scripts/jupyter-with-bigdl-b200.sh [--port 8888] [jupyter args...]
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export PYTHONPATH="$ROOT${PYTHONPATH:+:$PYTHONPATH}"
python -c "import bigdl_b200.build as b; b.build_all()" >/dev/null
command -v jupyter >/dev/null 2>&1 || { echo "jupyter is not installed in this environment" >&2; exit 1; }
exec jupyter notebook --notebook-dir="${NOTEBOOK_DIR:-$PWD}" --ip=127.0.0.1 --no-browser "$@"
