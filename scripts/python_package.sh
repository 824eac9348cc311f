#!/usr/bin/env bash
# Build the distributable wheel with the native libraries inside (reference: scripts/python_package.sh builds the jar
# and packages the Python API with it). Output: dist/bigdl_b200-<version>-*.whl
#   scripts/python_package.sh [linux]
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
cd "$ROOT"
python -c "import bigdl_b200.build as b; b.build_all()"
exec make dist
