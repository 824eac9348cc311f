#!/usr/bin/env bash
# Set the package version everywhere it is recorded (reference: scripts/bump-version.sh rewrites the Maven poms and
# pyspark/bigdl/version.py). Here the single source is bigdl_b200/__init__.py; setup.py and bigdl/version.py read it.
#   scripts/bump-version.sh 0.2.0
set -euo pipefail
[ $# -eq 1 ] || { echo "usage: $0 NEW_VERSION" >&2; exit 2; }
case "$1" in *[!0-9A-Za-z.+-]*|"") echo "not a version: $1" >&2; exit 2;; esac
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
F="$ROOT/bigdl_b200/__init__.py"
OLD="$(sed -n 's/^__version__ = "\(.*\)"/\1/p' "$F")"
sed -i "s/^__version__ = \".*\"/__version__ = \"$1\"/" "$F"
echo "version $OLD -> $1"
