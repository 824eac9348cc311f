#!/usr/bin/env bash
# Interactive Python with the framework importable, the native libraries built and conf/bigdl-b200.conf applied
# (reference: scripts/pyspark-with-bigdl.sh / spark-shell-with-bigdl.sh). Extra arguments go to python.
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export PYTHONPATH="$ROOT${PYTHONPATH:+:$PYTHONPATH}"
export PYTHONSTARTUP="$ROOT/scripts/.startup.py"
cat > "$PYTHONSTARTUP" <<'PY'
from bigdl_b200.utils.engine import Engine
Engine.init()
from bigdl.nn.layer import *          # noqa: F401,F403
from bigdl.nn.criterion import *      # noqa: F401,F403
from bigdl.optim.optimizer import *   # noqa: F401,F403
from bigdl.util.common import *       # noqa: F401,F403
print("bigdl_b200 ready: engine on", Engine.device)
PY
python -c "import bigdl_b200.build as b; b.build_all()" >/dev/null
exec python -i "$@"
