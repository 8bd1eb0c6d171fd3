#!/usr/bin/env bash
# Environment check + native build (reference: scripts/initial_install.sh installs a venv and
# hivemind 1.1.11; neither is needed here — the control plane is self-contained TCP).
set -euo pipefail
cd "$(dirname "$0")/.."
python - <<'PY'
import torch, msgpack, numpy
print("torch", torch.__version__, "| cuda", torch.version.cuda, "| gpus", torch.cuda.device_count())
PY
python csrc/build.py
python -c "import __graft_entry__ as g; g.build(); print('build ok')"
