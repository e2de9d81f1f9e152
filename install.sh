#!/usr/bin/env bash
# Build the sm_100a kernel library and install the package in editable mode.
#   ./install.sh            # uses the active Python environment
set -euo pipefail
cd "$(dirname "$0")"
command -v nvcc >/dev/null || { echo "nvcc (CUDA >= 12.8) is required"; exit 1; }
python build_kernels.py                 # csrc/ -> llm_training_b200/_C/*.so (no GPU needed)
python -m pip install --no-build-isolation -e .
python - <<'PY'
import llm_training_b200 as m
from llm_training_b200.ops import _lib
print("llm_training_b200", m.__version__, "kernels:", "ok" if _lib.kernels_available() else "not loaded (CPU only)")
PY
