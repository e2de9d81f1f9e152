#!/usr/bin/env bash
# compute-sanitizer targets for the sm_100a kernels.  Needs one B200.
#   scripts/sanitize.sh [memcheck|synccheck|racecheck|initcheck] [pytest -k expression]
# memcheck / synccheck cover every kernel test; racecheck only sees shared-memory hazards between generic-proxy
# accesses, so the TMA / tcgen05 (async proxy) traffic of the GEMM and attention kernels is outside its model — use it
# for the elementwise / reduction kernels.
set -euo pipefail
cd "$(dirname "$0")/.."
TOOL="${1:-memcheck}"
EXPR="${2:-rmsnorm or silu or rope or embedding or cross_entropy or adamw or gemm or attention or emulated}"
mkdir -p gpurun_out
compute-sanitizer --tool "$TOOL" --error-exitcode 7 --launch-timeout 120 \
    python -m pytest tests/test_kernels_gpu.py -x -q -k "$EXPR" 2>&1 | tee "gpurun_out/sanitize_${TOOL}.log" | tail -25
