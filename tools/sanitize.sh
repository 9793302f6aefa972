#!/bin/bash
# compute-sanitizer passes over the hand-rolled mbarrier / TMEM / TMA pipelines and the grouped kernels:
#   tools/sanitize.sh [memcheck|racecheck|synccheck] [pytest -k expression]
# Run on the GPU box (gpurun -- 'bash tools/sanitize.sh racecheck'); the summary lines go to stdout.
TOOL=${1:-memcheck}
SEL=${2:-"tcgen05 or wgrad or grouped or modconv"}
cd "$(dirname "$0")/.."
timeout ${SANITIZE_TIMEOUT:-1500} compute-sanitizer --tool "$TOOL" --print-limit 20 --error-exitcode 9 \
  python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "$SEL" 2>&1 | grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Error|hazard" | tail -20
echo "compute-sanitizer --tool $TOOL exit: ${PIPESTATUS[0]}"
