#!/bin/bash
# Race / memory checks for the kernels.  ONE compute-sanitizer tool per GPU call
# (see the profiling recipe): pass the tool name.  Uses the smallest cases of the GPU tests.
#   gpurun -- bash scripts/sanitize.sh memcheck|racecheck|synccheck|initcheck
set -e
TOOL=${1:-memcheck}
compute-sanitizer --tool "$TOOL" --error-exitcode 7 \
  python -m pytest tests/test_ops_gpu.py -q -x -m gpu -k "philox_matches or update_stats and 3-1024 or bcast or xent or mask_add"
