#!/bin/bash
# Race / memory checks for the kernels.  ONE compute-sanitizer tool per GPU call
# (see the profiling recipe): pass the tool name.  Uses the smallest cases of the GPU tests.
#   gpurun -- bash scripts/sanitize.sh memcheck|racecheck|synccheck|initcheck
set -e
TOOL=${1:-memcheck}
if [ "${2:-}" != "tc" ]; then
compute-sanitizer --tool "$TOOL" --error-exitcode 7 \
  python -m pytest tests/test_ops_gpu.py -q -x -m gpu -k "philox_matches or update_stats and 3-1024 or bcast or xent or mask_add"
fi
# tcgen05 / TMA kernels of the flagship step (small shapes, one process):
#   gpurun -- bash scripts/sanitize.sh memcheck tc
if [ "${2:-}" = "tc" ]; then
  CASES="conv1_fwd conv_pool fc_fwd fc_dgrad fc_wgrad_sgd conv_wgrad_p conv_dgrad_wgrad1 conv_dgrad_p"
  python tools/tc_check.py --inproc $CASES
  compute-sanitizer --tool "$TOOL" --error-exitcode 7 python tools/tc_check.py --inproc $CASES
fi
