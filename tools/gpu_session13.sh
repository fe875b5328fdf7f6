#!/bin/bash
# conv wgrad with a single x tile (kw via descriptor LBO = one row)
mkdir -p gpurun_out
timeout 120 python tools/tc_check.py conv_wgrad_p > gpurun_out/s13_wgrad.log 2>&1; echo "wgrad rc=$?"; tail -6 gpurun_out/s13_wgrad.log
timeout 300 python -m pytest tests/test_mnist_cohort_gpu.py tests/test_tc_gpu.py -x -q > gpurun_out/s13_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/s13_tests.log
timeout 200 python tools/prof_step.py > gpurun_out/s13_step.log 2>&1; tail -1 gpurun_out/s13_step.log
timeout 100 python tools/conv_probe.py 2>&1 | grep wgrad
