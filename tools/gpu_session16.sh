#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/tc_check.py conv_pool 2>&1 | tail -2
timeout 300 python -m pytest tests/test_mnist_cohort_gpu.py tests/test_tc_gpu.py -x -q 2>&1 | tail -2
timeout 200 python tools/prof_step.py 2>&1 | tail -1
timeout 100 python tools/conv_probe.py 2>&1 | grep "pool"
