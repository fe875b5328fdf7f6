#!/bin/bash
# MMA issue path: descriptor address fields precomputed on the host
mkdir -p gpurun_out
for k in conv_pool conv_wgrad_p conv_fwd_p conv_dgrad_p conv_dgrad_wgrad1; do timeout 120 python tools/tc_check.py $k >> gpurun_out/s14_check.log 2>&1; echo "$k rc=$?"; done; tail -5 gpurun_out/s14_check.log
timeout 300 python -m pytest tests/test_mnist_cohort_gpu.py tests/test_tc_gpu.py -x -q > gpurun_out/s14_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/s14_tests.log
timeout 200 python tools/prof_step.py > gpurun_out/s14_step.log 2>&1; tail -1 gpurun_out/s14_step.log
timeout 100 python tools/conv_probe.py 2>&1 | grep debug
