#!/bin/bash
# conv1 hi/lo split, fc_head split over row blocks, pool_bwd batched loads
mkdir -p gpurun_out
timeout 120 python tools/tc_check.py conv1_fwd > gpurun_out/s11_conv1.log 2>&1; echo "conv1 rc=$?"; tail -3 gpurun_out/s11_conv1.log
timeout 400 python -m pytest tests/test_mnist_cohort_gpu.py tests/test_tc_gpu.py tests/test_ops_gpu.py -x -q > gpurun_out/s11_tests.log 2>&1; echo "tests rc=$?"; tail -12 gpurun_out/s11_tests.log
timeout 200 python tools/prof_step.py > gpurun_out/s11_step.log 2>&1; tail -1 gpurun_out/s11_step.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 44 --csv --log-file gpurun_out/launches_step20.csv python tools/prof_step.py 1 > gpurun_out/s11_ncu.log 2>&1; echo "ncu rc=$?"
python tools/launch_list.py gpurun_out/launches_step20.csv 11
