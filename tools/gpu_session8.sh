#!/bin/bash
# tensor-core conv1 forward + conflict-free pool_bwd transposition
mkdir -p gpurun_out
timeout 120 python tools/tc_check.py conv1_fwd > gpurun_out/s8_conv1.log 2>&1; echo "conv1 rc=$?"; tail -6 gpurun_out/s8_conv1.log
timeout 300 python -m pytest tests/test_mnist_cohort_gpu.py tests/test_tc_gpu.py -x -q > gpurun_out/s8_tests.log 2>&1; echo "tests rc=$?"; tail -12 gpurun_out/s8_tests.log
timeout 200 python tools/prof_step.py > gpurun_out/s8_step.log 2>&1; tail -1 gpurun_out/s8_step.log
NFB_NO_TC_CONV1=1 timeout 200 python tools/prof_step.py > gpurun_out/s8_step_old.log 2>&1; tail -1 gpurun_out/s8_step_old.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 44 --csv --log-file gpurun_out/launches_step17.csv python tools/prof_step.py 1 > gpurun_out/s8_ncu.log 2>&1; echo "ncu rc=$?"
python tools/launch_list.py gpurun_out/launches_step17.csv 11
