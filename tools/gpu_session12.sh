#!/bin/bash
# conv2+pool epilogue with four warpgroups
mkdir -p gpurun_out
timeout 120 python tools/tc_check.py conv_pool > gpurun_out/s12_pool.log 2>&1; echo "pool rc=$?"; tail -3 gpurun_out/s12_pool.log
for k in conv_fwd_p conv_dgrad_p conv_dgrad_wgrad1; do timeout 120 python tools/tc_check.py $k >> gpurun_out/s12_pool.log 2>&1; echo "$k rc=$?"; done
timeout 400 python -m pytest tests/test_mnist_cohort_gpu.py tests/test_tc_gpu.py -x -q > gpurun_out/s12_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/s12_tests.log
timeout 200 python tools/prof_step.py > gpurun_out/s12_step.log 2>&1; tail -1 gpurun_out/s12_step.log
timeout 100 python tools/conv_probe.py 2>&1 | grep debug
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 44 --csv --log-file gpurun_out/launches_step21.csv python tools/prof_step.py 1 > gpurun_out/s12_ncu.log 2>&1; echo "ncu rc=$?"
python tools/launch_list.py gpurun_out/launches_step21.csv 11
