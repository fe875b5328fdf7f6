#!/bin/bash
# bring-up of the persistent fc data-gradient kernel (tc_fc.cu)
mkdir -p gpurun_out
for k in fc_dgrad fc_dgrad_ragged; do timeout 120 python tools/tc_check.py $k > gpurun_out/s7_$k.log 2>&1; echo "$k rc=$?"; tail -4 gpurun_out/s7_$k.log; done
timeout 300 python -m pytest tests/test_mnist_cohort_gpu.py tests/test_tc_gpu.py -x -q > gpurun_out/s7_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/s7_tests.log
timeout 200 python tools/prof_step.py > gpurun_out/s7_step.log 2>&1; tail -1 gpurun_out/s7_step.log
NFB_NO_PERSISTENT_FC=1 timeout 200 python tools/prof_step.py > gpurun_out/s7_step_old.log 2>&1; tail -1 gpurun_out/s7_step_old.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 44 --csv --log-file gpurun_out/launches_step16.csv python tools/prof_step.py 1 > gpurun_out/s7_ncu.log 2>&1; echo "ncu rc=$?"
python tools/launch_list.py gpurun_out/launches_step16.csv 11
