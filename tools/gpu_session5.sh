#!/bin/bash
# bring-up of conv2 dgrad fused with the conv1 weight gradient (tc_conv MODE 3)
mkdir -p gpurun_out
timeout 120 python tools/tc_check.py conv_dgrad_wgrad1 > gpurun_out/s5_check.log 2>&1; echo "check rc=$?"; tail -8 gpurun_out/s5_check.log
timeout 120 python tools/tc_check.py conv_dgrad_p >> gpurun_out/s5_check.log 2>&1; echo "check2 rc=$?"
timeout 400 python -m pytest tests/test_mnist_cohort_gpu.py tests/test_tc_gpu.py -x -q > gpurun_out/s5_tests.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/s5_tests.log
timeout 200 python tools/prof_step.py > gpurun_out/s5_step.log 2>&1; tail -2 gpurun_out/s5_step.log
NFB_NO_FUSED_WGRAD1=1 timeout 200 python tools/prof_step.py > gpurun_out/s5_step_unfused.log 2>&1; tail -2 gpurun_out/s5_step_unfused.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/s5_bench.json 2> gpurun_out/s5_bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/s5_bench.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_step14.csv python tools/prof_step.py 1 > gpurun_out/s5_ncu.log 2>&1; echo "ncu rc=$?"
