#!/bin/bash
# bring-up of the fused fc1 wgrad + SGD epilogue: isolated numerics, engine tests, timings
mkdir -p gpurun_out
timeout 120 python tools/tc_check.py fc_wgrad_sgd > gpurun_out/s4_check.log 2>&1; echo "check rc=$?"; tail -5 gpurun_out/s4_check.log
timeout 120 python tools/tc_check.py fc_wgrad >> gpurun_out/s4_check.log 2>&1; echo "check2 rc=$?"
timeout 400 python -m pytest tests/test_mnist_cohort_gpu.py tests/test_ops_gpu.py tests/test_tc_gpu.py -x -q > gpurun_out/s4_tests.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/s4_tests.log
timeout 200 python tools/prof_step.py > gpurun_out/s4_step.log 2>&1; tail -2 gpurun_out/s4_step.log
NFB_NO_FUSED_SGD=1 timeout 200 python tools/prof_step.py > gpurun_out/s4_step_unfused.log 2>&1; tail -2 gpurun_out/s4_step_unfused.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/s4_bench.json 2> gpurun_out/s4_bench.err; echo "bench rc=$?"; cat gpurun_out/s4_bench.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_step13.csv python tools/prof_step.py 1 > gpurun_out/s4_ncu.log 2>&1; echo "ncu rc=$?"
