#!/bin/bash
# mask/window prefetch in the dgrad epilogue + full ncu capture of the step's tcgen05 kernels
mkdir -p gpurun_out
timeout 120 python tools/tc_check.py conv_dgrad_wgrad1 > gpurun_out/s6_check.log 2>&1; echo "check rc=$?"; tail -3 gpurun_out/s6_check.log
timeout 120 python tools/tc_check.py conv_dgrad_p >> gpurun_out/s6_check.log 2>&1; echo "check2 rc=$?"; tail -1 gpurun_out/s6_check.log
timeout 300 python -m pytest tests/test_mnist_cohort_gpu.py -x -q > gpurun_out/s6_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/s6_tests.log
timeout 200 python tools/prof_step.py > gpurun_out/s6_step.log 2>&1; tail -1 gpurun_out/s6_step.log
timeout 100 python tools/conv_probe.py 2>&1 | grep debug
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"tc_conv_kernel|tc_conv_wgrad|tc_gemm_kernel" -s 12 -c 6 -f -o gpurun_out/ncu_step15 python tools/prof_step.py 1 > gpurun_out/s6_ncu.log 2>&1; echo "ncu rc=$?"; ls -la gpurun_out/ncu_step15.ncu-rep
