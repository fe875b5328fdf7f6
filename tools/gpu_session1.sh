#!/bin/bash
# First GPU session: tests, smoke, bench (autograd engine), kernel microbench, ncu captures.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_autograd.json 2> gpurun_out/bench_autograd.err; echo "bench exit $?"; cat gpurun_out/bench_autograd.json; tail -5 gpurun_out/bench_autograd.err
timeout 300 python tools/prof_agg.py > gpurun_out/prof_agg.txt 2>&1; echo "prof exit $?"; cat gpurun_out/prof_agg.txt
timeout 300 python tools/prof_agg.py --iters 2 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'fedavg_reduce|rows_sumsq|bcast_kernel|sgd_kernel' -s 8 -c 6 -o gpurun_out/prof_agg python tools/prof_agg.py --iters 2 > gpurun_out/ncu_agg.log 2>&1; echo "ncu exit $?"; tail -3 gpurun_out/ncu_agg.log
