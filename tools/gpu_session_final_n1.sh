#!/bin/bash
# Final 1-GPU numbers of the round (no profiler attached): bench lines for every engine/model pair
# quoted in BASELINE.md / README.md, the ResNet per-kernel profile and the numerics checks.
O=gpurun_out/final_n1
mkdir -p "$O"
b() { name=$1; shift; python bench.py "$@" 2>/dev/null | grep '^{' > "$O/$name.json"; echo "$name: $(cut -c1-200 "$O/$name.json")"; }
b bench_mnist --steps 20 --warmup 3
b bench_mnist_dp --steps 20 --warmup 3 --dp
b bench_mnist_local_dp --steps 20 --warmup 3 --local-dp
b bench_resnet18 --model resnet18 --steps 5 --warmup 3
b bench_resnet18_dp --model resnet18 --steps 5 --warmup 3 --dp
b bench_resnet18_gn --model resnet18_gn --steps 5 --warmup 3
b bench_resnet18_gn_local_dp --model resnet18_gn --local-dp --steps 5 --warmup 3
b bench_mnist_adam --optimizer adam --steps 20 --warmup 3
b bench_resnet18_gn_autograd --model resnet18_gn --engine autograd --steps 5 --warmup 3
b bench_resnet18_autograd --model resnet18 --engine autograd --steps 5 --warmup 3
python tools/rn_prof.py --reps 2 > "$O/rn_prof_step.log" 2>&1; grep "local step" "$O/rn_prof_step.log"
python tools/rn_check.py > "$O/rn_check.log" 2>&1; tail -1 "$O/rn_check.log"
python tools/rn_step_check.py > "$O/rn_step_check.log" 2>&1; tail -1 "$O/rn_step_check.log"
python tools/rn_step_check.py --model resnet18_gn > "$O/rn_step_check_gn.log" 2>&1; tail -1 "$O/rn_step_check_gn.log"
python tools/tc_check.py > "$O/tc_check.log" 2>&1; tail -1 "$O/tc_check.log"
