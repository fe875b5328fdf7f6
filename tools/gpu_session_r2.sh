#!/bin/bash
# Round-2 multi-GPU evidence.  usage: tools/gpu_session_r2.sh G   (G = GPUs on this box: 2, 4 or 8)
# Writes gpurun_out/r2_n$G/: bench lines (MNIST CNN, ResNet-18, DP-ResNet-18), the multi-GPU test,
# and the config-5 aggregation sweep (back-to-back timing, P2P and NVLS, streaming waves).
G=${1:-8}
O=gpurun_out/r2_n$G
mkdir -p "$O"
port=29800
run() { port=$((port + 1)); timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$G" --master-addr 127.0.0.1 --master-port "$port" "$@"; }
run bench.py --gpus "$G" --steps 20 --warmup 3 2>/dev/null | grep '^{' > "$O/bench_mnist.json"
run bench.py --gpus "$G" --steps 5 --warmup 3 --model resnet18 2>/dev/null | grep '^{' > "$O/bench_resnet18.json"
run bench.py --gpus "$G" --steps 5 --warmup 3 --model resnet18 --dp 2>/dev/null | grep '^{' > "$O/bench_resnet18_dp.json"
timeout 300 python -m pytest tests/test_multi_gpu.py -q 2>&1 | tail -2 > "$O/test_multi_gpu.log"
sweep() { run tools/agg_sweep.py "$@" 2>/dev/null | grep '^{' | tee -a "$O/agg_sweep.log"; }
: > "$O/agg_sweep.log"
sweep --both --params 1199936 350000000 1000000000 --clients-per-gpu 8
sweep --both --params 1199936 --clients-per-gpu 64
sweep --both --params 350000000 --clients-per-gpu 16
# 512 / G x G clients of bf16 rows: GPT-2-medium size resident, 1 B in 4 streaming waves
C=$((512 / G)); [ "$C" -gt 64 ] && C=64
sweep --both --dtype bf16 --params 350000000 --clients-per-gpu "$C" --iters 3 --b2b 5
sweep --both --dtype bf16 --params 1000000000 --clients-per-gpu 16 --iters 3 --b2b 5
sweep --both --dtype bf16 --params 1000000000 --clients-per-gpu "$C" --waves 4 --iters 3 --b2b 3
for f in "$O"/bench_*.json; do echo "$f: $(cut -c1-260 "$f")"; done
cat "$O/test_multi_gpu.log"
