#!/bin/bash
# Final multi-GPU numbers of the round.  usage: tools/gpu_session_final_mgpu.sh G
# Bench lines (MNIST CNN, ResNet-18, DP-ResNet-18, GroupNorm ResNet-18), the multi-GPU test and the
# latency-regime points of the aggregation sweep with the in-kernel phase trace.  The 350 M / 1 B
# points are in profiles/r2_n8 (tools/gpu_session_r2.sh; the streaming path has not changed since).
G=${1:-8}
O=gpurun_out/final_n$G
mkdir -p "$O"
port=29850
run() { port=$((port + 1)); timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$G" --master-addr 127.0.0.1 --master-port "$port" "$@"; }
b() { name=$1; shift; run bench.py --gpus "$G" "$@" 2>/dev/null | grep '^{' > "$O/$name.json"; echo "$name: $(cut -c1-210 "$O/$name.json")"; }
b bench_mnist --steps 20 --warmup 3
b bench_resnet18 --steps 5 --warmup 3 --model resnet18
b bench_resnet18_dp --steps 5 --warmup 3 --model resnet18 --dp
b bench_resnet18_gn --steps 5 --warmup 3 --model resnet18_gn
timeout 300 python -m pytest tests/test_multi_gpu.py -q 2>&1 | tail -2 > "$O/test_multi_gpu.log"; cat "$O/test_multi_gpu.log"
: > "$O/agg_sweep_latency.log"
for c in 8 32 64; do
  NANOFED_B200_AGG_TRACE=1 run tools/agg_sweep.py --both --params 1199936 --clients-per-gpu "$c" 2>/dev/null | grep '^{' | tee -a "$O/agg_sweep_latency.log" | cut -c1-420
done
