#!/bin/bash
# 8-GPU session: scaling bench, peer-kernel correctness (P2P + NVLS), aggregation sweep.
mkdir -p gpurun_out
G=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29531 bench.py --gpus $G --steps 10 --warmup 3 > gpurun_out/bench_ours_n$G.json 2> gpurun_out/bench_ours_n$G.err; echo "bench rc=$?"; cat gpurun_out/bench_ours_n$G.json | cut -c1-700
NANOFED_B200_NVLS=1 timeout 300 $TR --master-port 29532 tools/check_peer.py > gpurun_out/check_peer_n$G.log 2>&1; echo "check rc=$?"; grep "::" gpurun_out/check_peer_n$G.log
timeout 400 $TR --master-port 29533 tools/agg_sweep.py > gpurun_out/agg_sweep_n$G.log 2>&1; echo "sweep rc=$?"; grep params gpurun_out/agg_sweep_n$G.log; tail -2 gpurun_out/agg_sweep_n$G.log | cut -c1-300
timeout 400 $TR --master-port 29534 tools/agg_sweep.py --nvls > gpurun_out/agg_sweep_nvls_n$G.log 2>&1; echo "sweep nvls rc=$?"; grep params gpurun_out/agg_sweep_nvls_n$G.log
