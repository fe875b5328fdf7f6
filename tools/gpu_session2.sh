#!/bin/bash
mkdir -p gpurun_out
NANOFED_B200_NVLS=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/check_peer.py > gpurun_out/check_peer_n2.log 2>&1; echo "check_peer rc=$?"; grep "::" gpurun_out/check_peer_n2.log; tail -3 gpurun_out/check_peer_n2.log
NANOFED_REF_BUDGET_S=240 timeout 900 python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 > gpurun_out/bench_reference_n1.json 2> gpurun_out/bench_reference_n1.err; echo "ref rc=$?"; cat gpurun_out/bench_reference_n1.json
timeout 300 python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/bench_ours_n1.json 2> gpurun_out/bench_ours_n1.err; cat gpurun_out/bench_ours_n1.json
