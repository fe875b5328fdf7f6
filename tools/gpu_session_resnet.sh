#!/bin/bash
# BASELINE configs 3 and 4 at their named scale: ResNet-18 (CIFAR shape), 64 clients on 8 GPUs, bf16,
# without and with central DP fused into the aggregation kernel
mkdir -p gpurun_out
run() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port "$1" bench.py --gpus 8 --model resnet18 --clients-per-gpu 8 --samples-per-client 256 --steps 10 --warmup 3 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err; echo "$2 rc=$?"; grep '^{' gpurun_out/$2.json | cut -c1-260; }
run 29801 bench_resnet18_n8
run 29802 bench_resnet18_dp_n8 --dp
