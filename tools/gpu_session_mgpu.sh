#!/bin/bash
# Final multi-GPU evidence.  usage: tools/gpu_session_mgpu.sh G   (G = GPUs on this box)
G=${1:-8}
mkdir -p gpurun_out
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
if [ "$G" = "8" ]; then
  bash tools/scale_bench.sh
  timeout 300 python -m pytest tests/test_multi_gpu.py -q 2>&1 | tail -2
  run 8 29710 tools/agg_sweep.py --both --params 1199936 350000000 1000000000 --clients-per-gpu 8 2>/dev/null | grep '^{' | tee gpurun_out/agg_sweep_final_n8.log
  run 8 29711 tools/agg_sweep.py --both --params 1199936 --clients-per-gpu 64 2>/dev/null | grep '^{' | tee gpurun_out/agg_sweep_512clients_n8.log
  run 8 29712 tools/agg_sweep.py --nvls --params 350000000 --clients-per-gpu 16 2>/dev/null | grep '^{' | tee -a gpurun_out/agg_sweep_final_n8.log
else
  run "$G" 29720 tools/agg_sweep.py --both --params 1199936 350000000 1000000000 --clients-per-gpu 8 2>/dev/null | grep '^{' | tee gpurun_out/agg_sweep_final_n$G.log
  run "$G" 29721 tools/agg_sweep.py --both --params 1199936 --clients-per-gpu 64 2>/dev/null | grep '^{' | tee -a gpurun_out/agg_sweep_final_n$G.log
fi
