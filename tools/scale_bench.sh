#!/bin/bash
# Mirror of the driver's scaling run: bench.py at N = 1, 2, 4, 8 back to back.
mkdir -p gpurun_out
for N in 1 2 4 8; do
  if [ "$N" = "1" ]; then
    timeout 300 python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  fi
  echo "N=$N rc=$?"; python -c "
import json,sys
try:
    d=json.loads([l for l in open('gpurun_out/scale_n$N.json') if l.startswith('{')][-1])
    print(d['n_gpus'], round(d['value']), round(d['ms_per_step'],3), d['config']['agg_path'], 'e2e', round(d['e2e']['value']), d['phase_ms'], d['clocks'])
except Exception as e:
    print('no json', e); print(open('gpurun_out/scale_n$N.err').read()[-800:])
"
done
