#!/bin/bash
# bench at N = 1, 2, 4, 8 (or the list given as \$1) on one box (what the driver does at round end)
mkdir -p gpurun_out
for n in ${1:-1 2 4 8}; do
  if [ $n -eq 1 ]; then
    timeout 900 python bench.py --gpus 1 --steps 6 --warmup 2 --no-cpu-baseline > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 6 --warmup 2 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  echo "N=$n rc=$?"; python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/scale_n$n.json')); print(d['n_gpus'], 'ms', round(d['ms_per_step'],4), 'value %.3e'%d['value'], 'e2e %.3e'%d['e2e']['value'], d['volume'], d['clocks'])
except Exception as e: print('parse failed', e)"
done
