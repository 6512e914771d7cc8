#!/bin/bash
# A/B of the launch size (keyframes per launch) at N ranks: scripts/ab_launch_size.sh N "40 120 200"
N=${1:-2}; shift
mkdir -p gpurun_out
for kf in ${1:-40 120}; do
  L=$((600 / kf))
  timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-sections --kf-per-launch $kf --launches-per-step $L \
    2> gpurun_out/ab_launch_${N}_${kf}.err | grep '^{' | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('N=$N kf/launch=$kf', 'ms/step', round(d['ms_per_step'],3), 'value %.4e'%d['value'], 'kf/s', round(d['keyframes_per_s']), 'e2e ms', round(d['e2e']['ms_per_step'],3), 'e2e %.4e'%d['e2e']['value'])"
done
