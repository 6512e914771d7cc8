#!/bin/bash
# usage: scripts/multi_rank_checks.sh N  -- the real-rank checks + the bench line at N ranks
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for s in sharded_pull_check sharded_mesh_check sharded_query_check sharded_reintegrate_check; do
  timeout 300 $TR --master-port 296$((RANDOM % 90 + 10)) scripts/$s.py > gpurun_out/${s}_n$N.log 2>&1
  echo "$s N=$N rc=$?"; tail -2 gpurun_out/${s}_n$N.log | cut -c1-600
done
timeout 600 $TR --master-port 29577 bench.py --gpus $N --steps 4 --warmup 2 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench N=$N rc=$?"
python - <<P
import json
try:
    d=json.load(open('gpurun_out/bench_n$N.json'))
    print('N', d['n_gpus'], 'ms/step', round(d['ms_per_step'],3), 'launch_ms', round(d['roofline']['launch_ms'],4), 'value %.4e'%d['value'], 'e2e ms', round(d['e2e']['ms_per_step'],3))
    for k in ('extract','sweep','reintegrate'):
        print(k, json.dumps(d.get(k))[:500])
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/bench_n$N.err').read()[-3000:])
P
