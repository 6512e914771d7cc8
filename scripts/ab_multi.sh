#!/bin/bash
# A/B of the multi-GPU fuse at N ranks: scripts/ab_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { local name=$1; shift
  env "$@" timeout 300 $TR --master-port 295$((RANDOM % 90 + 10)) bench.py --gpus $N --steps 4 --warmup 2 --no-sections --launches-per-step 5 --passes-per-step 1 ${AB_ARGS} 2>gpurun_out/abm_$name.err | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('$name', 'launch_ms', round(d['roofline']['launch_ms'],4), 'value %.4e'%d['value'], 'e2e_ms', round(d['e2e']['ms_per_step'],3), d['volume'])
except Exception as e: print('$name failed', e)"
}
run pull X=1
AB_ARGS="--no-pipeline" run pull_serial X=1
AB_ARGS="--ingest gather" run gather X=1
run pull_nofast B200TSDF_NO_FAST=1
for f in build/variants/lib_*.so; do [ -e "$f" ] || continue; n=$(basename $f .so); run pull_$n B200TSDF_LIB=$PWD/$f; done
