#!/bin/bash
# A/B of kernel build variants on the bench workload (one gpurun call).
#   build/variants/lib_*.so   : python -c "from mast3r_slam_quality_dualtsdf_b200 import build; build.build_variant('minb3', ['FUSE_MINB=3'])"
#   env knobs                 : B200TSDF_NO_FAST=1 (generic march only)
mkdir -p gpurun_out
run() { # name, env...
  local name=$1; shift
  env "$@" timeout 600 python bench.py --steps 6 --warmup 2 --launches-per-step 5 --passes-per-step 1 --no-cpu-baseline --no-sections 2>gpurun_out/ab_$name.err | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('$name', 'launch_ms', round(d['roofline']['launch_ms'],4), 'value %.4e'%d['value'], 'e2e_ms', round(d['e2e']['ms_per_step'],3), d['clocks']['sm_mhz'])
except Exception as e: print('$name failed', e)"
}
run default X=1
run nofast B200TSDF_NO_FAST=1
for o in ${AB_ORDERS:-2 4 8 20 40}; do run order$o B200TSDF_KF_ORDER=$o; done
for f in build/variants/lib_*.so; do
  [ -e "$f" ] && [ "${AB_LIBS:-1}" = "1" ] || continue
  n=$(basename $f .so)
  run $n B200TSDF_LIB=$PWD/$f
done
# marching cubes A/B on the 594 MB volume (bit-plane kernel vs per-cube kernel)
if [ "${AB_MC:-1}" = "1" ]; then
  timeout 600 python scripts/big_volume_sweeps.py --kf 600 --batch 40 2>gpurun_out/ab_mc2.err | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('mc_kernel2', d['extract_mesh'], d['extract_mesh_indexed'], d['stats'])"
  B200TSDF_MC_V1=1 timeout 600 python scripts/big_volume_sweeps.py --kf 600 --batch 40 2>gpurun_out/ab_mc1.err | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('mc_kernel ', d['extract_mesh'])"
fi
