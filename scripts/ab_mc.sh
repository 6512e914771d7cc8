#!/bin/bash
# marching-cubes A/B on the 594 MB Target volume: kernel generations and build variants
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mesh.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
run() { local name=$1; shift
  env "$@" timeout 600 python scripts/big_volume_sweeps.py --kf 600 --batch 40 2>gpurun_out/ab_mc_$name.err | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$name', d['extract_mesh'], d['extract_mesh_indexed'])"
}
run mc3 X=1
run mc2 B200TSDF_MC=2
for f in build/variants/lib_mc*.so; do [ -e "$f" ] || continue; run $(basename $f .so) B200TSDF_LIB=$PWD/$f; done
