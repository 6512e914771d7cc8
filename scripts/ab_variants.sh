#!/bin/bash
# A/B of kernel build variants (build/variants/lib_*.so) on the bench workload.
for f in build/variants/lib_*.so; do
  n=$(basename $f .so)
  B200TSDF_LIB=$PWD/$f timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$n', 'ms_per_step', round(d['ms_per_step'],4), 'value', '%.3e'%d['value'], 'e2e_ms', round(d['e2e']['ms_per_step'],3))"
done
