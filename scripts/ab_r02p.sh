#!/bin/bash
# round-2 A/B: CTA order of the fuse launch, marching-cubes prefetch depth, DRAM traffic per order
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mesh.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -3
AB_MC=0 AB_LIBS=0 bash scripts/ab_variants.sh 2>&1
for lib in default build/variants/lib_mc2s.so build/variants/lib_mc4s.so; do
  if [ $lib = default ]; then unset B200TSDF_LIB; else export B200TSDF_LIB=$PWD/$lib; fi
  timeout 600 python scripts/big_volume_sweeps.py --kf 600 --batch 40 2>gpurun_out/ab_mc_$(basename $lib .so).err | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('mc $lib', d['extract_mesh'], d['extract_mesh_indexed'])"
done
unset B200TSDF_LIB
for o in 1 8 40; do
  B200TSDF_KF_ORDER=$o timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sectors.sum,smsp__inst_executed.sum,lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:integrate_keyframes_kernel -s 12 -c 3 --csv --log-file gpurun_out/ncu_order$o.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-sections > gpurun_out/ncu_order$o.log 2>&1
  echo "ncu order $o rc=$?"
  python - <<P
import csv
rows=[r for r in csv.reader(open('gpurun_out/ncu_order$o.csv')) if len(r)>10]
hdr=rows[0]; mi=hdr.index('Metric Name'); vi=hdr.index('Metric Value'); ii=hdr.index('ID')
agg={}
for r in rows[1:]:
    agg.setdefault(r[mi],[]).append(float(r[vi].replace(',','')))
for k,v in agg.items(): print('  order $o', k, sum(v)/len(v))
P
done
