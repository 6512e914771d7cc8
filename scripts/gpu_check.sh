#!/bin/bash
# One gpurun call: GPU tests, the N = 1 bench line, the ncu launch list and one
# ncu --set full capture of the fuse kernel.  Outputs under gpurun_out/.
#   scripts/gpu_check.sh [tag] [pytest-args...]
tag=${1:-run}; shift
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q "$@" > gpurun_out/${tag}_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.txt
timeout 900 python bench.py --steps 6 --warmup 2 > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err; echo "bench rc=$?"
python - <<P
import json
try:
    d=json.load(open('gpurun_out/${tag}_bench_n1.json'))
    print('ms/step', d['ms_per_step'], 'launch_ms', d['roofline']['launch_ms'], 'value %.4e'%d['value'], 'kf/s', d['keyframes_per_s'], 'e2e ms', d['e2e']['ms_per_step'], 'frac', d['roofline']['frac'], d['clocks'])
    for k in ('extract','sweep','reintegrate','config2','manager'):
        print(k, json.dumps(d.get(k))[:600])
except Exception as e: print('parse failed', e); print(open('gpurun_out/${tag}_bench_n1.err').read()[-3000:])
P
PROF="python bench.py --steps 1 --warmup 1 --passes-per-step 1 --no-cpu-baseline --no-sections"
$PROF > gpurun_out/${tag}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${tag}_launches.csv $PROF > gpurun_out/${tag}_ncu1.log 2>&1
echo "ncu launches rc=$?"
$PROF > gpurun_out/${tag}_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:integrate_keyframes -s 16 -c 1 -o gpurun_out/${tag}_prof -f $PROF > gpurun_out/${tag}_ncu2.log 2>&1
echo "ncu full rc=$?"
python scripts/ncu_summary.py gpurun_out/${tag}_prof.ncu-rep 40 > gpurun_out/${tag}_integrate_ncu_full.txt 2>&1
# marching cubes on the 594 MB Target volume: plain run, then the launch list and one full capture
MC="python scripts/big_volume_sweeps.py --kf 600 --batch 40"
$MC > gpurun_out/${tag}_big_volume.json 2> gpurun_out/${tag}_big_volume.err && \
ncu --set full --clock-control none --import-source on -k regex:mc_kernel3 -s 2 -c 1 -o gpurun_out/${tag}_mc3 -f $MC > gpurun_out/${tag}_ncu3.log 2>&1
echo "ncu mc rc=$?"; cat gpurun_out/${tag}_big_volume.json | cut -c1-900
python scripts/ncu_summary.py gpurun_out/${tag}_mc3.ncu-rep 40 > gpurun_out/${tag}_mc3_ncu_full.txt 2>&1
