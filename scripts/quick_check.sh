#!/bin/bash
# One short gpurun call: GPU tests, smoke, the default N = 1 bench line.
#   scripts/quick_check.sh [tag]
tag=${1:-run}
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q -rf > gpurun_out/${tag}_pytest.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${tag}_pytest.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/${tag}_smoke.txt
timeout 600 python bench.py > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench_n1.err; echo "bench rc=$?"
python - <<P
import json
try:
    d=json.load(open('gpurun_out/${tag}_bench_n1.json'))
    print('ms/step', d['ms_per_step'], 'launch_ms', d['roofline']['launch_ms'], 'value %.4e'%d['value'], 'kf/s', d['keyframes_per_s'], 'e2e ms', d['e2e']['ms_per_step'], 'frac', d['roofline']['frac'], d['clocks'])
    for k in ('extract','sweep','reintegrate','config2','manager'):
        print(k, json.dumps(d.get(k))[:700])
except Exception as e: print('parse failed', e); print(open('gpurun_out/${tag}_bench_n1.err').read()[-3000:])
P
