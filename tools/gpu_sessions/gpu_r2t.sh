#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -x -m gpu > gpurun_out/t_all.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/t_all.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.json').read())
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, 'e2e', d['e2e'], 'frac', d['roofline']['frac'], 'traffic', d['roofline']['traffic'])
for k,v in d['extras'].items(): print(' ', k, {a:v[a] for a in ('value','ms_per_step','e2e_ms_per_step','verified')})
print(' cpu', d['cpu_baseline']['value'], d['cpu_baseline_reference_python'].get('c1',{}).get('ms'))
PY
tail -3 gpurun_out/bench_default.err
