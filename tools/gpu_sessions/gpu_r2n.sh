#!/bin/bash
# the driver's scaling commands on an N-GPU box: bench.py at the given N values (+ the slab tool with per-rank counters)
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=8000
for N in "$@"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_x$N.json 2> gpurun_out/bench_x$N.err; echo "bench x$N rc=$?"; python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_x$N.json') if l.startswith('{')][-1])
    print({k:d[k] for k in ('value','ms_per_step','n_gpus','scaling')}, d['config']['parallelism'], d['config']['verified'], 'e2e', d['e2e']['ms_per_step'], 'frac', d['roofline']['frac'])
    print('  stats', {k:d['kernel_stats'][k] for k in ('tile_activations','cells_lowered','remote_pushes','kernel_ms','rank0_busy_ms','nvlink_bytes_read_rank0')}, d['config'].get('algorithmic_halo_bytes_rank0'))
    for k,v in d.get('extras',{}).items(): print('  extra',k,{a:v[a] for a in ('value','ms_per_step','queries_per_s','verified')})
except Exception as e: print('parse failed',e)
PY
  tail -3 gpurun_out/bench_x$N.err | cut -c1-300
done
