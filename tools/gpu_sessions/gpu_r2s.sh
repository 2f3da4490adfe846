#!/bin/bash
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=8000
timeout 600 python -m pytest tests/test_bfs_gpu.py tests/test_zz_no_alloc.py -q -x > gpurun_out/t_bfs.log 2>&1; echo "bfs tests rc=$?"; tail -5 gpurun_out/t_bfs.log
timeout 600 python bench.py --steps 5 --no-cpu --no-extras > gpurun_out/bench_stream.json 2> gpurun_out/bench_stream.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_stream.json').read())
print({k:d[k] for k in ('value','ms_per_step')}, 'e2e', d['e2e'], d['config']['verified'])
PY
tail -3 gpurun_out/bench_stream.err
PMP_B200_NO_STREAMED_UPLOAD=1 timeout 600 python bench.py --steps 5 --no-cpu --no-extras 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('no streaming: ms', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'])"
