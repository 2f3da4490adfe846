#!/bin/bash
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=8000
N=$1; shift
run() { n=$1; shift; echo "== N=$N n=$n $*"; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 tools/run_bfs_slab.py --size $n --steps 3 --verify 2> gpurun_out/err.log | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); pr=d.pop('per_rank',[]); print({k:d[k] for k in ('ms_per_search','tile_activations','cells_lowered','remote_pushes','violations','ok')})
print('    busy', [round(r['busy_ms'],1) for r in pr], 'first', [round(r['first_ms'],1) for r in pr], 'lowered(M)', [r['lowered']//1000000 for r in pr])" || tail -5 gpurun_out/err.log; }
for pw in "$@"; do run 1536 PMP_B200_BFS_PACE=$pw; done
run 768 PMP_B200_BFS_PACE=$1
