#!/bin/bash
# slab solver at N GPUs with fewer resident CTAs per SM (lower hop latency, less throughput)
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=8000
N=$1; shift
for c in "$@"; do echo "== N=$N CTAS_PER_SM=$c"; PMP_B200_BFS_CTAS_PER_SM=$c timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 tools/run_bfs_slab.py --size 1536 --steps 3 --verify 2> gpurun_out/err.log | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); pr=d.pop('per_rank',[]); print({k:d[k] for k in ('ms_per_search','tile_activations','cells_lowered','violations','ok')})
print('    busy', [round(r['busy_ms'],1) for r in pr])"; done
