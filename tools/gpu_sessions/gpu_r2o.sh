#!/bin/bash
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=8000
N=$1
for n in 1536; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 tools/run_bfs_slab.py --size $n --steps 3 --verify > gpurun_out/slab_${n}_x$N.json 2> gpurun_out/slab_${n}_x$N.err; echo "slab $n x$N rc=$?"; tail -1 gpurun_out/slab_${n}_x$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); pr=d.pop('per_rank',[]); print(d)
for r in pr: print('   ',r)"
done
