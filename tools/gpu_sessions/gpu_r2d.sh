#!/bin/bash
# round 2, 2-GPU session: the z-slab solver with device-side halo pulls over NVLink (IPC-mapped arenas)
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=8000
nvidia-smi --query-gpu=index,name --format=csv,noheader
N=${1:-2}
timeout 300 python -m pytest tests/test_bfs_gpu.py -q -x -k "multi_gpu or group" > gpurun_out/t_slab.log 2>&1; echo "slab tests rc=$?"; tail -15 gpurun_out/t_slab.log
for n in 768 1536; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 tools/run_bfs_slab.py --size $n --steps 3 --verify > gpurun_out/slab_${n}_x$N.json 2> gpurun_out/slab_${n}_x$N.err; echo "slab $n x$N rc=$?"; tail -1 gpurun_out/slab_${n}_x$N.json; tail -3 gpurun_out/slab_${n}_x$N.err | cut -c1-300
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_x$N.json 2> gpurun_out/bench_x$N.err; echo "bench x$N rc=$?"; cut -c1-3000 gpurun_out/bench_x$N.json; tail -5 gpurun_out/bench_x$N.err | cut -c1-400
for nt in 64 32; do echo "== NT=$nt"; PMP_B200_BFS_NT=$nt timeout 200 python tools/gpu_bfs_check.py --quick 768 1536 2>&1 | grep "c5 "; done
