#!/bin/bash
# round 2: re-validated kernel; CTAs/SM sweep; final ncu captures of k_bfs
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=8000
timeout 240 python tools/gpu_bfs_check.py 256 768 1536 > gpurun_out/bfs_check.log 2>&1; echo "bfs_check rc=$?"; tail -5 gpurun_out/bfs_check.log
timeout 300 python -m pytest tests/test_bfs_gpu.py tests/test_zz_no_alloc.py -q -x > gpurun_out/t_bfs.log 2>&1; echo "bfs tests rc=$?"; tail -3 gpurun_out/t_bfs.log
for c in 1 2 3; do echo "== CTAS_PER_SM=$c"; PMP_B200_BFS_CTAS_PER_SM=$c timeout 200 python tools/gpu_bfs_check.py --quick 256 768 1536 2>&1 | grep "c5 " | cut -c1-400; done
timeout 120 python tools/prof_bfs.py 768 2 > gpurun_out/prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_bfs -s 1 -c 1 -o gpurun_out/r02_k_bfs_768 python tools/prof_bfs.py 768 2 > gpurun_out/ncu_bfs.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_bfs.log
