#!/bin/bash
# round 2 GPU session: phase-locked k_bfs vs free-running, launch shapes; ncu instruction-cache metrics
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=5000
timeout 240 python tools/gpu_bfs_check.py 256 > gpurun_out/bfs_check.log 2>&1; echo "bfs_check rc=$?"; tail -12 gpurun_out/bfs_check.log
timeout 300 python -m pytest tests/test_bfs_gpu.py -q -x > gpurun_out/t_bfs.log 2>&1; echo "bfs tests rc=$?"; tail -3 gpurun_out/t_bfs.log
run() { echo "== $*"; env "$@" timeout 300 python tools/gpu_bfs_check.py --quick 768 1536 2>&1 | grep "c5 "; }
run PMP_B200_BFS_NT=512
run PMP_B200_BFS_NT=256
run PMP_B200_BFS_NT=128
run PMP_B200_BFS_NT=1024
run PMP_B200_BFS_FREE=1
run PMP_B200_BFS_FREE=1 PMP_B200_BFS_CTAS_PER_SM=2
M=smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,sm__icc_request_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum,smsp__inst_executed.sum,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio,smsp__average_warps_issue_stalled_membar_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
timeout 120 python tools/prof_bfs.py 768 2 > gpurun_out/prof_plain.log 2>&1 &&
timeout 600 ncu --metrics $M --clock-control none -k regex:k_bfs -s 1 -c 1 --csv --log-file gpurun_out/r02_k_bfs_sync_768_metrics.csv python tools/prof_bfs.py 768 2 > gpurun_out/ncu_bfs.log 2>&1; echo "ncu rc=$?"; cat gpurun_out/r02_k_bfs_sync_768_metrics.csv | tail -14
