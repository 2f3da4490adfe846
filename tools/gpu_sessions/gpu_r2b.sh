#!/bin/bash
# round 2 GPU session: phase timings of k_bfs for several register budgets, then an ncu capture
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=5000
D=python_motion_planning_b200
for v in "" _m5 _m6 _m8; do
  echo "== variant libpmp_b200$v.so"
  PMP_B200_LIB=$PWD/$D/libpmp_b200$v.so timeout 300 python tools/gpu_bfs_check.py --quick 512 768 1536 2>&1 | tail -4
done
echo "== 2 CTAs/SM (128 regs)"
PMP_B200_BFS_CTAS_PER_SM=2 timeout 300 python tools/gpu_bfs_check.py --quick 768 2>&1 | tail -2
timeout 120 python tools/prof_bfs.py 512 2 > gpurun_out/prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_bfs -s 1 -c 1 -o gpurun_out/r02_k_bfs_512 python tools/prof_bfs.py 512 2 > gpurun_out/ncu_bfs.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_bfs.log
ls -la gpurun_out/*.ncu-rep
