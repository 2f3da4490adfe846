#!/bin/bash
export PMP_B200_BFS_TIMEOUT_MS=8000
D=python_motion_planning_b200
for v in _m5 _m6 _m8; do
  echo "== variant libpmp_b200$v.so"
  PMP_B200_LIB=$PWD/$D/libpmp_b200$v.so timeout 300 python tools/gpu_bfs_check.py --quick 768 1536 2>&1 | grep "c5 " | cut -c1-420
done
