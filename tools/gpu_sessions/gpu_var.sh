#!/bin/bash
# run tools/gpu_bfs_check.py --quick on the given experiment builds (tags of tools/build_variants.sh)
export PMP_B200_BFS_TIMEOUT_MS=8000
D=python_motion_planning_b200
echo "== product"; timeout 200 python tools/gpu_bfs_check.py --quick 768 1536 2>&1 | grep "c5 " | cut -c1-130,330-460
for v in "$@"; do echo "== $v"; PMP_B200_LIB=$PWD/$D/libpmp_b200_$v.so timeout 200 python tools/gpu_bfs_check.py --quick 768 1536 2>&1 | grep "c5 " | cut -c1-130,330-460; done
