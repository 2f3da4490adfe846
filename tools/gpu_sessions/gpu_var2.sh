#!/bin/bash
# A/B of experiment builds (tags of tools/build_variants.sh) at 1536^3: median / min kernel time of 8 searches + phase times
export PMP_B200_BFS_TIMEOUT_MS=8000
D=python_motion_planning_b200
R=${REPS:-9}
run() { timeout 300 python tools/gpu_bfs_check.py --quick --reps $R ${SIZES:-1536} 2>&1 | grep "c5 " | sed -e 's/certificate violations/viol/' -e 's/, reached [0-9]*//' -e 's/lowered.*us\/activation/us/' | cut -c1-260; }
echo "== product"; run
for v in "$@"; do echo "== $v"; PMP_B200_LIB=$PWD/$D/libpmp_b200_$v.so run; done
echo "== product (again)"; run
