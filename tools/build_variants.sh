#!/bin/bash
# experiment builds of the unit-weight solver with other register budgets: libpmp_b200_m<k>.so
# (PMP_B200_LIB=<path> selects one at run time).  Not part of the product build.
set -e
cd "$(dirname "$0")/../python_motion_planning_b200/csrc"
make -s > /dev/null
for k in "$@"; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall -Xptxas -v -DBFS_MINB=$k -DBFS_SYNC_MINB=$k -c pmp_bfs.cu -o /tmp/pmp_bfs_m$k.o 2> /tmp/pmp_bfs_m$k.log
  grep -A2 "k_bfs7" /tmp/pmp_bfs_m$k.log | tail -2
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libpmp_b200_m$k.so pmp_grid.o pmp_sssp.o /tmp/pmp_bfs_m$k.o pmp_path.o pmp_los.o pmp_bestfirst.o
done
