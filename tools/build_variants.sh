#!/bin/bash
# experiment builds of the unit-weight solver: libpmp_b200_<tag>.so for each "tag:-Dflags" argument
# (PMP_B200_LIB=<path> selects one at run time).  Not part of the product build.
set -e
cd "$(dirname "$0")/../python_motion_planning_b200/csrc"
make -s > /dev/null
for spec in "$@"; do
  tag=${spec%%:*}; flags=${spec#*:}
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall -Xptxas -v $flags -c pmp_bfs.cu -o /tmp/pmp_bfs_$tag.o 2> /tmp/pmp_bfs_$tag.log
  grep -A2 "k_bfs7" /tmp/pmp_bfs_$tag.log | tail -1
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libpmp_b200_$tag.so pmp_grid.o pmp_sssp.o /tmp/pmp_bfs_$tag.o pmp_path.o pmp_los.o pmp_bestfirst.o
done
