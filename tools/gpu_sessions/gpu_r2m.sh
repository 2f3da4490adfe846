#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/prof_bfs.py 1536 2 > gpurun_out/prof_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_bfs -s 1 -c 1 -o gpurun_out/r02_k_bfs_1536 python tools/prof_bfs.py 1536 2 > gpurun_out/ncu_bfs.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_bfs.log
