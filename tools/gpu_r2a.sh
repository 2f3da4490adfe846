#!/bin/bash
# round 2, first GPU session: the new unit-weight solver against the oracle, then the suite, then c5 bench lines
mkdir -p gpurun_out
timeout 300 python tools/gpu_bfs_check.py > gpurun_out/bfs_check.log 2>&1; echo "bfs_check rc=$?"; tail -20 gpurun_out/bfs_check.log
timeout 900 python -m pytest tests -q -m gpu -x --durations=5 > gpurun_out/t_all.log 2>&1; echo "suite rc=$?"; tail -12 gpurun_out/t_all.log
timeout 300 python bench.py --workload c5 --n 768 --steps 5 --no-cpu > gpurun_out/bench_c5_768.json 2> gpurun_out/bench_c5_768.err; echo "c5 768 rc=$?"; cut -c1-1500 gpurun_out/bench_c5_768.json; tail -3 gpurun_out/bench_c5_768.err
timeout 600 python bench.py --workload c5 --n 1536 --steps 3 --no-cpu > gpurun_out/bench_c5_1536.json 2> gpurun_out/bench_c5_1536.err; echo "c5 1536 rc=$?"; cut -c1-1500 gpurun_out/bench_c5_1536.json; tail -3 gpurun_out/bench_c5_1536.err
