#!/bin/bash
# round 2 GPU session: the unit-weight solver against the oracle, the suite, then bench lines
mkdir -p gpurun_out
export PMP_B200_BFS_TIMEOUT_MS=5000
timeout 240 python tools/gpu_bfs_check.py > gpurun_out/bfs_check.log 2>&1; echo "bfs_check rc=$?"; tail -14 gpurun_out/bfs_check.log
timeout 600 python -m pytest tests/test_bfs_gpu.py -q -x --durations=5 > gpurun_out/t_bfs.log 2>&1; echo "bfs tests rc=$?"; tail -8 gpurun_out/t_bfs.log
timeout 900 python -m pytest tests -q -m gpu --durations=5 > gpurun_out/t_all.log 2>&1; echo "suite rc=$?"; tail -12 gpurun_out/t_all.log
timeout 200 python bench.py --n 256 --steps 3 --no-cpu > gpurun_out/bench_c5_256.json 2> gpurun_out/bench_c5_256.err; echo "c5 256 rc=$?"; cut -c1-2500 gpurun_out/bench_c5_256.json; tail -5 gpurun_out/bench_c5_256.err
timeout 300 python bench.py --n 768 --steps 5 --no-cpu --no-extras > gpurun_out/bench_c5_768.json 2> gpurun_out/bench_c5_768.err; echo "c5 768 rc=$?"; cut -c1-2000 gpurun_out/bench_c5_768.json; tail -3 gpurun_out/bench_c5_768.err
timeout 600 python bench.py --steps 3 --no-cpu --no-extras > gpurun_out/bench_c5_1536.json 2> gpurun_out/bench_c5_1536.err; echo "c5 1536 rc=$?"; cut -c1-2000 gpurun_out/bench_c5_1536.json; tail -3 gpurun_out/bench_c5_1536.err
