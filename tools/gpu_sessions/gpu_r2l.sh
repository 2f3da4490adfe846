#!/bin/bash
# the driver's commands: default bench (N=1) + reference arm, then the ncu launch list of the same command
mkdir -p gpurun_out
( time timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2>&1 | grep real; echo "bench rc=$?"; cut -c1-6000 gpurun_out/bench_default.json; tail -5 gpurun_out/bench_default.err
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err ) 2>&1 | grep real; cut -c1-1500 gpurun_out/bench_reference.json; tail -3 gpurun_out/bench_reference.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-extras > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_c5.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extras > gpurun_out/ncu_l.log 2>&1; echo "ncu rc=$?"; wc -l gpurun_out/r02_launches_c5.csv
