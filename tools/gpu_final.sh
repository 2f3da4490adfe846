#!/bin/bash
# final measurements of the round: bench lines of every workload + ncu launch list of the default bench
mkdir -p gpurun_out
timeout 200 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "c4 rc=$?"; cut -c1-200 gpurun_out/bench_c4.json
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c4_reference.json 2> gpurun_out/bench_c4_reference.err; echo "c4 ref rc=$?"; cut -c1-300 gpurun_out/bench_c4_reference.json
timeout 200 python bench.py --workload c3 --steps 5 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "c3 rc=$?"; cut -c1-200 gpurun_out/bench_c3.json
timeout 200 python bench.py --workload c5 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"; cut -c1-200 gpurun_out/bench_c5.json
timeout 100 python bench.py --workload c2 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?"; cut -c1-200 gpurun_out/bench_c2.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c4.csv python bench.py --no-cpu --steps 2 --warmup 1 > gpurun_out/ncu_c4.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/launches_c4.csv
