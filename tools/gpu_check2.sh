#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_incr_gpu.py tests/test_examples.py tests/test_replan.py tests/test_jps.py -x -q -m gpu > gpurun_out/t_incr.log 2>&1; echo "incr rc=$?"; tail -15 gpurun_out/t_incr.log
timeout 120 python tools/summarize_benchmark.py --order reference --iterations 20 --repeats 2 > gpurun_out/bench_examples_reference.json 2> gpurun_out/bench_examples_reference.err; echo "examples rc=$?"
cat gpurun_out/bench_examples_reference.json; tail -3 gpurun_out/bench_examples_reference.err
timeout 100 python bench.py --no-cpu --workload c3 --queries 1024 --steps 3 --warmup 3 > gpurun_out/c3_1024_newsub.json 2>/dev/null
python - <<PY
import json
d=json.load(open("gpurun_out/c3_1024_newsub.json"))
k=d["kernel_stats"]
print("c3 1024q default ms/step", round(d["ms_per_step"],3), "iters", k["sub_iterations"], "check", d["config"].get("check_value"))
PY
