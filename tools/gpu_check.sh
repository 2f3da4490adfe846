#!/bin/bash
# One GPU-box session (gpurun -- 'bash tools/gpu_check.sh'): smoke, the whole -m gpu suite with
# durations, the drop-in benchmark summary in reference order and the default bench line.
# Everything lands in gpurun_out/.
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python -m pytest tests -q -m gpu --durations=10 > gpurun_out/t_all.log 2>&1; echo "suite rc=$?"
tail -16 gpurun_out/t_all.log
timeout 120 python tools/summarize_benchmark.py --order reference --iterations 20 --repeats 2 > gpurun_out/bench_examples_reference.json 2> gpurun_out/bench_examples_reference.err; echo "examples rc=$?"
cat gpurun_out/bench_examples_reference.json
timeout 200 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/bench_c4.json
