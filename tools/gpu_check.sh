#!/bin/bash
# One GPU-box session: new reference-order tests first, then the whole -m gpu suite, the default
# bench line and the drop-in benchmark summary.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 300 python -m pytest tests/test_bestfirst_gpu.py tests/test_examples.py -x -q -m gpu > gpurun_out/t_bestfirst.log 2>&1; echo "bestfirst rc=$?"
tail -5 gpurun_out/t_bestfirst.log
timeout 420 python -m pytest tests -q -m gpu --deselect tests/test_bestfirst_gpu.py --deselect tests/test_examples.py --durations=8 > gpurun_out/t_all.log 2>&1; echo "suite rc=$?"
tail -14 gpurun_out/t_all.log
timeout 120 python tools/summarize_benchmark.py --order reference --iterations 20 --repeats 2 > gpurun_out/bench_examples_reference.json 2> gpurun_out/bench_examples_reference.err; echo "examples rc=$?"
cat gpurun_out/bench_examples_reference.json
timeout 200 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench rc=$?"
cut -c1-600 gpurun_out/bench_c4.json
