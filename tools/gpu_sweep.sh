#!/bin/bash
# tuning sweep on the GPU box: in-tile sub-bucket width (PMP_B200_SUB) for c4 / c3; new JPS tests
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_bestfirst_gpu.py tests/test_jps.py -x -q -m gpu -k "jps or reference_order_equals" > gpurun_out/t_jps.log 2>&1; echo "jps rc=$?"; tail -3 gpurun_out/t_jps.log
for sub in 1 2 3 4 6; do
  PMP_B200_SUB=$sub timeout 100 python bench.py --no-cpu --steps 8 --warmup 3 > gpurun_out/sweep_c4_sub$sub.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/sweep_c4_sub$sub.json"))
print("c4 sub=$sub ms/step", round(d["ms_per_step"],3), "search", round(d["phases"]["search_ms"],3), "act", d["kernel_stats"]["tile_activations"], "exp", d["kernel_stats"]["cells_expanded"], "iters", d["kernel_stats"]["sub_iterations"], "check", d["config"]["check_value"])
PY
done
for sub in 1 2 4; do
  PMP_B200_SUB=$sub timeout 100 python bench.py --no-cpu --workload c3 --queries 1024 --steps 3 --warmup 3 > gpurun_out/sweep_c3_sub$sub.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/sweep_c3_sub$sub.json"))
print("c3 sub=$sub ms/step", round(d["ms_per_step"],3), "act", d["kernel_stats"]["tile_activations"], "exp", d["kernel_stats"]["cells_expanded"], "check", d["config"].get("check_value"))
PY
done
for sub in 6 8 12 16; do
  PMP_B200_SUB=$sub timeout 100 python bench.py --no-cpu --workload c3 --queries 1024 --steps 3 --warmup 3 > gpurun_out/sweep_c3_sub$sub.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/sweep_c3_sub$sub.json"))
k=d["kernel_stats"]
print("c3 sub=$sub ms/step", round(d["ms_per_step"],3), "act", k["tile_activations"], "exp", k["cells_expanded"], "iters", k["sub_iterations"], "check", d["config"].get("check_value"))
PY
done
for w in 48 192; do
  PMP_B200_SUB=8 PMP_B200_WINDOW=$w timeout 100 python bench.py --no-cpu --workload c3 --queries 1024 --steps 3 --warmup 3 > gpurun_out/sweep_c3_sub8_w$w.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/sweep_c3_sub8_w$w.json"))
k=d["kernel_stats"]
print("c3 sub=8 window=$w ms/step", round(d["ms_per_step"],3), "act", k["tile_activations"], "exp", k["cells_expanded"], "iters", k["sub_iterations"])
PY
done
