#!/bin/bash
mkdir -p gpurun_out
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
