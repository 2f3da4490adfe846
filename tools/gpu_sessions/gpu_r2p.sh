#!/bin/bash
export PMP_B200_BFS_TIMEOUT_MS=8000
timeout 300 python -m pytest tests/test_bfs_gpu.py tests/test_zz_no_alloc.py -q -x 2>&1 | tail -2
for pw in 0 4096 16384 65536; do echo "== PACE=$pw"; PMP_B200_BFS_PACE=$pw timeout 120 python tools/gpu_slab_group.py 1536 2 8 2>&1 | tail -2; done
