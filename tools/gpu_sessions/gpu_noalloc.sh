#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -x > gpurun_out/t_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/t_noalloc_alone.log
timeout 1200 python -m pytest tests -q -x -m gpu > gpurun_out/t_all.log 2>&1; echo "suite rc=$?"; tail -6 gpurun_out/t_all.log
