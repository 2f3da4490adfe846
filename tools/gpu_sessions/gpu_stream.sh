#!/bin/bash
# e2e of the streamed upload (tools/prof_stream.py) on experiment builds
D=python_motion_planning_b200
echo "== product"; timeout 200 python tools/prof_stream.py 1536 2>&1 | grep "streamed\|plain:" | cut -c1-110
for v in "$@"; do echo "== $v"; PMP_B200_LIB=$PWD/$D/libpmp_b200_$v.so timeout 200 python tools/prof_stream.py 1536 2>&1 | grep "streamed" | cut -c1-110; done
