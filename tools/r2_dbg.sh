#!/bin/bash
# debug statistics of one C3 step (tie-group rounds)
set -u
TAG=${1:-r2_dbg}
OUT=gpurun_out
mkdir -p $OUT
KISS_B200_DEBUG_GROUPS=1 python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline > $OUT/${TAG}_v.json 2> $OUT/${TAG}_v.err
KISS_B200_REAL_PIVOT=1 KISS_B200_DEBUG_GROUPS=1 python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline > $OUT/${TAG}_r.json 2> $OUT/${TAG}_r.err
grep "medium class" $OUT/${TAG}_v.err | head -8
