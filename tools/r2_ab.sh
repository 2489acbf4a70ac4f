#!/bin/bash
# round 2: GPU test suite, then C3 bench lines under the A/B switches given as arguments ("NAME=VAL,NAME2=VAL2" each,
# "default" = none), e.g.  bash tools/r2_ab.sh r2_c default KISS_B200_NO_BUCKETS=1
set -u
TAG=${1:-r2_a}
shift
OUT=gpurun_out
mkdir -p $OUT
if [ "${SKIP_TESTS:-0}" != "1" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee $OUT/${TAG}_pytest.txt
fi
for v in "$@"; do
  name=$(echo "$v" | tr ',=' '__')
  envs=""
  if [ "$v" != "default" ]; then envs=$(echo "$v" | tr ',' ' '); fi
  if [ -n "${DEBUG_GROUPS:-}" ]; then envs="$envs KISS_B200_DEBUG_GROUPS=1"; fi
  env $envs python bench.py --steps ${STEPS:-5} --warmup 3 --no-e2e --no-cpu-baseline ${BENCH_ARGS:-} > $OUT/${TAG}_${name}.json 2> $OUT/${TAG}_${name}.err
  python tools/bench_show.py $OUT/${TAG}_${name}.json | grep -v "wave\|pack\|k_lms\|short"
done
