#!/bin/bash
# Evidence at the headline workload (C3: chm13-sized 3.1 Gbp with repeats) on one B200, run through gpurun:
# GPU tests, bench lines for c3 and c2, tie-group histograms, ncu launch list (time + DRAM bytes) of one c3 step.
set -u
TAG=${1:-r2_c3}
OUT=gpurun_out
mkdir -p $OUT
if [ "${SKIP_TESTS:-0}" != "1" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $OUT/${TAG}_pytest.txt
fi
python bench.py --steps ${STEPS:-5} --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
tail -c 600 $OUT/${TAG}_bench.json
if [ "${SKIP_C2:-0}" != "1" ]; then
  python bench.py --workload c2 > $OUT/${TAG}_bench_c2.json 2> $OUT/${TAG}_bench_c2.err
  tail -c 300 $OUT/${TAG}_bench_c2.json
fi
if [ "${SKIP_NCU:-0}" != "1" ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
      --log-file $OUT/${TAG}_launches.csv python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1
  # (also regenerates the per-launch DRAM table bench.py's roofline.traffic reads; copy it to profiles/ with the list)
  python tools/ncu_dram_per_kernel.py $OUT/${TAG}_launches.csv $OUT/ncu_dram_per_launch.json c3 > $OUT/${TAG}_launches.txt 2>&1
  head -30 $OUT/${TAG}_launches.txt
fi
ls -la $OUT/${TAG}_*
if [ "${SKIP_REF:-0}" != "1" ]; then
  python bench.py --impl reference --steps 5 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err
  tail -c 400 $OUT/${TAG}_bench_reference.json
fi
