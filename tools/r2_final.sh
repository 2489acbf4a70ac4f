#!/bin/bash
# Round-2 evidence on one B200 (run through gpurun): bench lines (C3 with e2e / cpu baseline, C2, reference arm with its
# full-size run), ncu launch list of one C3 step (time + DRAM bytes per kernel -> profiles/ncu_dram_per_launch.json),
# (the `ncu --set full` captures: tools/r2_ncu.sh)
set -u
TAG=${1:-r2_final}
OUT=gpurun_out
mkdir -p $OUT
python bench.py > $OUT/${TAG}_bench_c3.json 2> $OUT/${TAG}_bench_c3.err
tail -c 300 $OUT/${TAG}_bench_c3.json
python bench.py --workload c2 > $OUT/${TAG}_bench_c2.json 2> $OUT/${TAG}_bench_c2.err
python bench.py --impl reference --steps 5 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err
# launch list (the plain run of the same command line exits 0 first)
python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $OUT/${TAG}_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file $OUT/${TAG}_launches.csv python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1
python tools/ncu_dram_per_kernel.py $OUT/${TAG}_launches.csv $OUT/ncu_dram_per_launch.json c3 > $OUT/${TAG}_launches.txt 2>&1
head -12 $OUT/${TAG}_launches.txt
ls -la $OUT/${TAG}_*
