set -u
OUT=gpurun_out; mkdir -p $OUT
KISS_B200_DEBUG_WAVES=1 python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > $OUT/r1_c3j_bench.json 2> $OUT/r1_c3j_bench.err
grep "tail bucket" $OUT/r1_c3j_bench.err
