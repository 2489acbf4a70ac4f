set -u
OUT=gpurun_out; mkdir -p $OUT
for rs in 2 8 16 64; do
KISS_B200_RANK_SCALE=$rs python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e > $OUT/r1_c3h_bench_rs$rs.json 2> $OUT/r1_c3h_bench_rs$rs.err
tail -c 200 $OUT/r1_c3h_bench_rs$rs.json
done
