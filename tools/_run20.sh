set -u
OUT=gpurun_out; mkdir -p $OUT
python bench.py > $OUT/r1_c3q_bench.json 2> $OUT/r1_c3q_bench.err
tail -c 300 $OUT/r1_c3q_bench.json
python bench.py --workload c2 > $OUT/r1_c3q_bench_c2.json 2> $OUT/r1_c3q_bench_c2.err
