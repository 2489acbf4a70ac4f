set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $OUT/r1_c3d_pytest.txt
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/r1_c3d_bench.json 2> $OUT/r1_c3d_bench.err
tail -c 300 $OUT/r1_c3d_bench.json
KISS_B200_L2_FETCH=32 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/r1_c3d_bench_l2f32.json 2> $OUT/r1_c3d_bench_l2f32.err
tail -c 300 $OUT/r1_c3d_bench_l2f32.json; grep L2Fetch $OUT/r1_c3d_bench_l2f32.err
KISS_B200_L2_FETCH=128 python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/r1_c3d_bench_c2_l2f128.json 2> $OUT/r1_c3d_bench_c2_l2f128.err
KISS_B200_L2_FETCH=32 python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/r1_c3d_bench_c2_l2f32.json 2> $OUT/r1_c3d_bench_c2_l2f32.err
python bench.py --workload c2 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/r1_c3d_bench_c2.json 2> $OUT/r1_c3d_bench_c2.err
