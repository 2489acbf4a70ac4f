set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $OUT/r1_c3r_pytest.txt
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/r1_c3r_bench.json 2> $OUT/r1_c3r_bench.err
tail -c 300 $OUT/r1_c3r_bench.json
