set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee $OUT/r1_c3o_pytest.txt
timeout 600 python tools/validate_large.py c3h --profile > $OUT/r1_c3h_validate2.jsonl 2> $OUT/r1_c3h_validate2.err
cut -c1-400 $OUT/r1_c3h_validate2.jsonl
KISS_B200_DEBUG_WAVES=1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/r1_c3o_bench.json 2> $OUT/r1_c3o_bench.err
tail -c 200 $OUT/r1_c3o_bench.json
