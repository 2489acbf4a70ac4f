set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $OUT/r1_c3b_pytest.txt
KISS_B200_DEBUG_GROUPS=1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/r1_c3b_bench.json 2> $OUT/r1_c3b_bench.err
tail -c 300 $OUT/r1_c3b_bench.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_groups_medium -s 5 -c 1 -f -o $OUT/r1_c3b_medium2048 \
   python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline > $OUT/r1_c3b_ncu_medium.log 2>&1
tail -3 $OUT/r1_c3b_ncu_medium.log
ls -la $OUT/r1_c3b_*
