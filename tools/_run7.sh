set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_groups_medium -s 5 -c 1 -f -o $OUT/r1_c3f_medium2048 \
   python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline > $OUT/r1_c3f_ncu_medium.log 2>&1
tail -2 $OUT/r1_c3f_ncu_medium.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_groups_tiny -s 0 -c 1 -f -o $OUT/r1_c3f_tiny \
   python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline > $OUT/r1_c3f_ncu_tiny.log 2>&1
tail -2 $OUT/r1_c3f_ncu_tiny.log
