set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee $OUT/r1_c3n_pytest.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wave_fused -s 8 -c 3 -f -o $OUT/r1_c3n_waves \
   python bench.py --steps 1 --warmup 0 --no-e2e --no-cpu-baseline > $OUT/r1_c3n_ncu_waves.log 2>&1
tail -2 $OUT/r1_c3n_ncu_waves.log
