set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "prefix_doubling" 2>&1 | tail -60 > $OUT/r1_pd_test.txt
tail -40 $OUT/r1_pd_test.txt
