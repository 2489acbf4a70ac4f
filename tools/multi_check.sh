#!/bin/bash
# Multi-GPU evidence on an N-GPU box (gpurun --gpus N): parity of the sharded path (tests/test_gpu_multi.py) and the
# bench line at N ranks.   bash tools/multi_check.sh N
set -u
N=${1:-2}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3 | tee $OUT/r1_multi_pytest_n$N.txt
bash tools/bench_n.sh $N
