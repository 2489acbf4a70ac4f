set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -4 | tee $OUT/r1_c3l_pytest_multi.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 4 --steps 3 --warmup 3 > $OUT/r1_c3l_n4_bench.json 2> $OUT/r1_c3l_n4_bench.err
tail -c 300 $OUT/r1_c3l_n4_bench.json; grep -i "error\|Traceback" $OUT/r1_c3l_n4_bench.err | head -5
KISS_B200_MGPU_TIMING=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus 4 --steps 2 --warmup 2 --no-e2e > $OUT/r1_c3l_n4_bench_marks.json 2> $OUT/r1_c3l_n4_bench_marks.err
