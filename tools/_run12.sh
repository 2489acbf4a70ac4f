set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee $OUT/r1_c3k_pytest.txt
KISS_B200_MGPU_TIMING=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 3 --warmup 3 --no-e2e > $OUT/r1_c3k_n2_bench.json 2> $OUT/r1_c3k_n2_bench.err
tail -c 600 $OUT/r1_c3k_n2_bench.json; grep -i "error\|Traceback" $OUT/r1_c3k_n2_bench.err | head -5
