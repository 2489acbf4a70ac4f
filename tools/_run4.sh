set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 3 --warmup 3 > $OUT/r1_c3_n2_bench.json 2> $OUT/r1_c3_n2_bench.err
tail -c 1500 $OUT/r1_c3_n2_bench.json; tail -5 $OUT/r1_c3_n2_bench.err
