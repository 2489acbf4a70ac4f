set -u
N=${1:-4}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 3 --warmup 3 > $OUT/r1_c3_n${N}_bench.json 2> $OUT/r1_c3_n${N}_bench.err
tail -c 600 $OUT/r1_c3_n${N}_bench.json; grep -i "error\|Traceback" $OUT/r1_c3_n${N}_bench.err | head -5
