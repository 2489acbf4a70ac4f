set -u
N=${1:-2}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 tests/mgpu_check.py > $OUT/r1_peer_check.txt 2>&1
echo "mgpu_check rc=$?"; grep -v "^$" $OUT/r1_peer_check.txt | grep -v Warning | tail -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus $N --steps 3 --warmup 3 --no-e2e > $OUT/r1_c3m_n${N}_bench.json 2> $OUT/r1_c3m_n${N}_bench.err
tail -c 300 $OUT/r1_c3m_n${N}_bench.json; grep -i "error\|Traceback" $OUT/r1_c3m_n${N}_bench.err | head -5
KISS_B200_MGPU_TIMING=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29573 bench.py --gpus $N --steps 2 --warmup 2 --no-e2e > $OUT/r1_c3m_n${N}_bench_marks.json 2> $OUT/r1_c3m_n${N}_bench_marks.err
