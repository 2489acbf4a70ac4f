set -u
OUT=gpurun_out; mkdir -p $OUT
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $OUT/r1_final_pytest.txt
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/r1_final_bench_reference.json 2>$OUT/r1_final_bench_reference.err; tail -c 300 $OUT/r1_final_bench_reference.json
python bench.py > $OUT/r1_final_bench.json 2> $OUT/r1_final_bench.err; tail -c 400 $OUT/r1_final_bench.json
