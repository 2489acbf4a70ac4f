set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_ingest.py -m gpu -x -q 2>&1 | tail -25 | tee $OUT/r1_ingest_pytest.txt
